#!/usr/bin/env python3
"""Summarise gpurun_out/*.ncu-rep + launches.csv into profiles/<tag>/ (raw csv per kernel + summary.md + traffic.json)."""
import collections, csv, io, json, os, subprocess, sys
tag = sys.argv[1]
out = os.path.join("profiles", tag)
os.makedirs(out, exist_ok=True)
WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__grid_size", "launch__block_size",
        "launch__registers_per_thread", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "gpu__compute_memory_throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "l1tex__t_sector_hit_rate.pct",
        "lts__t_sector_hit_rate.pct", "launch__waves_per_multiprocessor",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio"]
UNIT = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}
lines = ["# ncu summary (%s)" % tag, ""]
traffic = {}
# launch list
if os.path.exists("gpurun_out/launches.csv"):
    rows = [l for l in open("gpurun_out/launches.csv") if not l.startswith("==")]
    agg = collections.OrderedDict()
    for row in csv.DictReader(rows):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(row["Metric Value"].replace(",", ""))
        v = v / 1000 if row["Metric Unit"] == "ns" else (v * 1000 if row["Metric Unit"] == "ms" else v)
        agg.setdefault((row["Kernel Name"].split("(")[0][:60], row.get("Grid Size", "").replace(" ", "")), []).append(v)
    tot = sum(sum(v) for v in agg.values())
    lines += ["## launch list (`ncu --metrics gpu__time_duration.sum --clock-control none`, cold-cache, serialised)", "",
              "Grouped by grid: grid z (or x for the one-CTA-per-image kernels) = images in the batch; the rows with the largest batch are",
              "the bench's timed batches, the 2-image rows its single-frame latency measurement.", "",
              "| kernel | grid | launches | mean us | share |", "|---|---|---|---|---|"]
    for (k, g), v in agg.items():
        lines.append("| %s | %s | %d | %.2f | %.1f %% |" % (k, g, len(v), sum(v) / len(v), 100 * sum(v) / tot))
    open(os.path.join(out, "launches.csv"), "w").writelines(rows)
    lines.append("")
for f in sorted(os.listdir("gpurun_out")):
    if not f.endswith(".ncu-rep"):
        continue
    name = f[len("prof_"):-len(".ncu-rep")]
    raw = subprocess.run(["ncu", "-i", os.path.join("gpurun_out", f), "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    open(os.path.join(out, name + "_raw.csv"), "w").write(raw)
    r = list(csv.reader(io.StringIO(raw)))
    if len(r) < 3:
        continue
    hdr, units, vals = r[0], r[1], r[2]
    lines += ["## %s (`ncu --set full`, one launch)" % name, "", "| metric | value | unit |", "|---|---|---|"]
    d = {}
    for h, u, v in zip(hdr, units, vals):
        if h in WANT:
            lines.append("| %s | %s | %s |" % (h, v, u))
            d[h] = (v, u)
    try:
        rd = float(d["dram__bytes_read.sum"][0].replace(",", "")) * UNIT[d["dram__bytes_read.sum"][1]]
        wr = float(d["dram__bytes_write.sum"][0].replace(",", "")) * UNIT[d["dram__bytes_write.sum"][1]]
        traffic[name] = {"dram_bytes_per_launch": rd + wr, "grid": d["launch__grid_size"][0]}
    except (KeyError, ValueError):
        pass
    lines.append("")
open(os.path.join(out, "summary.md"), "w").write("\n".join(lines))
# images per launch = 2 x frames_per_batch of the profiled command (its plain run's JSON line is in ncu_plain.log)
frames = 32
try:
    for l in open("gpurun_out/ncu_plain.log"):
        if l.startswith("{"):
            frames = int(json.loads(l)["config"]["frames_per_batch"])
except (OSError, ValueError, KeyError):
    pass
json.dump({"_note": "dram__bytes_read.sum + dram__bytes_write.sum per launch from one `ncu --set full` capture per kernel of "
                    "`python bench.py --steps 6 --warmup 3 --no-cpu-baseline` (tools/ncu_round.sh; frames_per_batch %d = %d "
                    "images per launch); hamming_best_lb from the 10k x 1M sweep of the same run" % (frames, 2 * frames),
           "images_per_launch": 2 * frames, "kernels": traffic}, open(os.path.join(out, "traffic.json"), "w"), indent=1)
print("\n".join(lines))
