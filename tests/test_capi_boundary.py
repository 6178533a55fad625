"""CPU-only: the C-ABI library loads, exports every symbol include/okvis_cuda.h declares, and refuses to compute
without a CUDA device (no CPU fallback)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import ROOT


def declared_symbols(headers=("okvis_cuda.h", "okvis_cuda_debug.h")):
    names = set()
    for h in headers:
        src = open(os.path.join(ROOT, "include", h)).read()
        src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
        names |= set(re.findall(r"\b(okc_[a-z0-9_]+)\s*\(", src))
    return sorted(names)


def test_header_symbols_are_exported(mod):
    L = mod.lib()
    names = declared_symbols()
    assert len(names) >= 20
    for n in names:
        assert hasattr(L, n), "libokvis_cuda.so does not export %s" % n
    for n in mod.EXPORTS:
        assert n in names, "%s is bound by the host mirror but not declared in the header" % n


def test_exports_are_exactly_the_declared_symbols(mod):
    """The library is built with -fvisibility=hidden: nothing but the entry points of include/*.h is exported (no
    undeclared test switches, no internal C++ symbols)."""
    import subprocess
    out = subprocess.run(["nm", "-D", "--defined-only", mod.LIB_PATH], capture_output=True, text=True).stdout
    exported = sorted(l.split()[2] for l in out.splitlines() if len(l.split()) == 3 and l.split()[1] == "T")
    assert exported == declared_symbols()


def test_struct_layouts_match_header(mod):
    assert mod.KP_DTYPE.itemsize == 28  # cv::KeyPoint
    assert C.sizeof(mod.DetectParams) == 32
    assert C.sizeof(mod.DescribeParams) == 20
    assert C.sizeof(mod.StereoView) == 4 * 8 + 8 + 9 * 8 + 3 * 8 + 8


def test_library_has_sm100a_code_and_no_oracle_dependency(mod):
    import subprocess
    out = subprocess.run(["cuobjdump", "--list-elf", mod.LIB_PATH], capture_output=True, text=True).stdout
    assert "sm_100a" in out
    ldd = subprocess.run(["ldd", mod.LIB_PATH], capture_output=True, text=True).stdout
    assert "oracle" not in ldd and "brisk_ref" not in ldd


def test_no_cpu_fallback_without_device(mod):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    h = C.c_void_p()
    rc = mod.lib().okc_context_create(0, C.byref(h))
    assert rc == 2  # OKC_ERR_CUDA
    assert b"no CUDA device" in mod.lib().okc_last_error_string()
    with pytest.raises(mod.OkcError):
        mod.Context(0)


def test_host_mirror_rejects_what_the_reference_rejects(mod):
    """scale-space-feature-detector.h:95-97 throws for non-continuous / non-u8 / multi-channel images."""
    cam = mod.Camera.__new__(mod.Camera)
    cam.W, cam.H = 64, 48
    with pytest.raises(mod.OkcError):
        cam._img(np.zeros((48, 64), np.float32))
    with pytest.raises(mod.OkcError):
        cam._img(np.zeros((48, 64, 3), np.uint8))
    with pytest.raises(mod.OkcError):
        cam._img(np.zeros((48, 128), np.uint8)[:, ::2])
    with pytest.raises(mod.OkcError):
        cam._img(np.zeros((40, 64), np.uint8))
    assert cam._img(np.zeros((48, 64), np.uint8)).shape == (48, 64)
