"""CPU-only: the built library contains no contracted multiply-add (tools/sass_audit.py; profiles/r02_sass_audit.md).

--fmad=false is what makes the float / double paths round like the x86 reference; this guards the flag and any
explicit fma() sneaking into a kernel."""
import os
import shutil
import subprocess
import sys

import pytest

from conftest import ROOT


@pytest.mark.skipif(shutil.which("nvdisasm") is None or shutil.which("cuobjdump") is None, reason="CUDA binary tools missing")
def test_no_contracted_or_explicit_fma_in_the_library(mod, tmp_path):
    md = tmp_path / "audit.md"
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "sass_audit.py"), "--md", str(md)],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    text = md.read_text()
    assert "| OTHER (inspect) |" not in text and "| explicit fma() |" not in text
    assert "DFMA" in text      # the audit saw the kernels (division / sqrt sequences are always there)
