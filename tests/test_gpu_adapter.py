"""The compiled reference-side binding: adapter/okvis_cuda_adapter.hpp (cv::FeatureDetector / brisk::BriskDescriptorExtractor
subclasses on the C ABI) run next to the reference's own classes by tests/cpp/adapter_test.cc, through the base-class
pointers okvis::Frame uses. The program is built by __graft_entry__.build() against the reference's brisk headers
(this container) and travels prebuilt to the GPU box."""
import os
import subprocess

import pytest

from conftest import ROOT

BIN = os.path.join(ROOT, "tests", "cpp", "_build", "adapter_test")


def test_adapter_test_program_is_built(mod):
    """CPU: build() produced the program wherever the reference headers exist (the adapter header compiles)."""
    if not os.path.isdir("/root/reference/external/brisk") and not os.path.exists(BIN):
        pytest.skip("no reference tree and no prebuilt program")
    assert os.path.exists(BIN) and os.access(BIN, os.X_OK)


@pytest.mark.gpu
def test_adapter_matches_reference_classes(mod):
    assert os.path.exists(BIN), "tests/cpp/_build/adapter_test missing: run __graft_entry__.build() where /root/reference exists"
    r = subprocess.run([BIN], capture_output=True, text=True, timeout=600)
    print(r.stdout)
    assert r.returncode == 0 and "ADAPTER_TEST PASS" in r.stdout, r.stdout + r.stderr
    assert r.stdout.count("ok  ") >= 10      # 3 configurations x 3 frames + the error cases
