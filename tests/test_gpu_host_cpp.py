"""Drop-in test of the C++ host adapter: the reference's own DevStep (CPU) and host/GpuStep
(B200), both driven through the reference's Iterator/IModelStep interface over
std::vector<ElementData>, must leave the same state (1e-10 relative).  The binary is built in
the container against the reference tree (host/Makefile) and travels to the GPU box."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "host", "_build", "test_gpustep")


@pytest.mark.gpu
def test_gpustep_is_a_drop_in_for_devstep():
    if not os.path.exists(BIN):
        pytest.skip("host/_build/test_gpustep not built (needs the reference tree at build time)")
    r = subprocess.run([BIN, "200000", "3", "600"], capture_output=True, text=True, timeout=600)
    print(r.stdout, r.stderr)
    assert r.returncode == 0 and "PASS" in r.stdout, r.stdout + r.stderr
    assert "albedo_kind=2 freezing_kind=1" in r.stdout  # module selection came from [Modules]


def test_no_gpu_means_loud_failure():
    """Without a device the adapter throws; it never computes on the CPU."""
    if not os.path.exists(BIN) or os.path.exists("/dev/nvidia0"):
        pytest.skip("needs the prebuilt binary and a box without a GPU")
    r = subprocess.run([BIN, "10", "1"], capture_output=True, text=True, timeout=120)
    assert r.returncode != 0 and "no CPU fallback" in (r.stdout + r.stderr)
