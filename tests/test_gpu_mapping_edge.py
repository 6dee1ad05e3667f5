"""Stray accesses fault instead of hiding in allocator slack (tests/vmm_probe.py): result buffers
and array ends are placed on the last bytes of a VMM mapping followed by an unmapped page.
Runs in a subprocess because an illegal address poisons the CUDA context of the process."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PROBE = os.path.join(ROOT, "tests", "vmm_probe.py")
pytestmark = pytest.mark.gpu


def run_probe(*extra):
    return subprocess.run([sys.executable, PROBE, *extra], cwd=ROOT, capture_output=True, text=True, timeout=900)


def test_no_access_past_the_promised_bytes():
    r = run_probe()
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "vmm_probe ok" in r.stdout


def test_probe_traps_a_stray_access():
    """The harness itself: a result pointer 8 bytes late must NOT pass."""
    r = run_probe("--inject")
    assert r.returncode != 0, r.stdout[-2000:]


def test_round1_library_is_caught():
    """build/r1_regress/libpint_cuda_r1.so = this tree with round 1's pairs_reduce
    (atomicXor(sum_dev + 1, 0)), built by `make tools`: the probe must fail on it."""
    old = os.path.join(ROOT, "build", "r1_regress", "libpint_cuda_r1.so")
    if not os.path.exists(old):
        pytest.skip("regression library not built (make tools)")
    r = run_probe("--lib", old)
    assert r.returncode != 0, "round 1's out-of-bounds atomic went unnoticed:\n" + r.stdout[-2000:]
    assert "pairs_reduce" in r.stdout + r.stderr
