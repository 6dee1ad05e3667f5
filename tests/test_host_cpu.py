"""CPU-side checks: the C-ABI library loads without a GPU and exports every
symbol include/pint_cuda.h declares; layout validation, the mask generator and
the shard planner (host logic) agree with the oracle.  No kernel is launched."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import pint_b200
from pint_b200 import _lib, sharded


def declared_symbols():
    src = open(_lib.HEADER_PATH).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(pint_cuda_\w+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    lib = _lib.lib()
    names = declared_symbols()
    assert len(names) >= 24
    for name in names:
        assert hasattr(lib, name), f"{name} declared in include/pint_cuda.h but not exported"
    assert sorted(lib._pint_symbols) == names, "pint_b200/_lib.py binds exactly the declared ABI"
    assert lib.pint_cuda_abi_version() == 1


def test_no_cpu_fallback():
    """Without a device a compute call returns a CUDA error; it never computes on the host."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("has a GPU")
    L = pint_b200.Layout.make(5, 6, 5)
    a = np.arange(64, dtype=np.uint16)
    out = np.zeros_like(a)
    rc = _lib.lib().pint_cuda_binary(0, L.ref, a.ctypes.data, a.ctypes.data, out.ctypes.data, a.size, None)
    assert rc > 0, "expected a cudaError_t"
    assert not out.any()
    from pint_b200 import ops
    with pytest.raises(TypeError):
        ops.add_wrap(L, torch.zeros(4, dtype=torch.int16), torch.zeros(4, dtype=torch.int16))


def test_callers_argument_checks_need_no_device():
    """pint_cuda_binary_scalar / _pairs reject bad arguments before any CUDA call."""
    lib = _lib.lib()
    L = pint_b200.Layout.make(5, 6, 5)
    a = np.zeros(64, dtype=np.uint16)
    p = a.ctypes.data
    assert lib.pint_cuda_binary_scalar(0, L.ref, p, 1 << 16, 0, p, 16, None) == -15
    assert lib.pint_cuda_binary_scalar(99, L.ref, p, 1, 0, p, 16, None) == -10
    assert lib.pint_cuda_binary_scalar(0, L.ref, p, 1, 0, p + 2, 16, None) == -14
    assert lib.pint_cuda_binary_scalar(0, L.ref, p, 1, 1, None, 16, None) == -1
    assert lib.pint_cuda_binary_pairs(0, L.ref, p, p + 16, 16, None) == -14
    assert lib.pint_cuda_binary_pairs(13, L.ref, p, p + 64, 16, None) == -10
    assert lib.pint_cuda_binary_pairs(0, L.ref, p, p + 64, 0, None) == 0
    assert lib.pint_cuda_equal(L.ref, p, p, 16, None, None, None) == -1


def test_layout_validation_matches_oracle(oracle):
    rng = np.random.default_rng(7)
    lib = _lib.lib()
    for _ in range(2000):
        eb = int(rng.choice([1, 2, 3, 4, 8, 16]))
        nf = int(rng.integers(0, 12))
        bits = [int(x) for x in rng.integers(0, 20, nf)]
        l = _lib.PintLayout()
        l.elem_bytes, l.nfields = eb, nf
        for i, b in enumerate(bits):
            l.bits[i] = b
        assert lib.pint_cuda_layout_validate(C.byref(l)) == oracle.validate(eb, bits), (eb, bits)


def test_mask_generator_matches_oracle(oracle, refvec):
    for _, eb, bits in refvec.layouts():
        got = pint_b200.Layout(eb, *bits).masks()
        _, m = oracle.masks(eb, bits)
        assert got == {"H": m.H, "L": m.L, "A": m.A, "sat_type": m.sat_type}, (eb, bits)
    rng = np.random.default_rng(11)
    seen = set()
    for _ in range(500):
        eb = int(rng.choice([1, 2, 4, 8]))
        bits, room = [], 8 * eb
        while room > 0 and len(bits) < 64 and rng.random() < 0.85:
            w = int(rng.integers(1, min(room, 20) + 1))
            bits.append(w)
            room -= w
        if not bits:
            continue
        got = pint_b200.Layout(eb, *bits).masks()
        _, m = oracle.masks(eb, bits)
        assert got == {"H": m.H, "L": m.L, "A": m.A, "sat_type": m.sat_type}, (eb, bits)
        seen.add(m.sat_type)
    assert seen == {0, 1, 2}


def test_storage_bytes_matches_reference_static_asserts(kats):
    for k in kats:
        if k["kind"] == "storage":
            assert pint_b200.storage_bytes(k["total_bits"]) == k["elem_bytes"]
    assert pint_b200.storage_bytes(0) == 0 and pint_b200.storage_bytes(65) == 0
    assert pint_b200.Layout.make(5, 6, 5).elem_bytes == 2
    assert pint_b200.Layout.make(3, 7, 11, 13, 17, 13).elem_bytes == 8


def test_registry_has_baseline_layouts():
    for eb, bits in [(2, (5, 6, 5)), (4, (8, 8, 8, 8)), (8, (16, 16, 16, 16)), (8, (3, 7, 11, 13, 17, 13))]:
        assert pint_b200.Layout(eb, *bits).is_specialized()
    assert not pint_b200.Layout(4, 7, 9, 3).is_specialized()


def test_slice_layout():
    L = pint_b200.Layout(2, 1, 2, 3, 4, 5)
    assert L.sliced(2, 4) == pint_b200.Layout(2, 3, 4)   # tests/pint_test.cpp:135-146
    with pytest.raises(ValueError):
        L.sliced(3, 3)
    res = _lib.PintLayout()
    assert _lib.lib().pint_cuda_slice_layout(L.ref, 2, 4, C.byref(res)) == 0
    assert (res.elem_bytes, res.nfields, res.bits[0], res.bits[1]) == (2, 2, 3, 4)
    assert _lib.lib().pint_cuda_slice_layout(L.ref, 4, 6, C.byref(res)) == -12


@pytest.mark.parametrize("n", [0, 1, 127, 128, 1000003, 2 ** 31, 2 ** 33 + 5])
@pytest.mark.parametrize("eb", [1, 2, 4, 8])
@pytest.mark.parametrize("k", [1, 2, 3, 4, 8])
def test_plan_shards(n, eb, k):
    plan = sharded.plan_shards(n, eb, k)
    assert len(plan) == k
    pos = 0
    for i, (b, c) in enumerate(plan):
        assert b == pos
        assert (b * eb) % 128 == 0, "shard starts on a 128-byte boundary"
        pos += c
    assert pos == n
    if n >= k * 128:
        counts = [c for _, c in plan]
        assert max(counts) - min(counts) < k * (128 // eb) + 128 // eb


def test_options_roundtrip():
    from pint_b200 import ops
    ops.set_option("force_generic", 1)
    assert ops.get_option("force_generic") == 1
    assert not pint_b200.Layout(4, 8, 8, 8, 8).is_specialized()
    ops.set_option("force_generic", 0)
    assert pint_b200.Layout(4, 8, 8, 8, 8).is_specialized()
    with pytest.raises(_lib.PintCudaError):
        ops.set_option("no_such_option", 1)


def test_bench_reference_arm_line():
    """`bench.py --impl reference` needs no GPU: one JSON line with the contract's keys; under torchrun
    (RANK != 0) it prints nothing and exits 0."""
    import json
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--steps", "2", "--warmup", "1", "--cpu-log2", "18"]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=120, check=True).stdout
    lines = [l for l in out.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "packed elements/sec" and d["unit"] == "elements/s"
    assert d["value"] > 0 and d["higher_is_better"] is True and d["steps"] == 2 and d["vs_baseline"] is None
    assert d["cpu_baseline"]["kind"] in ("reference", "port") and d["cpu_baseline"]["cores"] >= 1
    assert d["e2e"] == {"value": d["value"], "unit": "elements/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert "workload" in d["config"] and "model" not in d["config"]
    env = dict(os.environ, RANK="1", LOCAL_RANK="1", WORLD_SIZE="2")
    r = subprocess.run(cmd + ["--gpus", "2"], capture_output=True, text=True, timeout=120, env=env)
    assert r.returncode == 0 and not r.stdout.strip()


def test_jit_prebuild_without_a_device(tmp_path, monkeypatch):
    """pint_cuda_jit_prebuild: the NVRTC specialisation of the embedded kernel templates for an
    unregistered layout compiles for sm_100a on a machine with no GPU and fills the disk cache."""
    import ctypes.util
    import pint_b200
    from pint_b200 import _lib
    if not (os.path.exists("/usr/local/cuda/lib64/libnvrtc.so.12") or ctypes.util.find_library("nvrtc")):
        pytest.skip("libnvrtc not installed")
    monkeypatch.setenv("PINT_CUDA_JIT_CACHE", str(tmp_path / "jit"))
    lib = _lib.lib()
    L = pint_b200.Layout(4, 7, 9, 3)
    rc = lib.pint_cuda_jit_prebuild(L.ref, 10, 0)
    assert rc == 0, lib.pint_cuda_jit_last_error().decode()[:2000]
    files = sorted(os.listdir(tmp_path / "jit"))
    # padded families (binary, fused, pair), the caller shapes (scalar, pairsmap) and, since sum(Bits) = 19 is a
    # whole-chunk dense layout, the dense ones
    fams = ["binary", "dense", "densepair", "fused", "pair", "pairsmap", "scalar"]
    assert sorted(f.split("_")[1] for f in files) == sorted(f"{x}.{e}" for x in fams for e in ("cubin", "names"))
    def kernel_names(suffix):
        """the lowered kernel names of a cache entry; its first line is "#cubin <bytes> <fnv1a>" (checked against the image on load)"""
        path = tmp_path / "jit" / [f for f in files if f.endswith(suffix)][0]
        lines = open(path).read().splitlines()
        head = lines[0].split()
        assert head[0] == "#cubin" and int(head[1]) == os.path.getsize(str(path)[:-len(".names")] + ".cubin"), head
        return lines[1:]
    sn = kernel_names("_scalar.names")
    assert len(sn) == 26 and all("map1_stream_kernel" in x and "scalar_fn" in x for x in sn)
    dn = kernel_names("_dense.names")
    assert len(dn) == 13 and all("dense_binary_kernel" in x and "chunk_layout" in x for x in dn)
    names = kernel_names("binary.names")
    assert len(names) == 26 and all("binary_stream" in x and "static_layout" in x for x in names)
    assert lib.pint_cuda_jit_prebuild(L.ref, 7, 0) == -15                      # not a target this library has
    bad = pint_b200._lib.PintLayout(4, 2, (pint_b200._lib.C.c_uint8 * 64)(20, 20))
    assert lib.pint_cuda_jit_prebuild(bad, 10, 0) == -5                        # the reference's static_assert (pint.hpp:763)


def test_mask_generator_equals_reference_metafunctions_at_compile_time():
    """tests/cpp/static_masks_check.cpp static_asserts H / L / A / saturation-mask type of
    static_layout<Integer, Bits...> against the reference's metafunctions (pint.hpp:324-348, 443-456)
    for every layout of oracle/ref_layouts.def: compiling it IS the test."""
    import shutil
    import subprocess
    ref = "/root/reference/include/pint/pint.hpp"
    if not os.path.exists(ref) or not shutil.which("g++"):
        pytest.skip("needs /root/reference and g++ (build container only)")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run(["g++", "-std=c++17", "-fsyntax-only", "-Iinclude", "-I/root/reference/include", "-Ioracle",
                        "tests/cpp/static_masks_check.cpp"], cwd=root, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-3000:]
