"""GPU: the host-buffer (end-to-end) entry points, the single-process multi-device
partitioner, and the C++ template layer (include/pint/cuda.hpp) binary."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from gpu_util import to_dev, to_np  # noqa: E402

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_binary_host_matches_oracle(oracle):
    import pint_b200
    from pint_b200 import ops
    ops.set_option("host_chunk_bytes", 1 << 20)   # force many chunks and slot reuse
    try:
        for eb, bits in [(4, (8, 8, 8, 8)), (8, (3, 7, 11, 13, 17, 13)), (2, (5, 6, 5)), (1, (2, 3, 3))]:
            L = pint_b200.Layout(eb, *bits)
            n = 3 * (1 << 20) // eb + 12345
            a, b = oracle.fill(eb, n, 41), oracle.fill(eb, n, 42)
            out = np.zeros_like(a)
            ops.binary_host("sub_signed_saturate", L, a, b, out)          # pageable numpy memory
            np.testing.assert_array_equal(out, oracle.binary("sub_signed_saturate", eb, bits, a, b))
            outs = [np.zeros_like(a), np.zeros_like(a)]
            ops.binary_multi_host(["add_unsigned_saturate", "add_signed_saturate"], L, a, b, outs)
            np.testing.assert_array_equal(outs[0], oracle.binary("add_unsigned_saturate", eb, bits, a, b))
            np.testing.assert_array_equal(outs[1], oracle.binary("add_signed_saturate", eb, bits, a, b))
    finally:
        ops.set_option("host_chunk_bytes", 32 << 20)


def test_binary_host_pinned(oracle):
    import pint_b200
    from pint_b200 import _lib, ops
    L = pint_b200.Layout(4, 8, 8, 8, 8)
    n = (1 << 22) + 7
    lib = _lib.lib()
    ptrs = [C.c_void_p() for _ in range(3)]
    for p in ptrs:
        assert lib.pint_cuda_host_alloc(C.byref(p), 4 * n) == 0
    a = np.ctypeslib.as_array((C.c_uint32 * n).from_address(ptrs[0].value))
    b = np.ctypeslib.as_array((C.c_uint32 * n).from_address(ptrs[1].value))
    out = np.ctypeslib.as_array((C.c_uint32 * n).from_address(ptrs[2].value))
    a[:], b[:] = oracle.fill(4, n, 1), oracle.fill(4, n, 2)
    ops.binary_host("add_signed_saturate", L, a, b, out)
    np.testing.assert_array_equal(out, oracle.binary("add_signed_saturate", 4, (8, 8, 8, 8), a, b))
    del a, b, out
    for p in ptrs:
        assert lib.pint_cuda_host_free(p) == 0


def test_sharded_binary_single_process(oracle):
    """pint_cuda_sharded_binary over the visible devices (1 device: both shards on it)."""
    import pint_b200
    from pint_b200 import sharded
    L = pint_b200.Layout(8, 16, 16, 16, 16)
    n = (1 << 20) + 99
    ndev = torch.cuda.device_count()
    k = max(2, ndev)
    plan = sharded.plan_shards(n, 8, k)
    a_h, b_h = oracle.fill(8, n, 5), oracle.fill(8, n, 6)
    a_s, b_s, o_s = [], [], []
    for i, (beg, cnt) in enumerate(plan):
        with torch.cuda.device(i % ndev):
            a_s.append(to_dev(a_h[beg:beg + cnt]).to(f"cuda:{i % ndev}"))
            b_s.append(to_dev(b_h[beg:beg + cnt]).to(f"cuda:{i % ndev}"))
            o_s.append(torch.empty_like(a_s[-1]))
    sharded.sharded_binary("sub_signed_saturate", L, a_s, b_s, o_s)
    for d in range(ndev):
        torch.cuda.synchronize(d)
    got = np.concatenate([to_np(o) for o in o_s])
    np.testing.assert_array_equal(got, oracle.binary("sub_signed_saturate", 8, (16, 16, 16, 16), a_h, b_h))


def test_sharded_fused_unpack_repack(oracle):
    """C4's shape: the fused kernel and unpack/repack over per-device shards (round-robin over the visible devices)."""
    import pint_b200
    from pint_b200 import sharded
    bits = (3, 7, 11, 13, 17, 13)
    L = pint_b200.Layout(8, *bits)
    n = (1 << 19) + 77
    ndev = torch.cuda.device_count()
    k = max(3, ndev)
    plan = sharded.plan_shards(n, 8, k)
    a_h, b_h = oracle.fill(8, n, 51), oracle.fill(8, n, 52)
    a_s, b_s, o_s, p_s = [], [], [], []
    for i, (beg, cnt) in enumerate(plan):
        dev = f"cuda:{i % ndev}"
        a_s.append(to_dev(a_h[beg:beg + cnt]).to(dev))
        b_s.append(to_dev(b_h[beg:beg + cnt]).to(dev))
        o_s.append(torch.empty_like(a_s[-1]))
        p_s.append([torch.empty(cnt, dtype=torch.int32, device=dev) for _ in bits])
    sharded.sharded_binary("add_signed_saturate", L, a_s, b_s, o_s, fused=True)
    sharded.sharded_unpack(L, o_s, p_s, signed=True)
    back = [torch.empty_like(o) for o in o_s]
    sharded.sharded_repack(L, p_s, back)
    for d in range(ndev):
        torch.cuda.synchronize(d)
    want = oracle.binary("add_signed_saturate", 8, bits, a_h, b_h)
    np.testing.assert_array_equal(np.concatenate([to_np(o) for o in o_s]), want)
    np.testing.assert_array_equal(np.concatenate([to_np(o) for o in back]), want)
    wp = oracle.unpack(8, bits, want, True, 4)
    for f in range(len(bits)):
        np.testing.assert_array_equal(np.concatenate([to_np(p[f], np.int32) for p in p_s]), wp[f])


def test_persistent_grid_option_gives_same_results(oracle):
    import pint_b200
    from pint_b200 import ops
    L = pint_b200.Layout(4, 1, 2, 3, 4, 5, 6, 11)
    n = (1 << 21) + 3
    a_h, b_h = oracle.fill(4, n, 8), oracle.fill(4, n, 9)
    a, b = to_dev(a_h), to_dev(b_h)
    want = oracle.binary("add_signed_saturate", 4, L.bits, a_h, b_h)
    for cps in (1, 8, 0):
        ops.set_option("ctas_per_sm", cps)
        np.testing.assert_array_equal(to_np(ops.add_signed_saturate(L, a, b)), want)
        planes = ops.unpack(L, a, signed=True)
        np.testing.assert_array_equal(to_np(ops.repack(L, planes)), a_h)
    ops.set_option("ctas_per_sm", 0)


def test_ordered_persistent_mode_matches_one_cta_per_tile(oracle):
    """Large light ops run the ordered persistent kernel (atomic tile counter); same bits either way,
    also on repeated launches (the counter re-arms itself) and on concurrent streams (counter ring)."""
    import pint_b200
    from pint_b200 import ops
    for eb, bits in [(4, (8, 8, 8, 8)), (2, (5, 6, 5)), (8, (16, 16, 16, 16))]:
        L = pint_b200.Layout(eb, *bits)
        n = (24 << 20) // eb + 1029     # > 2 * 148 * 4 tiles of 8 KiB, ragged tail
        a_h, b_h = oracle.fill(eb, n, 31), oracle.fill(eb, n, 32)
        a, b = to_dev(a_h), to_dev(b_h)
        want = {op: oracle.binary(op, eb, bits, a_h, b_h) for op in ("add_unsigned_saturate", "sub_wrap", "bit_xor")}
        want_shift = oracle.shift(0, eb, bits, 1, a_h)
        for mode in (4, 0, 4):
            ops.set_option("ordered_ctas_per_sm", mode)
            for rep in range(3):
                for op, w in want.items():
                    np.testing.assert_array_equal(to_np(ops.binary(op, L, a, b)), w, err_msg=f"{L} {op} ordered={mode}")
                np.testing.assert_array_equal(to_np(ops.shift_right_unsigned(L, a, 1)), want_shift)
        ops.set_option("ordered_ctas_per_sm", 4)
        streams = [torch.cuda.Stream() for _ in range(3)]
        outs = [torch.empty_like(a) for _ in streams]
        torch.cuda.synchronize()
        for _ in range(4):
            for s, o in zip(streams, outs):
                ops.binary("add_unsigned_saturate", L, a, b, out=o, stream=s)
        torch.cuda.synchronize()
        for o in outs:
            np.testing.assert_array_equal(to_np(o), want["add_unsigned_saturate"])


def test_cpp_template_layer():
    exe = os.path.join(ROOT, "build", "cpp_api_test")
    if not os.path.exists(exe):
        pytest.skip("build/cpp_api_test not built (make tools)")
    out = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "0 failures" in out.stdout


def test_cpp_template_layer_compiled_by_nvcc():
    """include/pint/cuda.hpp under nvcc instantiates the kernels for static_layout<Integer, Bits...> in the
    user's translation unit (compiled-in constants for ANY layout): every op on four unregistered layouts
    equals the library's parity-tested run-time-mask path bit for bit (tests/cpp/cuda_api_nvcc_test.cu)."""
    exe = os.path.join(ROOT, "build", "cuda_api_nvcc_test")
    if not os.path.exists(exe):
        pytest.skip("build/cuda_api_nvcc_test not built (make tools)")
    out = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "total: 0 failures" in out.stdout


def test_sharded_host_arrays_all_devices(oracle):
    """pint_cuda_sharded_binary_multi_host: HOST arrays split over every visible device, each shard
    through its device's ring; ragged n, several layouts, small chunks (many slot reuses)."""
    import pint_b200
    from pint_b200 import ops, sharded
    ndev = torch.cuda.device_count()
    ops.set_option("host_chunk_bytes", 1 << 19)
    try:
        for eb, bits in [(4, (8, 8, 8, 8)), (8, (16, 16, 16, 16)), (2, (5, 6, 5)), (4, (7, 9, 3))]:
            L = pint_b200.Layout(eb, *bits)
            for devices in ([0], list(range(ndev)), list(range(ndev))[::-1]):
                n = 5 * (1 << 19) // eb * len(devices) + 4321
                a, b = oracle.fill(eb, n, 61), oracle.fill(eb, n, 62)
                pair = ["add_unsigned_saturate", "add_signed_saturate"]
                outs = [np.zeros_like(a), np.zeros_like(a)]
                sharded.sharded_binary_multi_host(pair, L, devices, a, b, outs)
                for op, o in zip(pair, outs):
                    np.testing.assert_array_equal(o, oracle.binary(op, eb, bits, a, b))
                one = [np.zeros_like(a)]
                sharded.sharded_binary_multi_host(["sub_signed_saturate"], L, devices, a, b, one)
                np.testing.assert_array_equal(one[0], oracle.binary("sub_signed_saturate", eb, bits, a, b))
        # tiny and empty arrays: fewer elements than devices * 128 bytes
        L = pint_b200.Layout(4, 8, 8, 8, 8)
        for n in (0, 1, 31, 33):
            a, b = oracle.fill(4, n, 1), oracle.fill(4, n, 2)
            out = [np.zeros_like(a)]
            sharded.sharded_binary_multi_host(["add_wrap"], L, list(range(ndev)), a, b, out)
            np.testing.assert_array_equal(out[0], oracle.binary("add_wrap", 4, (8, 8, 8, 8), a, b))
    finally:
        ops.set_option("host_chunk_bytes", 32 << 20)


def test_sharded_host_argument_errors():
    import pint_b200
    from pint_b200 import _lib
    lib = _lib.lib()
    L = pint_b200.Layout(4, 8, 8, 8, 8)
    a = np.zeros(16, dtype=np.uint32)
    outs = (C.c_void_p * 1)(a.ctypes.data)
    codes = (C.c_int * 1)(0)
    dup = (C.c_int * 2)(0, 0)
    f = lib.pint_cuda_sharded_binary_multi_host
    assert f(1, codes, L.ref, 2, dup, a.ctypes.data, a.ctypes.data, outs, 16) == -15      # a device twice
    assert f(1, codes, L.ref, 0, dup, a.ctypes.data, a.ctypes.data, outs, 16) == -15
    assert f(1, codes, L.ref, 1, dup, None, a.ctypes.data, outs, 16) == -1
    bad = (C.c_int * 1)(99)
    assert f(1, bad, L.ref, 1, dup, a.ctypes.data, a.ctypes.data, outs, 16) == -10


def test_host_calls_on_two_threads_and_release(oracle):
    """Per-device ring lock: two threads on the same device serialise but stay correct, threads on
    different devices run concurrently; pint_cuda_host_release frees the rings and the next call
    allocates them again."""
    import threading
    import pint_b200
    from pint_b200 import _lib, ops
    L = pint_b200.Layout(4, 8, 8, 8, 8)
    ndev = torch.cuda.device_count()
    n = (1 << 21) + 17
    a, b = oracle.fill(4, n, 71), oracle.fill(4, n, 72)
    want = oracle.binary("add_signed_saturate", 4, (8, 8, 8, 8), a, b)
    outs = [np.zeros_like(a) for _ in range(4)]
    errs = []

    def work(i):
        try:
            torch.cuda.set_device(i % ndev)
            for _ in range(3):
                ops.binary_host("add_signed_saturate", L, a, b, outs[i])
        except Exception as e:  # noqa: BLE001
            errs.append(e)
    th = [threading.Thread(target=work, args=(i,)) for i in range(4)]
    [t.start() for t in th]
    [t.join() for t in th]
    assert not errs, errs
    for o in outs:
        np.testing.assert_array_equal(o, want)
    free0 = torch.cuda.mem_get_info(0)[0]
    assert _lib.lib().pint_cuda_host_release() == 0
    assert torch.cuda.mem_get_info(0)[0] > free0, "release must return the staging memory"
    out = np.zeros_like(a)
    ops.binary_host("add_signed_saturate", L, a, b, out)
    np.testing.assert_array_equal(out, want)


def test_scatter_compute_gather_from_one_device(oracle):
    """An array resident on device 0 -> shards on all devices (one peer copy each) -> op per shard ->
    gathered back onto device 0; equals the op run on the whole array."""
    import pint_b200
    from pint_b200 import ops, sharded
    L = pint_b200.Layout(8, 3, 7, 11, 13, 17, 13)
    bits = (3, 7, 11, 13, 17, 13)
    ndev = torch.cuda.device_count()
    devices = list(range(ndev)) if ndev > 1 else [0, 0, 0]
    sharded.enable_peer_access(list(range(ndev)))
    n = (1 << 20) + 777
    a_h, b_h = oracle.fill(8, n, 81), oracle.fill(8, n, 82)
    a, b = to_dev(a_h), to_dev(b_h)
    torch.cuda.synchronize()
    a_s, b_s = sharded.scatter_shards(a, devices), sharded.scatter_shards(b, devices)
    for d in range(ndev):
        torch.cuda.synchronize(d)
    o_s = [torch.empty_like(t) for t in a_s]
    sharded.sharded_binary("add_signed_saturate", L, a_s, b_s, o_s, fused=True)
    for d in range(ndev):
        torch.cuda.synchronize(d)
    out = torch.zeros_like(a)
    sharded.gather_shards(o_s, out)
    for d in range(ndev):
        torch.cuda.synchronize(d)
    np.testing.assert_array_equal(to_np(out), oracle.binary("add_signed_saturate", 8, bits, a_h, b_h))
