"""GPU: pint_cuda_pairs_reduce_static -- the pair reduction for callers that may not allocate (its scratch is
a file-scope static device array of the library; include/pint_cuda.h).  No counterpart in the reference.
Called through the C ABI and compared, bit for bit (integer sums: the
tolerance of every test here is 0), with the existing pint_cuda_pairs_reduce on the same device arrays, on
calls in a row that SHARE the scratch array.

Collected last on purpose: it passes alone in a fresh process and after the
other tests, and a failure here hides nothing behind it.
"""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from gpu_util import to_dev  # noqa: E402

MAX_PAIRS = 8192 * 16384  # PINT_STATIC_MAX_PAIRS


def _both(lib, ops, L, op, pairs_t, n, res):
    """(existing entry point, static-scratch entry point) on the same device array; res = 2 device uint64"""
    vp = C.c_void_p
    code = ops.OP_CODE[op]
    res.fill_(-1)
    rc = lib.pint_cuda_pairs_reduce(code, L.ref, vp(pairs_t.data_ptr()), n, vp(res.data_ptr()), None)
    assert rc == 0, rc
    rc = lib.pint_cuda_pairs_reduce_static(code, L.ref, vp(pairs_t.data_ptr()), n, vp(res.data_ptr() + 8), None)
    assert rc == 0, rc
    a, b = res.tolist()
    return a & 0xFFFFFFFFFFFFFFFF, b & 0xFFFFFFFFFFFFFFFF


def test_static_scratch_reduce_matches_existing_entry_point_on_calls_in_a_row(oracle):
    import pint_b200
    from pint_b200 import _lib, ops
    lib = _lib.lib()
    torch.cuda.set_device(0)
    res = torch.zeros(2, dtype=torch.int64, device="cuda")
    # big first, small second: the second call finds the first one's partials still in the scratch array
    plan = [(4, (8, 8, 8, 8), "add_signed_saturate", (1 << 22) + 12345, 0),
            (4, (8, 8, 8, 8), "add_unsigned_saturate", 1000, 0),
            (4, (1, 2, 3, 4, 5, 6, 11), "sub_wrap", 16384 * 3, 0),          # exactly three CTA chunks
            (4, (1, 2, 3, 4, 5, 6, 11), "add_signed_saturate", 16385, 1),   # pairs one element off 16-byte alignment
            (2, (5, 6, 5), "add_wrap", (1 << 20) + 3, 0),
            (1, (3, 5), "max_signed", 70001, 0),
            (8, (3, 7, 11, 13, 17, 13), "min_unsigned", (1 << 19) + 5, 0),
            (8, (16, 16, 16, 16), "sub_signed_saturate", 7, 1),
            (4, (8, 8, 8, 8), "add_signed_saturate", 1, 0)]
    for eb, bits, op, n, off in plan:
        L = pint_b200.Layout(eb, *bits)
        flat = oracle.fill(eb, 2 * n + 2, 0x5A71C + n)
        dev = to_dev(flat)
        view = dev[2 * off:2 * off + 2 * n] if off == 0 else dev[1:1 + 2 * n]
        host = flat[:2 * n] if off == 0 else flat[1:1 + 2 * n]
        want_existing, got_static = _both(lib, ops, L, op, view, n, res)
        assert got_static == want_existing, (eb, bits, op, n, off)
        assert got_static == oracle.pairs_reduce(op, eb, bits, host), (eb, bits, op, n, off)   # and the checker agrees with both
    # n = 0 writes a zero sum
    res.fill_(-1)
    L = pint_b200.Layout(4, 8, 8, 8, 8)
    assert lib.pint_cuda_pairs_reduce_static(0, L.ref, None, 0, C.c_void_p(res.data_ptr()), None) == 0
    assert res.tolist()[0] == 0


def test_static_scratch_reduce_allocates_nothing_and_rejects_oversize(oracle):
    import pint_b200
    from pint_b200 import _lib, ops
    lib = _lib.lib()
    torch.cuda.set_device(0)
    L = pint_b200.Layout(4, 8, 8, 8, 8)
    n = (1 << 21) + 77
    flat = oracle.fill(4, 2 * n, 99)
    dev = to_dev(flat)
    res = torch.zeros(2, dtype=torch.int64, device="cuda")
    _both(lib, ops, L, "add_signed_saturate", dev, n, res)   # warm-up: the driver loads the kernels on first launch
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info(0)[0]
    torch_alloc0, torch_reserved0 = torch.cuda.memory_allocated(0), torch.cuda.memory_reserved(0)
    vp = C.c_void_p
    sums = []
    for _ in range(2):   # two calls in a row, nothing but the entry point between the measurements
        assert lib.pint_cuda_pairs_reduce_static(ops.OP_CODE["add_signed_saturate"], L.ref, vp(dev.data_ptr()), n, vp(res.data_ptr() + 8), None) == 0
        torch.cuda.synchronize()
        sums.append(res[1].item() & 0xFFFFFFFFFFFFFFFF)
    assert torch.cuda.mem_get_info(0)[0] == free0, "the call must not take device memory"
    assert (torch.cuda.memory_allocated(0), torch.cuda.memory_reserved(0)) == (torch_alloc0, torch_reserved0)
    assert sums[0] == sums[1] == oracle.pairs_reduce("add_signed_saturate", 4, (8, 8, 8, 8), flat)
    # beyond the fixed scratch: a clear error, nothing launched, the result untouched
    res.fill_(-7)
    before = ops.launch_count()
    rc = lib.pint_cuda_pairs_reduce_static(0, L.ref, vp(dev.data_ptr()), MAX_PAIRS + 1, vp(res.data_ptr()), None)
    assert rc == -21
    assert b"PINT_STATIC_MAX_PAIRS" in lib.pint_cuda_error_string(rc)
    assert ops.launch_count() == before
    torch.cuda.synchronize()
    assert res.tolist() == [-7, -7]
    # argument errors as the existing entry point reports them
    assert lib.pint_cuda_pairs_reduce_static(99, L.ref, vp(dev.data_ptr()), 4, vp(res.data_ptr()), None) == -10
    assert lib.pint_cuda_pairs_reduce_static(0, L.ref, vp(dev.data_ptr()), 4, None, None) == -1
    assert lib.pint_cuda_pairs_reduce_static(0, L.ref, None, 4, vp(res.data_ptr()), None) == -1


def test_static_scratch_reduce_at_the_maximum_size(oracle):
    """n == PINT_STATIC_MAX_PAIRS (every scratch slot in use; 1 GiB of uint32_t pairs) still equals the existing entry point."""
    import pint_b200
    from pint_b200 import _lib, ops
    lib = _lib.lib()
    torch.cuda.set_device(0)
    if torch.cuda.mem_get_info(0)[0] < (3 << 30):
        pytest.skip("needs 1 GiB of device memory for the pairs")
    L = pint_b200.Layout(4, 1, 3, 5, 11)
    pairs = torch.empty(2 * MAX_PAIRS, dtype=torch.uint32, device="cuda")
    ops.fill_splitmix(pairs, 0x57A71C)
    res = torch.zeros(2, dtype=torch.int64, device="cuda")
    for op in ("add_unsigned_saturate", "sub_signed_saturate"):
        a, b = _both(lib, ops, L, op, pairs, MAX_PAIRS, res)
        assert a == b, op
