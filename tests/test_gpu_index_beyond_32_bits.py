"""GPU, maximum sizes: arrays with MORE THAN 2^32 ELEMENTS (uint8_t packs; the BASELINE configs top out at 2^31
elements, so only byte offsets pass 2^32 there).  An element index, tile count or group count held in 32 bits
anywhere on the path shows up here as a result that differs from the same work done in four separate calls, or
from the oracle on windows at the head, across the 2^32 boundary and at the tail.  Bit-exact (integer work)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from gpu_util import to_np  # noqa: E402

N = (1 << 32) + 4099
CUTS = [0, (1 << 30) + 16, (1 << 31) + 48, 3 * (1 << 30) + 80, N]   # pieces of about 2^30 elements
WINDOWS = (0, (1 << 32) - 2048, N - 4096)                            # head, across 2^32, tail
W = 4096


def _need(gib):
    free = torch.cuda.mem_get_info(0)[0]
    if free < gib << 30:
        pytest.skip(f"needs {gib} GiB of free device memory, {free >> 30} GiB free")


def _eq(x, y):
    return torch.equal(x.view(torch.int8), y.view(torch.int8))


def test_uint8_packs_beyond_2_32_elements(oracle):
    import pint_b200
    from pint_b200 import ops
    torch.cuda.set_device(0)
    _need(40)
    L = pint_b200.Layout(1, 3, 5)
    bits = (3, 5)
    a = torch.empty(N, dtype=torch.uint8, device="cuda")
    b = torch.empty_like(a)
    ops.fill_splitmix(a, 0xA11CE)
    ops.fill_splitmix(b, 0xB0B)
    pieces = torch.empty_like(a)
    for op in ("add_signed_saturate", "sub_unsigned_saturate"):
        out = ops.binary(op, L, a, b)
        for lo, hi in zip(CUTS, CUTS[1:]):
            ops.binary(op, L, a[lo:hi], b[lo:hi], out=pieces[lo:hi])
        assert _eq(out, pieces), op
        for lo in WINDOWS:
            want = oracle.binary(op, 1, bits, to_np(a[lo:lo + W]), to_np(b[lo:lo + W]))
            assert np.array_equal(to_np(out[lo:lo + W]), want), (op, lo)
    # two ops in one pass, the fused pipeline and the shift on the same arrays
    o1, o2 = ops.binary_multi(["add_signed_saturate", "sub_unsigned_saturate"], L, a, b)
    assert _eq(o2, out)
    assert _eq(ops.fused_unpack_arith_repack("sub_unsigned_saturate", L, a, b, out=pieces), out)
    sh = ops.shift_right_unsigned(L, a, 2, out=pieces)
    for lo in WINDOWS:
        assert np.array_equal(to_np(sh[lo:lo + W]), oracle.shift(False, 1, bits, 2, to_np(a[lo:lo + W]))), lo
    del o1, o2, sh
    # unpack to uint8 planes and back: 3 + 5 bits fill the byte, so the round trip is the identity
    planes = ops.unpack(L, out, signed=False, plane_bytes=1)
    for lo in WINDOWS:
        want = oracle.unpack(1, bits, to_np(out[lo:lo + W]), False, 1)
        for f in range(2):
            assert np.array_equal(to_np(planes[f][lo:lo + W]), want[f]), (f, lo)
    assert _eq(ops.repack(L, planes, out=pieces), out)
    del planes
    # operator== over the arrays: differences planted on both sides of 2^32
    c = a.clone()
    spots = torch.tensor([5, (1 << 31) + 7, (1 << 32) - 1, 1 << 32, (1 << 32) + 4098], device="cuda")
    c[spots] ^= 0x80
    flags, count = ops.equal(L, a, c)
    assert count == spots.numel()
    assert int((flags == 0).sum().item()) == spots.numel()
    assert bool((flags[spots] == 0).all().item())
    # checksum of the whole array == the four pieces' checksums combined
    s, x = ops.checksum(out)
    ps = [ops.checksum(out[lo:hi]) for lo, hi in zip(CUTS, CUTS[1:])]
    assert s == sum(p[0] for p in ps) & 0xFFFFFFFFFFFFFFFF
    xr = 0
    for p in ps:
        xr ^= p[1]
    assert x == xr


def test_pair_reduction_beyond_2_32_pairs(oracle):
    """the reference benchmark's loop (tests/pint_bench.cpp:90-94) over more than 2^32 interleaved pairs"""
    import pint_b200
    from pint_b200 import ops
    torch.cuda.set_device(0)
    _need(20)
    L = pint_b200.Layout(1, 3, 5)
    pairs = torch.empty(2 * N, dtype=torch.uint8, device="cuda")
    ops.fill_splitmix(pairs, 0x9A125)
    total = ops.pairs_reduce("add_unsigned_saturate", L, pairs)
    out = ops.binary_pairs("add_unsigned_saturate", L, pairs)
    assert total == ops.checksum(out)[0]
    parts = sum(ops.pairs_reduce("add_unsigned_saturate", L, pairs[2 * lo:2 * hi]) for lo, hi in zip(CUTS, CUTS[1:]))
    assert total == parts & 0xFFFFFFFFFFFFFFFF
    for lo in WINDOWS:
        w = to_np(pairs[2 * lo:2 * (lo + W)])
        want = oracle.binary("add_unsigned_saturate", 1, (3, 5), w[0::2].copy(), w[1::2].copy())
        assert np.array_equal(to_np(out[lo:lo + W]), want), lo


def test_dense_stream_beyond_2_32_elements():
    """5-bit elements <3,2> on a dense stream: 2^32 + 4099 elements in 2.7 GB; the whole stream in one call equals
    the stream processed in pieces that start on 128-element (16-byte) boundaries, and the padded kernels"""
    import pint_b200
    from pint_b200 import dense, ops
    torch.cuda.set_device(0)
    _need(40)
    L = pint_b200.Layout(1, 3, 2)
    x = torch.empty(N, dtype=torch.uint8, device="cuda")
    ops.fill_splitmix(x, 21)
    da = dense.pack(L, x)
    ops.fill_splitmix(x, 22)
    db = dense.pack(L, x)
    assert da.numel() == ((N * 5 + 127) // 128) * 16
    for op in ("add_signed_saturate", "min_unsigned"):
        got = dense.binary(op, L, da, db, N)
        pieces = torch.zeros_like(got)
        for lo, hi in zip(CUTS, CUTS[1:]):          # CUTS are multiples of 16 elements; round to 128
            lo, hi = lo // 128 * 128, (hi // 128 * 128 if hi != N else N)
            blo, bhi = lo * 5 // 8, (hi * 5 + 127) // 128 * 16 if hi == N else hi * 5 // 8
            dense.binary(op, L, da[blo:bhi], db[blo:bhi], hi - lo, out=pieces[blo:bhi])
        assert torch.equal(got, pieces), op
        a_p, b_p = dense.unpack(L, da, N), dense.unpack(L, db, N)
        assert torch.equal(got, dense.pack(L, ops.binary(op, L, a_p, b_p))), op
        del got, pieces, a_p, b_p
