"""GPU parity for the two extra caller shapes of the 13 operations, through the C ABI:
pint_cuda_binary_pairs (interleaved std::pair operands, tests/pint_bench.cpp:22) and
pint_cuda_binary_scalar (one operand the same packed_int for every element).
Bit-exact against the compiled reference's vectors and the oracle."""
import ctypes as C

import numpy as np
import pytest
import torch

from oracle import OPS

pytestmark = pytest.mark.gpu

from gpu_util import to_dev, to_np  # noqa: E402

DT = {1: np.uint8, 2: np.uint16, 4: np.uint32, 8: np.uint64}
SPOT = [(4, (8, 8, 8, 8)), (2, (5, 6, 5)), (8, (3, 7, 11, 13, 17, 13)), (1, (2, 3, 3)), (4, (7, 9, 3)), (8, (64,)),
        (2, (1, 1, 1, 1, 1)), (4, (1, 2, 3, 4, 5, 6, 11)), (8, (16, 16, 16, 16)), (8, (8, 8, 8, 8, 8, 8, 8, 8))]


@pytest.fixture(scope="module")
def P():
    import pint_b200
    from pint_b200 import ops
    assert torch.cuda.is_available()
    return pint_b200, ops


@pytest.fixture(params=["static_or_generic", "force_generic"])
def mode(request, P):
    """Registered layouts run functors with compiled-in masks; force_generic sends them through the
    run-time-mask kernels every other layout uses."""
    _, ops = P
    ops.set_option("force_generic", 1 if request.param == "force_generic" else 0)
    yield request.param
    ops.set_option("force_generic", 0)


def interleave(a, b):
    return np.stack([a, b], axis=1).reshape(-1)


def test_golden_binary_pairs(P, refvec, mode):
    """13 ops x 33 layouts on interleaved operands == the unmodified reference header's results."""
    pint_b200, ops = P
    for key, eb, bits in refvec.layouts():
        L = pint_b200.Layout(eb, *bits)
        pairs = to_dev(interleave(refvec[key + "_a"], refvec[key + "_b"]))
        for op in OPS:
            np.testing.assert_array_equal(to_np(ops.binary_pairs(op, L, pairs)), refvec[f"{key}_{op}"],
                                          err_msg=f"pairs {L} {op}")


def test_golden_binary_scalar(P, refvec, oracle, ref, mode):
    """OP(a[i], k) and OP(k, a[i]) for k taken from the reference vectors (structured edge values
    and random words, unused high bits set), 13 ops x 33 layouts."""
    pint_b200, ops = P
    for key, eb, bits in refvec.layouts():
        L = pint_b200.Layout(eb, *bits)
        a_h, b_h = refvec[key + "_a"], refvec[key + "_b"]
        a = to_dev(a_h)
        for j in (0, 3, len(b_h) // 2, len(b_h) - 1):
            k = int(b_h[j])
            kk = np.full_like(a_h, k)
            for op in OPS:
                want = oracle.binary(op, eb, bits, a_h, kk)
                np.testing.assert_array_equal(to_np(ops.binary_scalar(op, L, a, k)), want, err_msg=f"{L} {op} k={k:#x}")
                want_first = oracle.binary(op, eb, bits, kk, a_h)
                np.testing.assert_array_equal(to_np(ops.binary_scalar(op, L, a, k, scalar_first=True)), want_first,
                                              err_msg=f"{L} {op} k={k:#x} first")
                if ref is not None and ref.has(eb, bits) and j == 3:
                    np.testing.assert_array_equal(want, ref.binary(op, eb, bits, a_h, kk))
                    np.testing.assert_array_equal(want_first, ref.binary(op, eb, bits, kk, a_h))


@pytest.mark.parametrize("eb,bits", SPOT)
def test_pairs_large_and_alignments(P, oracle, eb, bits, mode):
    """Whole tiles, partial tile, vector body with scalar head and tail, and operand addresses that
    force the element-granular kernel (pairs not reaching a 32-byte boundary where out reaches 16)."""
    pint_b200, ops = P
    L = pint_b200.Layout(eb, *bits)
    n = (1 << 18) + 4 * 512 + 7
    flat = oracle.fill(eb, 2 * n + 128, 0xA11CE)
    dev = to_dev(flat)
    obuf = torch.zeros(n + 128, dtype=dev.dtype, device="cuda")
    per = 16 // eb
    for p_off, o_off in ((0, 0), (2, 1), (2 * (per - 1), per - 1), (1, 0), (0, 1), (2 * per, per), (3, 5)):
        pairs_h = flat[p_off:p_off + 2 * n]
        for op in ("add_unsigned_saturate", "add_signed_saturate", "sub_signed_saturate", "sub_wrap", "min_signed", "bit_xor"):
            want = oracle.binary(op, eb, bits, pairs_h[0::2], pairs_h[1::2])
            out = obuf[o_off:o_off + n]
            before = ops.launch_count()
            ops.binary_pairs(op, L, dev[p_off:p_off + 2 * n], out=out)
            assert ops.launch_count() > before
            np.testing.assert_array_equal(to_np(out), want, err_msg=f"{L} {op} offsets {p_off},{o_off}")


@pytest.mark.parametrize("eb,bits", SPOT)
def test_scalar_large_alignments_in_place(P, oracle, eb, bits, mode):
    pint_b200, ops = P
    L = pint_b200.Layout(eb, *bits)
    n = (1 << 18) + 4 * 1024 + 5
    a_h = oracle.fill(eb, n + 64, 0x5CA1A)
    dev = to_dev(a_h)
    obuf = torch.zeros(n + 64, dtype=dev.dtype, device="cuda")
    ks = [int(x) for x in oracle.fill(eb, 2, 0xC0FFEE)] + [0, (1 << (8 * eb)) - 1]
    for i_off, o_off in ((0, 0), (1, 1), (3, 3 + 16 // eb), (1, 0)):
        src_h = a_h[i_off:i_off + n]
        for k in ks[:2] if (i_off, o_off) != (0, 0) else ks:
            kk = np.full_like(src_h, k)
            for op in ("add_unsigned_saturate", "add_signed_saturate", "sub_unsigned_saturate", "sub_wrap", "max_signed", "bit_and"):
                for first in (False, True):
                    want = oracle.binary(op, eb, bits, kk, src_h) if first else oracle.binary(op, eb, bits, src_h, kk)
                    out = obuf[o_off:o_off + n]
                    ops.binary_scalar(op, L, dev[i_off:i_off + n], k, out=out, scalar_first=first)
                    np.testing.assert_array_equal(to_np(out), want, err_msg=f"{L} {op} k={k:#x} first={first} {i_off},{o_off}")
    # in place
    x = to_dev(a_h[:n])
    k = ks[0]
    ops.binary_scalar("sub_signed_saturate", L, x, k, out=x, scalar_first=True)
    np.testing.assert_array_equal(to_np(x), oracle.binary("sub_signed_saturate", eb, bits, np.full_like(a_h[:n], k), a_h[:n]))


def test_callers_small_sizes(P, oracle):
    pint_b200, ops = P
    for eb, bits in SPOT[:4]:
        L = pint_b200.Layout(eb, *bits)
        for n in list(range(0, 40)) + [63, 64, 65, 511, 513]:
            a_h, b_h = oracle.fill(eb, n, 21), oracle.fill(eb, n, 22)
            got = ops.binary_pairs("add_signed_saturate", L, to_dev(interleave(a_h, b_h)) if n else
                                   torch.zeros(0, dtype=to_dev(a_h[:0]).dtype, device="cuda"))
            np.testing.assert_array_equal(to_np(got), oracle.binary("add_signed_saturate", eb, bits, a_h, b_h))
            k = 0x5A5A5A5A5A5A5A5A & ((1 << (8 * eb)) - 1)
            got = ops.binary_scalar("sub_unsigned_saturate", L, to_dev(a_h), k)
            np.testing.assert_array_equal(to_np(got), oracle.binary("sub_unsigned_saturate", eb, bits, a_h, np.full_like(a_h, k)))


def test_callers_error_codes(P):
    pint_b200, ops = P
    from pint_b200 import _lib
    lib = _lib.lib()
    L = pint_b200.Layout.make(5, 6, 5)
    x = torch.zeros(64, dtype=torch.int16, device="cuda")
    p = x.data_ptr()
    assert lib.pint_cuda_binary_scalar(0, L.ref, p, 0x1FFFF, 0, p, 16, None) == -15   # not a uint16_t value
    assert lib.pint_cuda_binary_scalar(13, L.ref, p, 1, 0, p, 16, None) == -10
    assert lib.pint_cuda_binary_scalar(0, L.ref, None, 1, 0, p, 16, None) == -1
    assert lib.pint_cuda_binary_scalar(0, L.ref, p, 1, 0, p + 2, 16, None) == -14
    assert lib.pint_cuda_binary_scalar(0, L.ref, p, 1, 0, p, 0, None) == 0
    assert lib.pint_cuda_binary_pairs(-1, L.ref, p, p + 64, 16, None) == -10
    assert lib.pint_cuda_binary_pairs(0, L.ref, p, None, 16, None) == -1
    assert lib.pint_cuda_binary_pairs(0, L.ref, p, p, 16, None) == -14               # results over the pairs
    assert lib.pint_cuda_binary_pairs(0, L.ref, p, p + 62, 16, None) == -14
    assert lib.pint_cuda_binary_pairs(0, L.ref, p, p + 64, 16, None) == 0
    assert lib.pint_cuda_binary_pairs(0, L.ref, p + 1, p + 64, 8, None) == -15       # not Integer-aligned
    bad = _lib.PintLayout()
    bad.elem_bytes, bad.nfields = 2, 2
    bad.bits[0], bad.bits[1] = 9, 8
    assert lib.pint_cuda_binary_pairs(0, C.byref(bad), p, p + 64, 16, None) == -5
    assert lib.pint_cuda_binary_scalar(0, C.byref(bad), p, 0, 0, p, 16, None) == -5


def test_callers_in_cuda_graph(P, oracle):
    """Both entry points are plain asynchronous launches on the caller's stream: capturable, replayable."""
    pint_b200, ops = P
    eb, bits = 2, (5, 6, 5)
    L = pint_b200.Layout(eb, *bits)
    n = (1 << 16) + 11
    a_h, b_h = oracle.fill(eb, n, 41), oracle.fill(eb, n, 42)
    pairs = to_dev(interleave(a_h, b_h))
    t1 = torch.empty(n, dtype=pairs.dtype, device="cuda")
    out = torch.empty_like(t1)
    k = 0x1234
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    g = torch.cuda.CUDAGraph()
    with torch.cuda.stream(s):
        ops.binary_pairs("add_wrap", L, pairs, out=t1)  # warm-up outside capture
        torch.cuda.synchronize()
        with torch.cuda.graph(g, stream=s):
            ops.binary_pairs("add_signed_saturate", L, pairs, out=t1)
            ops.binary_scalar("sub_unsigned_saturate", L, t1, k, out=out)
    want = oracle.binary("sub_unsigned_saturate", eb, bits, oracle.binary("add_signed_saturate", eb, bits, a_h, b_h),
                         np.full_like(a_h, k))
    for _ in range(3):
        out.zero_()
        g.replay()
        torch.cuda.synchronize()
        np.testing.assert_array_equal(to_np(out), want)


@pytest.mark.parametrize("eb", [1, 2, 4, 8])
def test_equal_matches_operator_eq(P, oracle, eb):
    """packed_int::operator== / != (pint.hpp:778-779) compare the whole Integer, unused high bits
    included: flags and the not-equal count against numpy on arrays that differ in scattered places
    (some only above sum(Bits)); vector and element-granular paths, flags / count / both."""
    pint_b200, ops = P
    bits = {1: (2, 3), 2: (5, 6, 4), 4: (8, 8, 8), 8: (3, 7, 11, 13, 17)}[eb]   # all leave unused high bits
    L = pint_b200.Layout(eb, *bits)
    rng = np.random.default_rng(eb)
    for n in (0, 1, 5, 16, 17, 1023, 4096, (1 << 18) + 4 * 1024 + 3):
        a_h = oracle.fill(eb, n + 24, 77)
        b_h = a_h.copy()
        if n:
            idx = rng.choice(n + 24, size=max(1, (n + 24) // 7), replace=False)
            b_h[idx] ^= (DT[eb](1) << DT[eb](8 * eb - 1 - (idx % 3)))   # top bits: outside every field
            idx2 = rng.choice(n + 24, size=max(1, (n + 24) // 11), replace=False)
            b_h[idx2] ^= DT[eb](1)
        da, db = to_dev(a_h), to_dev(b_h)
        for off in (0, 1, 16 // eb):
            x_h, y_h = a_h[off:off + n], b_h[off:off + n]
            want = (x_h == y_h).astype(np.uint8)
            f, c = ops.equal(L, da[off:off + n], db[off:off + n])
            np.testing.assert_array_equal(f.cpu().numpy(), want, err_msg=f"eb={eb} n={n} off={off}")
            assert c == int((want == 0).sum())
            f2, c2 = ops.equal(L, da[off:off + n], db[off:off + n], count=False)
            assert c2 is None
            np.testing.assert_array_equal(f2.cpu().numpy(), want)
            f3, c3 = ops.equal(L, da[off:off + n], db[off:off + n], flags=False)
            assert f3 is None and c3 == int((want == 0).sum())
            # flags at an odd address: element-granular path
            buf = torch.full((n + 5,), 7, dtype=torch.uint8, device="cuda")
            ops.equal(L, da[off:off + n], db[off:off + n], out=buf[3:3 + n], count=False)
            got = buf.cpu().numpy()
            np.testing.assert_array_equal(got[3:3 + n], want)
            assert (got[:3] == 7).all() and (got[3 + n:] == 7).all()
        f, c = ops.equal(L, da[:n], da[:n])
        assert c == 0 and bool((f == 1).all())


def test_equal_error_codes(P):
    pint_b200, ops = P
    from pint_b200 import _lib
    lib = _lib.lib()
    L = pint_b200.Layout.make(5, 6, 5)
    x = torch.zeros(64, dtype=torch.int16, device="cuda")
    f = torch.zeros(64, dtype=torch.uint8, device="cuda")
    c = torch.zeros(1, dtype=torch.int64, device="cuda")
    p = x.data_ptr()
    assert lib.pint_cuda_equal(L.ref, p, p, 16, None, None, None) == -1
    assert lib.pint_cuda_equal(L.ref, None, p, 16, f.data_ptr(), None, None) == -1
    assert lib.pint_cuda_equal(L.ref, p, p + 1, 16, f.data_ptr(), None, None) == -15
    assert lib.pint_cuda_equal(L.ref, p, p, 16, p + 8, None, None) == -14
    assert lib.pint_cuda_equal(L.ref, p, p, 0, None, c.data_ptr(), None) == 0
    assert lib.pint_cuda_equal(L.ref, p, p, 16, f.data_ptr(), c.data_ptr(), None) == 0
