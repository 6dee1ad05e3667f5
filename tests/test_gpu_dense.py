"""GPU parity of the dense bit-stream format (pint_cuda_dense_*), through the C ABI, against the
oracle (dense op == pack(op(unpack)) with the reference-pinned per-element op).  Bit-exact."""
import ctypes as C

import numpy as np
import pytest
import torch

from oracle import OPS

pytestmark = pytest.mark.gpu

from gpu_util import to_dev, to_np  # noqa: E402

# static dense kernels exist for the first nine (pint_b200/csrc/dense_layouts.def); the rest run the
# run-time-geometry kernel; S in {8,16,32,64} is the padded format itself
LAYOUTS = [(4, (8, 8, 8)), (4, (10, 10, 10)), (2, (12,)), (4, (7, 9, 3)), (8, (16, 16, 16)),
           (2, (14,)), (2, (5, 5, 5)), (4, (6, 6, 6)), (2, (10,)),
           (4, (5, 6, 5, 3)), (8, (10, 10, 10, 10)), (8, (21, 21, 21)), (1, (3,)), (4, (1,)), (8, (31, 2)), (8, (1, 62)),
           (8, (3, 7, 11, 13, 17, 12)), (4, (11, 11, 9)), (2, (5, 6, 5)), (8, (64,)), (4, (8, 8, 8, 8)), (1, (2, 3, 3)),
           (8, (12, 12, 12)), (8, (17, 17)), (1, (2, 2)), (2, (3, 4)), (4, (13, 13)), (8, (19, 19, 19)),
           (4, (8, 8)), (8, (16, 16)), (8, (4, 4))]  # the last three: S = 16 / 32 / 8 inside a wider Integer
SIZES = [1, 2, 5, 127, 128, 129, 4096, (1 << 18) + 77]  # the last: many whole tiles + a remainder


@pytest.fixture(scope="module")
def P():
    import pint_b200
    from pint_b200 import dense, ops
    assert torch.cuda.is_available()
    return pint_b200, ops, dense


# compiled-in kernels where they exist / run-time masks (whole-chunk kernels where the chunk is short
# enough) / run-time masks with the element-by-element window kernels forced
@pytest.fixture(params=["static_or_generic", "force_generic", "force_generic_windows"])
def mode(request, P):
    _, ops, _ = P
    ops.set_option("force_generic", 0 if request.param == "static_or_generic" else 1)
    ops.set_option("dense_windows", 1 if request.param == "force_generic_windows" else 0)
    yield request.param
    ops.set_option("force_generic", 0)
    ops.set_option("dense_windows", 0)


def dev_bytes(x):
    return torch.from_numpy(np.ascontiguousarray(x).copy()).cuda()


@pytest.mark.parametrize("eb,bits", LAYOUTS)
def test_dense_binary_vs_oracle(P, oracle, eb, bits, mode):
    pint_b200, ops, dense = P
    L = pint_b200.Layout(eb, *bits)
    for n in SIZES:
        x, y = oracle.fill(eb, n, 0xD0 + n), oracle.fill(eb, n, 0xE0 + n)
        dx_h, dy_h = oracle.dense_pack(eb, bits, x), oracle.dense_pack(eb, bits, y)
        assert dense.nbytes(L, n) == dx_h.size
        if (n * sum(bits)) % 128:
            dx_h[-1] |= 0x80  # garbage above n*S must not reach the output
        dx, dy = dev_bytes(dx_h), dev_bytes(dy_h)
        for op in OPS:
            got = dense.binary(op, L, dx, dy, n).cpu().numpy()
            np.testing.assert_array_equal(got, oracle.dense_binary(op, eb, bits, dx_h, dy_h, n),
                                          err_msg=f"{L} {op} n={n} {mode}")


@pytest.mark.parametrize("eb,bits", LAYOUTS)
def test_dense_shift_vs_oracle(P, oracle, eb, bits, mode):
    """shift_left / shift_right_unsigned on a dense stream == pack(shift(unpack)), every amount class:
    0, inside the narrowest field, between field widths, >= the widest field, beyond the word."""
    pint_b200, ops, dense = P
    L = pint_b200.Layout(eb, *bits)
    S = sum(bits)
    amounts = sorted({0, 1, 2, min(bits) - 1, min(bits), max(bits) - 1, max(bits), max(bits) + 1, 8 * eb, 8 * eb + 5} - {-1})
    for n in (1, 5, 129, 4096, (1 << 17) + 77):
        x = oracle.fill(eb, n, 0xA5 + n)
        x = x & np.array((1 << S) - 1, dtype=x.dtype)
        dx_h = oracle.dense_pack(eb, bits, x)
        if (n * S) % 128:
            dx_h[-1] |= 0x80
        dx = dev_bytes(dx_h)
        for left in (True, False):
            for amt in amounts:
                got = dense.shift(left, L, dx, amt, n).cpu().numpy()
                want = oracle.dense_pack(eb, bits, oracle.shift(left, eb, bits, amt, x))
                np.testing.assert_array_equal(got, want, err_msg=f"{L} left={left} amount={amt} n={n} {mode}")
        inplace = dev_bytes(dx_h)
        dense.shift(True, L, inplace, 1, n, out=inplace)
        np.testing.assert_array_equal(inplace.cpu().numpy(), oracle.dense_pack(eb, bits, oracle.shift(True, eb, bits, 1, x)))


@pytest.mark.parametrize("eb,bits", LAYOUTS)
def test_dense_shift_chunk_kernels_still_agree(P, oracle, eb, bits):
    """Shifts by less than 32 bits take the word-stream kernel by default (cuda_kernels_dense_shift.cuh), in place or
    not; the whole-chunk kernels they replaced still serve wide shifts, so they stay under test: option off, same
    streams, same oracle, out of place and in place."""
    pint_b200, ops, dense = P
    L = pint_b200.Layout(eb, *bits)
    S = sum(bits)
    try:
        for n in (129, (1 << 17) + 77):
            x = oracle.fill(eb, n, 0x5F + n)
            x = x & np.array((1 << S) - 1, dtype=x.dtype)
            dx_h = oracle.dense_pack(eb, bits, x)
            dx = dev_bytes(dx_h)
            for left in (True, False):
                for amt in sorted({0, 1, min(bits), max(bits) - 1}):
                    want = oracle.dense_pack(eb, bits, oracle.shift(left, eb, bits, amt, x))
                    for opt in (0, 1):
                        ops.set_option("dense_shift_stream", opt)
                        got = dense.shift(left, L, dx, amt, n).cpu().numpy()
                        np.testing.assert_array_equal(got, want, err_msg=f"{L} left={left} amount={amt} n={n} dense_shift_stream={opt}")
                        work = dev_bytes(dx_h)
                        dense.shift(left, L, work, amt, n, out=work)
                        np.testing.assert_array_equal(work.cpu().numpy(), want, err_msg=f"in place {L} left={left} amount={amt} n={n} dense_shift_stream={opt}")
    finally:
        ops.set_option("dense_shift_stream", 1)


def test_dense_shift_every_total_width(P, oracle):
    """sum(Bits) = 1 ... 63 with a seeded random split into fields: every period of the keep mask (1 ... 63 words) and
    every phase of a field against the 32-bit words the stream kernel shifts, out of place and in place."""
    pint_b200, ops, dense = P
    rng = np.random.default_rng(0x5417F7)
    for S in range(1, 64):
        cuts = sorted(set(rng.integers(1, S, size=int(rng.integers(0, 5))).tolist())) if S > 1 else []
        bits = tuple(b - a for a, b in zip([0] + cuts, cuts + [S]))
        eb = 1 if S <= 8 else 2 if S <= 16 else 4 if S <= 32 else 8
        L = pint_b200.Layout(eb, *bits)
        for n in (1, 77, 5000):
            x = oracle.fill(eb, n, S * 131 + n)
            x = x & np.array((1 << S) - 1, dtype=x.dtype)
            dx_h = oracle.dense_pack(eb, bits, x)
            dx = dev_bytes(dx_h)
            for left in (True, False):
                for amt in sorted({0, 1, min(bits), max(bits) - 1, min(max(bits), 31)}):
                    want = oracle.dense_pack(eb, bits, oracle.shift(left, eb, bits, amt, x))
                    got = dense.shift(left, L, dx, amt, n).cpu().numpy()
                    np.testing.assert_array_equal(got, want, err_msg=f"{L} left={left} amount={amt} n={n}")
            inplace = dev_bytes(dx_h)
            dense.shift(False, L, inplace, 1, n, out=inplace)
            np.testing.assert_array_equal(inplace.cpu().numpy(), oracle.dense_pack(eb, bits, oracle.shift(False, eb, bits, 1, x)))


@pytest.mark.parametrize("eb,bits", LAYOUTS[:12])
def test_dense_in_place_and_tiles(P, oracle, eb, bits):
    """In place on either operand; every tile geometry option gives the same stream."""
    pint_b200, ops, dense = P
    L = pint_b200.Layout(eb, *bits)
    n = (1 << 19) + 3
    x, y = oracle.fill(eb, n, 21), oracle.fill(eb, n, 22)
    dx_h, dy_h = oracle.dense_pack(eb, bits, x), oracle.dense_pack(eb, bits, y)
    want = oracle.dense_binary("sub_signed_saturate", eb, bits, dx_h, dy_h, n)
    for threads, tile in ((0, 0), (32, 4096), (64, 0), (256, 32768), (128, 65536)):
        ops.set_option("dense_threads", threads)
        ops.set_option("dense_tile_bytes", tile)
        try:
            dx, dy = dev_bytes(dx_h), dev_bytes(dy_h)
            dense.binary("sub_signed_saturate", L, dx, dy, n, out=dx)
            np.testing.assert_array_equal(dx.cpu().numpy(), want, err_msg=f"{L} threads={threads} tile={tile}")
            dx = dev_bytes(dx_h)
            dense.binary("sub_signed_saturate", L, dx, dy, n, out=dy)
            np.testing.assert_array_equal(dy.cpu().numpy(), want)
        finally:
            ops.set_option("dense_threads", 0)
            ops.set_option("dense_tile_bytes", 0)


@pytest.fixture(params=["compiled_in_S", "runtime_geometry"])
def convert_mode(request, P):
    """Conversions: kernels with S compiled in (dense_fixed.cu) / the run-time-geometry kernels only."""
    _, ops, _ = P
    ops.set_option("dense_windows", 1 if request.param == "runtime_geometry" else 0)
    yield request.param
    ops.set_option("dense_windows", 0)


@pytest.mark.parametrize("eb,bits", LAYOUTS)
def test_dense_pack_unpack_planes(P, oracle, eb, bits, convert_mode):
    pint_b200, ops, dense = P
    L = pint_b200.Layout(eb, *bits)
    S = sum(bits)
    for n in SIZES:
        x = oracle.fill(eb, n, 0xF0 + n)
        want = oracle.dense_pack(eb, bits, x)
        xd = to_dev(x)
        d = dense.pack(L, xd)
        np.testing.assert_array_equal(d.cpu().numpy(), want, err_msg=f"pack {L} n={n}")
        back = to_np(dense.unpack(L, d, n))
        np.testing.assert_array_equal(back, x & np.array((1 << S) - 1, dtype=x.dtype), err_msg=f"unpack {L} n={n}")
        masked = x & np.array((1 << S) - 1, dtype=x.dtype)
        for signed in (False, True):
            for pb in (4, 8):
                planes = dense.to_planes(L, d, n, signed=signed, plane_bytes=pb)
                ref_planes = oracle.unpack(eb, bits, masked, signed, pb)
                for f in range(len(bits)):
                    np.testing.assert_array_equal(to_np(planes[f], ref_planes[f].dtype), ref_planes[f],
                                                  err_msg=f"to_planes {L} f={f} signed={signed} pb={pb} n={n}")
        # planes -> dense (truncating constructor): arbitrary int32 plane values, not just in-range ones
        rng = np.random.default_rng(n)
        raw = [rng.integers(-2**31, 2**31, n, dtype=np.int64).astype(np.int32) for _ in bits]
        want2 = oracle.dense_pack(eb, bits, oracle.repack(eb, bits, raw, 4))
        got2 = dense.from_planes(L, [to_dev(p) for p in raw])
        np.testing.assert_array_equal(got2.cpu().numpy(), want2, err_msg=f"from_planes {L} n={n}")


def test_dense_round_trip_properties_large(P):
    """2^27 RGB24 elements (384 MiB per stream): pack/unpack round trip, op identities."""
    pint_b200, ops, dense = P
    L = pint_b200.Layout(4, 8, 8, 8)
    n = 1 << 27
    x = torch.empty(n, dtype=torch.uint32, device="cuda")
    y = torch.empty(n, dtype=torch.uint32, device="cuda")
    ops.fill_splitmix(x, 1)
    ops.fill_splitmix(y, 2)
    dx, dy = dense.pack(L, x), dense.pack(L, y)
    mask = torch.tensor(0xFFFFFF, dtype=torch.int64, device="cuda")
    assert torch.equal(dense.unpack(L, dx, n).view(torch.int32) , (x.view(torch.int32) & 0xFFFFFF))
    # dense op == padded op then pack
    for op in ("add_unsigned_saturate", "sub_signed_saturate", "max_signed"):
        got = dense.binary(op, L, dx, dy, n)
        want = dense.pack(L, ops.binary(op, L, x, y))
        assert torch.equal(got, want), op
    # (a - b) + b == a wherever nothing wraps differently: wrap ops are exact inverses
    d = dense.binary("sub_wrap", L, dx, dy, n)
    assert torch.equal(dense.binary("add_wrap", L, d, dy, n), dx)
    del mask


def test_dense_error_codes(P):
    pint_b200, ops, dense = P
    from pint_b200 import _lib
    lib = _lib.lib()
    L = pint_b200.Layout(4, 8, 8, 8)
    buf = torch.zeros(4096, dtype=torch.uint8, device="cuda")
    p = buf.data_ptr()
    assert lib.pint_cuda_dense_binary(0, L.ref, p + 4, p + 1024, p + 2048, 10, None) == -15   # misaligned stream
    assert lib.pint_cuda_dense_binary(99, L.ref, p, p + 1024, p + 2048, 10, None) == -10
    assert lib.pint_cuda_dense_binary(0, L.ref, p, p + 1024, p + 16, 100, None) == -14       # partial overlap
    assert lib.pint_cuda_dense_binary(0, L.ref, None, p, p, 10, None) == -1
    assert lib.pint_cuda_dense_binary(0, L.ref, p, p, p, 0, None) == 0
    assert lib.pint_cuda_dense_bytes(L.ref, 5) == 16 and lib.pint_cuda_dense_bytes(L.ref, 6) == 32
    planes = (C.c_void_p * 3)(p, p + 1024, p + 2048)
    assert lib.pint_cuda_dense_to_planes(L.ref, 0, p + 3072, planes, 3, 4, None) == -13


def test_sharded_dense_binary_single_process(P, oracle):
    """pint_cuda_sharded_dense_binary over the visible devices (the same device twice when there is one)."""
    pint_b200, ops, dense = P
    from pint_b200 import sharded
    eb, bits = 4, (5, 6, 5, 3)
    L = pint_b200.Layout(eb, *bits)
    S = sum(bits)
    n = 777_777
    ndev = torch.cuda.device_count()
    devs = [k % ndev for k in range(max(2, ndev))]
    plan = sharded.dense_plan_shards(n, len(devs))
    x, y = oracle.fill(eb, n, 31), oracle.fill(eb, n, 32)
    dx, dy = oracle.dense_pack(eb, bits, x), oracle.dense_pack(eb, bits, y)
    want = oracle.dense_binary("add_signed_saturate", eb, bits, dx, dy, n)
    a_sh, b_sh, o_sh = [], [], []
    for (beg, cnt), d in zip(plan, devs):
        lo, nb = beg * S // 8, dense.nbytes(L, cnt)
        pad = lambda v: np.concatenate([v[lo:lo + nb], np.zeros(max(0, lo + nb - v.size), dtype=np.uint8)])  # noqa: E731
        a_sh.append(torch.from_numpy(pad(dx)).to(f"cuda:{d}"))
        b_sh.append(torch.from_numpy(pad(dy)).to(f"cuda:{d}"))
        o_sh.append(torch.empty(nb, dtype=torch.uint8, device=f"cuda:{d}"))
    sharded.sharded_dense_binary("add_signed_saturate", L, a_sh, b_sh, o_sh, [c for _, c in plan])
    for d in set(devs):
        torch.cuda.synchronize(d)
    for (beg, cnt), o in zip(plan, o_sh):
        lo, valid = beg * S // 8, (cnt * S) // 8  # whole bytes of the shard (its last partial byte shares bits only with padding)
        np.testing.assert_array_equal(o.cpu().numpy()[:valid], want[lo:lo + valid])


@pytest.mark.parametrize("eb,bits", [(4, (8, 8, 8)), (4, (7, 9, 3)), (8, (21, 21, 21)), (2, (12,))])
def test_dense_conversions_misaligned_other_side(P, oracle, eb, bits, convert_mode):
    """The padded array / the planes at addresses that are not 16-byte aligned (element-granular
    fallback of the quad paths) give the same streams and values."""
    pint_b200, ops, dense = P
    L = pint_b200.Layout(eb, *bits)
    S = sum(bits)
    n = 70_001
    x = oracle.fill(eb, n, 0xAB)
    masked = x & np.array((1 << S) - 1, dtype=x.dtype)
    want = oracle.dense_pack(eb, bits, x)
    for off in (1, 3):
        xd = to_dev(x, off)
        assert xd.data_ptr() % 16 != 0
        d = dense.pack(L, xd)
        np.testing.assert_array_equal(d.cpu().numpy(), want)
        out = to_dev(np.zeros(n, dtype=x.dtype), off)
        dense.unpack(L, d, n, out=out)
        np.testing.assert_array_equal(to_np(out), masked)
        ref_planes = oracle.unpack(eb, bits, masked, True, 4)
        planes = [to_dev(np.zeros(n, dtype=np.int32), off) for _ in bits]
        assert planes[0].data_ptr() % 16 != 0
        dense.to_planes(L, d, n, signed=True, out=planes)
        for f in range(len(bits)):
            np.testing.assert_array_equal(to_np(planes[f], np.int32), ref_planes[f])
        got = dense.from_planes(L, planes)
        np.testing.assert_array_equal(got.cpu().numpy(), want)


@pytest.mark.parametrize("eb,bits", [(4, (8, 8, 8)), (2, (12,)), (4, (5, 6, 5, 3)), (8, (10, 10, 10, 10)), (4, (8, 8, 8, 8)),
                                     (8, (16, 16)), (2, (5, 6, 5)), (8, (4, 4)), (8, (16, 16, 16, 16))])  # S = 32 / 16 / 8 / 64: padded format
def test_dense_binary_multi(P, oracle, eb, bits, mode):
    """Op pairs in one pass over dense streams == the ops run one by one; lists mixing pairs and singles."""
    pint_b200, ops, dense = P
    L = pint_b200.Layout(eb, *bits)
    for n in (3, 4099, (1 << 18) + 5):
        x, y = oracle.fill(eb, n, 71), oracle.fill(eb, n, 72)
        dx_h, dy_h = oracle.dense_pack(eb, bits, x), oracle.dense_pack(eb, bits, y)
        dx, dy = dev_bytes(dx_h), dev_bytes(dy_h)
        for op_list in (["add_unsigned_saturate", "add_signed_saturate"], ["sub_signed_saturate", "add_signed_saturate"],
                        ["max_unsigned", "min_unsigned"], ["add_wrap", "bit_xor", "sub_wrap", "min_signed"], ["sub_unsigned_saturate"]):
            outs = dense.binary_multi(op_list, L, dx, dy, n)
            for op, o in zip(op_list, outs):
                np.testing.assert_array_equal(o.cpu().numpy(), oracle.dense_binary(op, eb, bits, dx_h, dy_h, n),
                                              err_msg=f"{L} {op} n={n} {mode}")
    import ctypes as C
    from pint_b200 import _lib
    codes = (C.c_int * 2)(1, 2)
    outs = (C.c_void_p * 2)(dx.data_ptr(), dy.data_ptr())
    assert _lib.lib().pint_cuda_dense_binary_multi(2, codes, L.ref, dx.data_ptr(), dy.data_ptr(), outs, n, None) == -14
