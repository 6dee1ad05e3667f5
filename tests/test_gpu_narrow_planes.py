"""GPU parity for 1- and 2-byte planes in unpack / repack (plane_bytes == sizeof(Integer) gives the
reference's own return types for 8 / 16-bit packs: get -> Integer, get_signed -> make_signed<Integer>,
pint.hpp:800, 810-811).  Against the compiled reference's 8-byte plane vectors narrowed by a C cast,
and against the oracle; bit-exact."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from gpu_util import to_dev, to_np  # noqa: E402

NDT = {(1, 0): np.uint8, (1, 1): np.int8, (2, 0): np.uint16, (2, 1): np.int16}


@pytest.fixture(scope="module")
def P():
    import pint_b200
    from pint_b200 import ops
    assert torch.cuda.is_available()
    return pint_b200, ops


@pytest.fixture(params=["static_or_generic", "force_generic"])
def mode(request, P):
    _, ops = P
    ops.set_option("force_generic", 1 if request.param == "force_generic" else 0)
    yield request.param
    ops.set_option("force_generic", 0)


def test_golden_narrow_unpack(P, refvec, oracle, mode):
    pint_b200, ops = P
    for key, eb, bits in refvec.layouts():
        L = pint_b200.Layout(eb, *bits)
        a_h = refvec[key + "_a"]
        a = to_dev(a_h)
        for signed in (0, 1):
            for pb in (1, 2):
                planes = ops.unpack(L, a, signed=bool(signed), plane_bytes=pb)
                want_o = oracle.unpack(eb, bits, a_h, bool(signed), pb)
                for f in range(len(bits)):
                    wide = refvec[f"{key}_unpack{'S' if signed else 'U'}8_{f}"]
                    want = wide.astype(NDT[(pb, signed)])           # the reference's value through a C cast
                    got = to_np(planes[f], NDT[(pb, signed)])
                    np.testing.assert_array_equal(got, want, err_msg=f"{L} f{f} s{signed} pb{pb} {mode}")
                    np.testing.assert_array_equal(want_o[f], want)


def test_narrow_repack_and_round_trip(P, refvec, oracle, mode):
    pint_b200, ops = P
    rng = np.random.default_rng(5)
    for key, eb, bits in refvec.layouts():
        L = pint_b200.Layout(eb, *bits)
        a_h = refvec[key + "_a"]
        n = a_h.size
        A = L.masks()["A"]
        for pb in (1, 2):
            sdt = NDT[(pb, 1)]
            planes_h = [rng.integers(np.iinfo(sdt).min, np.iinfo(sdt).max, n, dtype=sdt, endpoint=True) for _ in bits]
            got = to_np(ops.repack(L, [to_dev(p.view(NDT[(pb, 0)])).view({1: torch.int8, 2: torch.int16}[pb]) for p in planes_h]))
            np.testing.assert_array_equal(got, oracle.repack(eb, bits, planes_h, pb), err_msg=f"{L} pb{pb} {mode}")
            if max(bits) <= 8 * pb:
                back = to_np(ops.repack(L, ops.unpack(L, to_dev(a_h), signed=True, plane_bytes=pb)))
                np.testing.assert_array_equal(back, a_h & np.array(A).astype(back.dtype))


@pytest.mark.parametrize("eb,bits", [(2, (5, 6, 5)), (1, (2, 3, 3)), (4, (8, 8, 8, 8)), (8, (16, 16, 16, 16)), (4, (7, 9, 3))])
def test_narrow_planes_large_and_misaligned(P, oracle, eb, bits):
    """Whole quads plus a tail; planes and packed array at element offsets (element-granular kernels);
    a skipped field; sentinel bands around the planes."""
    pint_b200, ops = P
    L = pint_b200.Layout(eb, *bits)
    n = (1 << 18) + 7
    x_h = oracle.fill(eb, n + 8, 99)
    x = to_dev(x_h)
    for pb in (1, 2):
        tdt = {1: torch.int8, 2: torch.int16}[pb]
        for x_off, p_off in ((0, 0), (1, 0), (0, 1), (4, 4 if pb == 2 else 8)):
            want = oracle.unpack(eb, bits, x_h[x_off:x_off + n], True, pb, fields=[0, len(bits) - 1])
            bufs = [torch.full((n + 32,), 0x5B, dtype=tdt, device="cuda") for _ in bits]
            out = [b[p_off:p_off + n] if f in (0, len(bits) - 1) else None for f, b in enumerate(bufs)]
            ops.unpack(L, x[x_off:x_off + n], signed=True, plane_bytes=pb, out=out)
            for f in (0, len(bits) - 1):
                got = bufs[f].cpu().numpy()
                np.testing.assert_array_equal(got[p_off:p_off + n], want[f], err_msg=f"{L} pb{pb} f{f} off {x_off},{p_off}")
                assert (got[:p_off] == 0x5B).all() and (got[p_off + n:] == 0x5B).all()
            # repack from (mis)aligned narrow planes
            full = oracle.unpack(eb, bits, x_h[x_off:x_off + n], True, pb)
            src = [torch.empty(n + 32, dtype=tdt, device="cuda") for _ in bits]
            for f in range(len(bits)):
                src[f][p_off:p_off + n].copy_(torch.from_numpy(full[f].copy()))
            got = to_np(ops.repack(L, [s_[p_off:p_off + n] for s_ in src]))
            np.testing.assert_array_equal(got, oracle.repack(eb, bits, full, pb))


def test_plane_bytes_error_code(P):
    import ctypes as C
    pint_b200, ops = P
    from pint_b200 import _lib
    lib = _lib.lib()
    L = pint_b200.Layout.make(5, 6, 5)
    x = torch.zeros(16, dtype=torch.int16, device="cuda")
    planes = (C.c_void_p * 3)(x.data_ptr(), x.data_ptr(), x.data_ptr())
    for pb in (0, 3, 5, 16):
        assert lib.pint_cuda_unpack(L.ref, 0, x.data_ptr(), planes, pb, 4, None) == -13
        assert lib.pint_cuda_repack(L.ref, planes, pb, x.data_ptr(), 4, None) == -13


@pytest.mark.parametrize("eb,bits", [(4, (8, 8, 8)), (4, (7, 9, 3)), (2, (4, 4, 4)), (8, (12, 12, 12)), (8, (16, 16, 16)),
                                     (1, (2, 3, 2)), (8, (21, 21, 21)), (8, (10, 10, 10, 10)), (4, (5, 5, 5)), (8, (11, 11, 11)),
                                     # every field one plane value wide: the byte-permute kernels (interleaved <-> planar)
                                     (8, (8, 8, 8, 8, 8)), (8, (8, 8, 8, 8, 8, 8)), (8, (8, 8, 8, 8, 8, 8, 8)),
                                     # S = 8 / 16 / 32 / 64: the stream is the padded array of the natural Integer
                                     (4, (8, 8, 8, 8)), (2, (5, 6, 5)), (8, (16, 16, 16, 16)), (1, (3, 5)), (8, (8, 8, 8, 8))])
def test_dense_narrow_planes(P, oracle, eb, bits):
    """Dense streams <-> 1- and 2-byte planes (dense RGB24 <-> three uint8 channels is interleaved <->
    planar RGB at 3 + 3 bytes per pixel): to_planes against get / get_signed of the elements narrowed by
    a C cast, from_planes against pack(field constructor); aligned planes (quads) and planes at an
    element offset (element-granular), whole tiles plus a remainder.  Aligned planes with whole tiles run the
    kernels with S compiled in (16 / plane_bytes elements per thread, one 16-byte access per plane), the
    rest the run-time-geometry kernels; both must agree with the oracle."""
    pint_b200, ops = P
    from pint_b200 import dense
    L = pint_b200.Layout(eb, *bits)
    rng = np.random.default_rng(17)
    for n in (1, 127, 4097, (1 << 17) + 77):
        x_h = oracle.fill(eb, n, 55)
        d = torch.from_numpy(oracle.dense_pack(eb, bits, x_h).copy()).cuda()
        for pb in (1, 2):
            tdt = {(1, 0): torch.uint8, (1, 1): torch.int8, (2, 0): torch.uint16, (2, 1): torch.int16}
            for signed in (0, 1):
                want = oracle.unpack(eb, bits, x_h, bool(signed), pb)
                for p_off in (0, 1):
                    bufs = [torch.zeros(n + 16, dtype=tdt[(pb, signed)], device="cuda") for _ in bits]
                    dense.to_planes(L, d, n, signed=bool(signed), plane_bytes=pb, out=[b[p_off:p_off + n] for b in bufs])
                    for f in range(len(bits)):
                        got = to_np(bufs[f], NDT[(pb, signed)])
                        np.testing.assert_array_equal(got[p_off:p_off + n], want[f], err_msg=f"{L} n={n} pb={pb} s={signed} f={f} off={p_off}")
                        assert not got[:p_off].any() and not got[p_off + n:].any()
            sdt = NDT[(pb, 1)]
            planes_h = [rng.integers(np.iinfo(sdt).min, np.iinfo(sdt).max, n, dtype=sdt, endpoint=True) for _ in bits]
            want_d = oracle.dense_pack(eb, bits, oracle.repack(eb, bits, planes_h, pb))
            for p_off in (0, 1):
                src = [torch.zeros(n + 16, dtype=tdt[(pb, 1)], device="cuda") for _ in bits]
                for f in range(len(bits)):
                    src[f][p_off:p_off + n].copy_(torch.from_numpy(planes_h[f].copy()))
                got_d = dense.from_planes(L, [s_[p_off:p_off + n] for s_ in src])
                np.testing.assert_array_equal(got_d.cpu().numpy(), want_d, err_msg=f"from_planes {L} n={n} pb={pb} off={p_off}")


def test_unpack_bulk_store_kernel_matches(oracle):
    """Option "unpack_bulk" (the default): int32 planes assembled in shared memory and written by bulk (TMA) stores
    (unpack_bulk_kernel), whole 2048-element tiles + remainder through the register kernels.  Every
    layout class (compiled-in / run-time fields, 1 ... 12 fields, all Integer sizes), signed and
    unsigned, skipped planes, ragged sizes, against the oracle and against the default kernels."""
    import pint_b200
    from pint_b200 import ops
    cases = [(8, (3, 7, 11, 13, 17, 13)), (4, (8, 8, 8, 8)), (2, (5, 6, 5)), (1, (2, 3, 3)), (4, (7, 9, 3)),
             (8, (16, 16, 16, 16)), (8, (1, 63)), (4, (1, 2, 3, 4, 5, 6, 11)), (8, (5,) * 12), (8, (4,) * 13), (8, (33, 31))]
    try:
        for eb, bits in cases:
            L = pint_b200.Layout(eb, *bits)
            for n in (2048, 2048 * 5 + 1, 2048 * 37 + 1029, 100000):
                x_h = oracle.fill(eb, n, 0xB01C + n)
                x = to_dev(x_h)
                for signed in (True, False):
                    want = oracle.unpack(eb, bits, x_h, signed, 4)
                    ops.set_option("unpack_bulk", 0)
                    base = ops.unpack(L, x, signed=signed, plane_bytes=4)
                    ops.set_option("unpack_bulk", 1)
                    before = ops.launch_count()
                    got = ops.unpack(L, x, signed=signed, plane_bytes=4)
                    assert ops.launch_count() - before >= 1
                    for f in range(len(bits)):
                        np.testing.assert_array_equal(to_np(got[f], want[f].dtype), want[f], err_msg=f"{L} n={n} f={f} signed={signed}")
                        assert torch.equal(got[f], base[f])
                # only some planes wanted: the others are never touched
                fields = [0, len(bits) - 1]
                sent = [torch.full((n,), 0x5A5A5A5A, dtype=torch.int32, device="cuda") for _ in bits]
                ops.unpack(L, x, signed=True, plane_bytes=4, fields=fields, out=[sent[f] if f in fields else None for f in range(len(bits))])
                want = oracle.unpack(eb, bits, x_h, True, 4)
                for f in range(len(bits)):
                    if f in fields:
                        np.testing.assert_array_equal(to_np(sent[f], np.int32), want[f])
                    else:
                        assert bool((sent[f] == 0x5A5A5A5A).all())
    finally:
        ops.set_option("unpack_bulk", 1)
