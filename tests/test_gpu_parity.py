"""GPU parity: the CUDA path, called through the C ABI (libpint_cuda.so), against
the committed reference vectors, the reference's own known-answer tests and the
oracle.  Integer/byte work: every comparison is bit-exact (tolerance 0)."""
import numpy as np
import pytest
import torch

from oracle import OPS

pytestmark = pytest.mark.gpu

from gpu_util import to_dev, to_np  # noqa: E402

MASK64 = (1 << 64) - 1


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


def test_library_is_cuda_only(P):
    pint_b200, ops = P
    L = pint_b200.Layout.make(5, 6, 5)
    with pytest.raises(TypeError):
        ops.add_wrap(L, torch.zeros(4, dtype=torch.int16), torch.zeros(4, dtype=torch.int16))
    before = ops.launch_count()
    x = torch.zeros(1024, dtype=torch.int16, device="cuda")
    ops.add_wrap(L, x, x)
    assert ops.launch_count() > before


def test_golden_binary_all_layouts(P, refvec, mode):
    """All 13 binary ops x 33 layouts, SWAR and fused kernels, vs the unmodified reference header."""
    pint_b200, ops = P
    for key, eb, bits in refvec.layouts():
        L = pint_b200.Layout(eb, *bits)
        a, b = to_dev(refvec[key + "_a"]), to_dev(refvec[key + "_b"])
        for op in OPS:
            want = refvec[f"{key}_{op}"]
            got = to_np(ops.binary(op, L, a, b))
            np.testing.assert_array_equal(got, want, err_msg=f"swar {L} {op} {mode}")
            got = to_np(ops.fused_unpack_arith_repack(op, L, a, b))
            np.testing.assert_array_equal(got, want, err_msg=f"fused {L} {op} {mode}")


def test_golden_shift_slice(P, refvec):
    pint_b200, ops = P
    for key, eb, bits in refvec.layouts():
        L = pint_b200.Layout(eb, *bits)
        a = to_dev(refvec[key + "_a"])
        for left in (0, 1):
            for amt in range(0, max(bits) + 2):
                k = f"{key}_shift{'L' if left else 'R'}{amt}"
                if not refvec.has(k):
                    continue  # reference UB case, see oracle/pint_oracle.c
                f = ops.shift_left if left else ops.shift_right_unsigned
                np.testing.assert_array_equal(to_np(f(L, a, amt)), refvec[k], err_msg=k)
        for s in range(len(bits)):
            for e in range(s + 1, len(bits) + 1):
                out, RL = ops.slice(L, s, e, a)
                assert RL.bits == tuple(bits[s:e]) and RL.elem_bytes == eb
                np.testing.assert_array_equal(to_np(out), refvec[f"{key}_slice{s}_{e}"])


def test_golden_unpack_repack(P, refvec, mode):
    pint_b200, ops = P
    for key, eb, bits in refvec.layouts():
        L = pint_b200.Layout(eb, *bits)
        a = to_dev(refvec[key + "_a"])
        for signed in (0, 1):
            for pb in (4, 8):
                planes = ops.unpack(L, a, signed=bool(signed), plane_bytes=pb)
                for f in range(len(bits)):
                    want = refvec[f"{key}_unpack{'S' if signed else 'U'}{pb}_{f}"]
                    np.testing.assert_array_equal(to_np(planes[f], want.dtype), want, err_msg=f"{L} f{f} s{signed} pb{pb}")
        planes4 = [to_dev(refvec[f"{key}_planes4_{f}"]) for f in range(len(bits))]
        np.testing.assert_array_equal(to_np(ops.repack(L, planes4)), refvec[f"{key}_repack4"], err_msg=str(L))
        # repack(unpack(x)) == x & all_ones(sum(Bits)) for both plane widths (w <= 32 for 4-byte planes)
        A = L.masks()["A"]
        for pb in (4, 8):
            if pb == 4 and max(bits) > 32:
                continue
            back = to_np(ops.repack(L, ops.unpack(L, a, signed=True, plane_bytes=pb)))
            np.testing.assert_array_equal(back, refvec[key + "_a"] & np.array(A).astype(back.dtype))


def test_golden_pairs_reduce(P, refvec):
    pint_b200, ops = P
    for key, eb, bits in refvec.layouts():
        L = pint_b200.Layout(eb, *bits)
        pairs = to_dev(np.stack([refvec[key + "_a"], refvec[key + "_b"]], axis=1).reshape(-1))
        want = refvec[f"{key}_pairsum"]
        for generic in (0, 1):   # compiled-in masks where the layout is registered / run-time masks (bucketed smear steps)
            ops.set_option("force_generic", generic)
            try:
                for i, op in enumerate(OPS):
                    assert ops.pairs_reduce(op, L, pairs) == int(want[i]), (L, op, generic)
            finally:
                ops.set_option("force_generic", 0)


@pytest.mark.parametrize("eb,bits", [(4, (8, 8, 8, 8)), (2, (5, 6, 5)), (8, (3, 7, 11, 13, 17, 13)), (1, (2, 3, 3))])
def test_pairs_reduce_large_and_alignments(P, oracle, eb, bits):
    """Whole 256-bit tiles, 16-byte tiles, vector and scalar tails; arrays starting 0 / 16 / eb bytes
    past a 32-byte boundary; both load widths."""
    pint_b200, ops = P
    L = pint_b200.Layout(eb, *bits)
    n = (1 << 19) + 13
    flat = oracle.fill(eb, 2 * n + 64, 0xBEEF)
    dev = to_dev(flat)
    try:
        for off in (0, 16 // eb, 1):
            pairs_h = flat[off:off + 2 * n]
            for op in ("add_signed_saturate", "sub_wrap", "max_unsigned"):
                want = int(oracle.pairs_reduce(op, eb, bits, pairs_h))
                for vec256 in (0, 1):
                    ops.set_option("vec256", vec256)
                    assert ops.pairs_reduce(op, L, dev[off:off + 2 * n]) == want, (L, op, off, vec256)
    finally:
        ops.set_option("vec256", 1)


def run_kats_through_cuda(P, kats):
    pint_b200, ops = P
    ran = 0

    def pack(layout, fields):
        L = pint_b200.Layout(layout["elem_bytes"], *layout["bits"])
        planes = [to_dev(np.array([f & MASK64], dtype=np.uint64)) for f in fields]
        return L, ops.repack(L, planes)

    def val(t):
        return int(to_np(t)[0])

    for k in kats:
        kind = k["kind"]
        if kind == "storage":
            assert pint_b200.storage_bytes(k["total_bits"]) == k["elem_bytes"]
        elif kind == "make_truncate":
            L, v = pack(k["layout"], k["fields"])
            assert val(v) == k["expected_value"], k
        elif kind == "get":
            L, v = pack(k["layout"], k["fields"])
            got = to_np(ops.unpack(L, v, signed=k["signed"], plane_bytes=8)[k["index"]], np.int64 if k["signed"] else np.uint64)[0]
            if k.get("negated"):
                assert int(got) != k["expected"], k
            else:
                assert int(got) == k["expected"], k
        elif kind == "slice":
            L, v = pack(k["layout"], k["fields"])
            out, RL = ops.slice(L, k["start"], k["end"], v)
            _, want = pack(k["expected_layout"], k["expected_fields"])
            assert RL.bits == tuple(k["expected_layout"]["bits"])
            assert val(out) == val(want), k
        elif kind == "shift":
            L, v = pack(k["layout"], k["fields"])
            _, want = pack(k["layout"], k["expected_fields"])
            f = ops.shift_left if k["left"] else ops.shift_right_unsigned
            assert val(f(L, v, k["amount"])) == val(want), k
        elif kind == "binary":
            L, a = pack(k["layout"], k["a_fields"])
            _, b = pack(k["layout"], k["b_fields"])
            _, want = pack(k["layout"], k["expected_fields"])
            assert val(ops.binary(k["op"], L, a, b)) == val(want), k
            assert val(ops.fused_unpack_arith_repack(k["op"], L, a, b)) == val(want), k
        ran += 1
    return ran


def test_reference_kats_through_cuda(P, oracle, kats):
    """The reference's own 61 unit tests (tests/pint_test.cpp), executed by the CUDA kernels."""
    assert run_kats_through_cuda(P, kats) == 78


def test_readme_examples_through_cuda(P, readme_kats):
    """The worked examples of the reference's README (lines 100-410), executed by the CUDA kernels."""
    assert run_kats_through_cuda(P, readme_kats) == 18


LAYOUTS_BIG = [(2, (5, 6, 5)), (4, (8, 8, 8, 8)), (8, (16, 16, 16, 16)), (8, (3, 7, 11, 13, 17, 13)),
               (4, (1, 2, 3, 4, 5, 6, 11)), (1, (1, 1, 1)), (4, (10, 10, 10)), (8, (64,)), (1, (2, 3, 3)),
               (8, (1, 63)), (2, (7, 2, 3)), (4, (9, 5, 5, 1, 1, 1, 3, 2))]


@pytest.mark.parametrize("eb,bits", LAYOUTS_BIG)
def test_oracle_differential_ragged(P, oracle, eb, bits, mode):
    """1M+ elements, sizes that are not multiples of anything, misaligned views, in place."""
    pint_b200, ops = P
    L = pint_b200.Layout(eb, *bits)
    n = (1 << 20) + 37
    a_h, b_h = oracle.fill(eb, n, 0xA11CE), oracle.fill(eb, n, 0xB0B)
    for off in (0, 3):
        a, b = to_dev(a_h, off), to_dev(b_h, off)
        for op in OPS:
            want = oracle.binary(op, eb, bits, a_h, b_h)
            np.testing.assert_array_equal(to_np(ops.binary(op, L, a, b)), want, err_msg=f"{L} {op} off{off}")
        for op in ("add_signed_saturate", "sub_unsigned_saturate", "max_signed"):
            np.testing.assert_array_equal(to_np(ops.fused_unpack_arith_repack(op, L, a, b)),
                                          oracle.binary(op, eb, bits, a_h, b_h))
    # operands at different misalignments (element-granular kernel)
    a, b = to_dev(a_h, 1), to_dev(b_h, 2)
    np.testing.assert_array_equal(to_np(ops.add_unsigned_saturate(L, a, b)),
                                  oracle.binary("add_unsigned_saturate", eb, bits, a_h, b_h))
    # in place on a, then on b
    a, b = to_dev(a_h), to_dev(b_h)
    ops.sub_signed_saturate(L, a, b, out=a)
    np.testing.assert_array_equal(to_np(a), oracle.binary("sub_signed_saturate", eb, bits, a_h, b_h))
    a = to_dev(a_h)
    ops.add_wrap(L, a, b, out=b)
    np.testing.assert_array_equal(to_np(b), oracle.binary("add_wrap", eb, bits, a_h, b_h))
    # unpack / repack on ragged, misaligned input
    x = to_dev(a_h, 5)
    planes = ops.unpack(L, x, signed=True, plane_bytes=8)
    want = oracle.unpack(eb, bits, a_h, True, 8)
    for f in range(len(bits)):
        np.testing.assert_array_equal(to_np(planes[f], np.int64), want[f])
    np.testing.assert_array_equal(to_np(ops.repack(L, planes)), oracle.repack(eb, bits, want, 8))
    # shifts
    for amt in (0, 1, max(bits) - 1, max(bits)):
        if amt == 0 and len(bits) == 1 and bits[0] == 8 * eb and eb >= 4:
            continue
        np.testing.assert_array_equal(to_np(ops.shift_left(L, x, amt)), oracle.shift(1, eb, bits, amt, a_h))
        np.testing.assert_array_equal(to_np(ops.shift_right_unsigned(L, x, amt)), oracle.shift(0, eb, bits, amt, a_h))


def test_binary_multi_one_pass(P, oracle, mode):
    """Several ops on the same operands in one kernel == the ops run one by one."""
    import ctypes as C
    pint_b200, ops = P
    for eb, bits in [(4, (8, 8, 8, 8)), (2, (5, 6, 5)), (8, (3, 7, 11, 13, 17, 13)), (1, (2, 3, 3)), (4, (7, 9, 3))]:
        L = pint_b200.Layout(eb, *bits)
        for n, off in ((1 << 20) + 13, 0), (4099, 1), (3, 0):
            a_h, b_h = oracle.fill(eb, n, 61), oracle.fill(eb, n, 62)
            a, b = to_dev(a_h, off), to_dev(b_h, off)
            op_lists = [["add_unsigned_saturate", "add_signed_saturate"], list(OPS[:4]), ["max_signed"],
                        ["bit_xor", "sub_signed_saturate", "min_unsigned"],
                        # every pair that has a one-pass kernel, in the swapped order too, and lists that mix pairs and singles
                        ["sub_signed_saturate", "sub_unsigned_saturate"], ["add_signed_saturate", "sub_signed_saturate"],
                        ["sub_unsigned_saturate", "add_unsigned_saturate"], ["add_wrap", "sub_wrap"],
                        ["max_unsigned", "min_unsigned"], ["min_signed", "max_signed"],
                        ["add_signed_saturate", "bit_and", "add_unsigned_saturate", "sub_signed_saturate"],
                        ["min_signed", "min_unsigned", "max_signed", "max_unsigned"]]
            for op_list in op_lists:
                outs = [to_dev(np.zeros(n, dtype=a_h.dtype), off) for _ in op_list]
                ops.binary_multi(op_list, L, a, b, outs=outs)
                for op, o in zip(op_list, outs):
                    np.testing.assert_array_equal(to_np(o), oracle.binary(op, eb, bits, a_h, b_h), err_msg=f"{L} {op} n={n} {mode}")
    L = pint_b200.Layout(4, 8, 8, 8, 8)
    x = torch.zeros(64, dtype=torch.int32, device="cuda")
    y = torch.zeros(64, dtype=torch.int32, device="cuda")
    from pint_b200 import _lib
    codes = (C.c_int * 2)(1, 2)
    outs = (C.c_void_p * 2)(x.data_ptr(), y.data_ptr())
    # an output may not overlap an input or another output
    assert _lib.lib().pint_cuda_binary_multi(2, codes, L.ref, x.data_ptr(), y.data_ptr(), outs, 64, None) == -14
    codes5 = (C.c_int * 5)(0, 1, 2, 3, 4)
    assert _lib.lib().pint_cuda_binary_multi(5, codes5, L.ref, x.data_ptr(), y.data_ptr(), outs, 64, None) == -15


def test_small_sizes_and_empty(P, oracle):
    pint_b200, ops = P
    L = pint_b200.Layout(4, 8, 8, 8, 8)
    for n in (0, 1, 2, 3, 4, 5, 7, 8, 15, 16, 17, 31, 33, 255, 1023, 1025, 4097):
        a_h, b_h = oracle.fill(4, n, 1), oracle.fill(4, n, 2)
        a, b = to_dev(a_h), to_dev(b_h)
        for op in ("add_unsigned_saturate", "add_signed_saturate"):
            np.testing.assert_array_equal(to_np(ops.binary(op, L, a, b)), oracle.binary(op, 4, (8, 8, 8, 8), a_h, b_h))
        if n:
            planes = ops.unpack(L, a, signed=False)
            np.testing.assert_array_equal(to_np(ops.repack(L, planes)), a_h)


def test_error_codes(P):
    """The reference's compile-time rejections (pint.hpp:761-764, 748, 802) become error returns."""
    import ctypes as C
    pint_b200, ops = P
    from pint_b200 import _lib
    lib = _lib.lib()
    bad = _lib.PintLayout()
    bad.elem_bytes, bad.nfields = 2, 2
    bad.bits[0], bad.bits[1] = 9, 8
    x = torch.zeros(16, dtype=torch.int16, device="cuda")
    assert lib.pint_cuda_binary(0, C.byref(bad), x.data_ptr(), x.data_ptr(), x.data_ptr(), 16, None) == -5
    bad.elem_bytes = 3
    assert lib.pint_cuda_layout_validate(C.byref(bad)) == -2
    L = pint_b200.Layout.make(5, 6, 5)
    assert lib.pint_cuda_binary(99, L.ref, x.data_ptr(), x.data_ptr(), x.data_ptr(), 16, None) == -10
    assert lib.pint_cuda_slice(L.ref, 2, 2, x.data_ptr(), x.data_ptr(), 16, None) == -12
    assert lib.pint_cuda_slice(L.ref, 0, 4, x.data_ptr(), x.data_ptr(), 16, None) == -12
    assert lib.pint_cuda_binary(0, L.ref, None, x.data_ptr(), x.data_ptr(), 16, None) == -1
    assert lib.pint_cuda_binary(0, L.ref, x.data_ptr(), x.data_ptr(), x.data_ptr() + 2, 8, None) == -14  # partial overlap
    planes = (C.c_void_p * 3)(x.data_ptr(), x.data_ptr(), x.data_ptr())
    assert lib.pint_cuda_unpack(L.ref, 0, x.data_ptr(), planes, 3, 4, None) == -13
    with pytest.raises(ValueError):
        pint_b200.Layout(2, 9, 8)
    with pytest.raises(ValueError):
        pint_b200.Layout.make(33, 32)


def test_fill_and_checksum_match_oracle(P, oracle):
    pint_b200, ops = P
    for eb in (1, 2, 4, 8):
        n = 100003
        x = torch.empty(n, dtype={1: torch.uint8, 2: torch.uint16, 4: torch.uint32, 8: torch.uint64}[eb], device="cuda")
        ops.fill_splitmix(x, 0x5EED0001, start=12345)
        want = oracle.fill(eb, n, 0x5EED0001, 12345)
        np.testing.assert_array_equal(to_np(x), want)
        assert ops.checksum(x) == oracle.checksum(want)
        assert ops.checksum(x[1:]) == oracle.checksum(want[1:])  # misaligned start


@pytest.mark.parametrize("vec256", [0, 2])
def test_vector_width_option_same_results(P, oracle, vec256):
    """16-byte-only and 32-byte-everywhere kernel selection (option "vec256") are bit-identical to the
    oracle: single ops, op pairs and the fused kernel, aligned and 16-byte-but-not-32-byte-aligned operands."""
    pint_b200, ops = P
    ops.set_option("vec256", vec256)
    try:
        for eb, bits in [(4, (8, 8, 8, 8)), (8, (3, 7, 11, 13, 17, 13)), (2, (5, 6, 5)), (4, (1, 2, 3, 4, 5, 6, 11))]:
            L = pint_b200.Layout(eb, *bits)
            n = (1 << 19) + 11
            a_h, b_h = oracle.fill(eb, n, 91), oracle.fill(eb, n, 92)
            for off in (0, 16 // eb, 1):
                a, b = to_dev(a_h, off), to_dev(b_h, off)
                for op in ("add_unsigned_saturate", "sub_signed_saturate", "min_signed"):
                    want = oracle.binary(op, eb, bits, a_h, b_h)
                    np.testing.assert_array_equal(to_np(ops.binary(op, L, a, b)), want, err_msg=f"{L} {op} off{off}")
                    np.testing.assert_array_equal(to_np(ops.fused_unpack_arith_repack(op, L, a, b)), want)
                np.testing.assert_array_equal(to_np(ops.shift_left(L, a, 1)), oracle.shift(1, eb, bits, 1, a_h))
                np.testing.assert_array_equal(to_np(ops.slice(L, 1, len(bits), a)[0]), oracle.slice(eb, bits, 1, len(bits), a_h))
                # int32 planes: eight elements per thread with 256-bit plane stores (vec256 >= 1, planes on
                # 32-byte boundaries) or four with 16-byte stores (vec256 = 0, or planes only 16-byte aligned)
                for signed in (False, True):
                    want_planes = oracle.unpack(eb, bits, a_h, signed, 4)
                    for p_off in (0, 4):
                        bufs = [torch.zeros(n + 8, dtype=torch.int32 if signed else torch.uint32, device="cuda") for _ in bits]
                        ops.unpack(L, a, signed=signed, out=[b_[p_off:p_off + n] for b_ in bufs])
                        for f in range(len(bits)):
                            np.testing.assert_array_equal(to_np(bufs[f][p_off:p_off + n], want_planes[f].dtype), want_planes[f],
                                                          err_msg=f"unpack {L} f{f} off{off} p_off{p_off} vec256={vec256}")
                            assert int(bufs[f][p_off + n:].view(torch.int32).abs().sum()) == 0
                pair = ["add_unsigned_saturate", "add_signed_saturate"]
                outs = [to_dev(np.zeros(n, dtype=a_h.dtype), off) for _ in pair]
                ops.binary_multi(pair, L, a, b, outs=outs)
                for op, o in zip(pair, outs):
                    np.testing.assert_array_equal(to_np(o), oracle.binary(op, eb, bits, a_h, b_h))
    finally:
        ops.set_option("vec256", 1)
