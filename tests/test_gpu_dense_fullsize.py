"""GPU, full size: dense streams of 2^30 RGB24 elements (3 GiB per stream) and 2^29 19-bit elements,
checked through size-independent properties and against the padded kernels (themselves pinned to the
reference); no CPU oracle at this size."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("eb,bits,log2n", [(4, (8, 8, 8), 30), (4, (7, 9, 3), 29)])
def test_dense_fullsize_properties(eb, bits, log2n):
    import pint_b200
    from pint_b200 import dense, ops
    L = pint_b200.Layout(eb, *bits)
    S = sum(bits)
    n = (1 << log2n) - 37  # not a multiple of the chunk: the tail tile is exercised at full size too
    x = torch.empty(n, dtype=torch.uint32, device="cuda")
    ops.fill_splitmix(x, 11)
    da = dense.pack(L, x)
    ops.fill_splitmix(x, 12)
    db = dense.pack(L, x)
    assert da.numel() == ((n * S + 127) // 128) * 16
    # wrap ops are exact inverses; xor twice is the identity
    d = dense.binary("sub_wrap", L, da, db, n)
    assert torch.equal(dense.binary("add_wrap", L, d, db, n, out=d), da)
    d = dense.binary("bit_xor", L, da, db, n, out=d)
    assert torch.equal(dense.binary("bit_xor", L, d, db, n, out=d), da)
    # min + max == a + b (wrapping) per field;  max(a, a) == a
    mn, mx = dense.binary_multi(["min_unsigned", "max_unsigned"], L, da, db, n)
    assert torch.equal(dense.binary("add_wrap", L, mn, mx, n, out=mn), dense.binary("add_wrap", L, da, db, n, out=mx))
    assert torch.equal(dense.binary("max_signed", L, da, da, n, out=mn), da)
    del mn, mx, d
    # dense op == pack(padded op(unpack)), with the padded kernel pinned to the reference elsewhere
    a_p, b_p = dense.unpack(L, da, n), dense.unpack(L, db, n)
    for op in ("add_signed_saturate", "sub_unsigned_saturate"):
        got = dense.binary(op, L, da, db, n)
        want = dense.pack(L, ops.binary(op, L, a_p, b_p))
        assert torch.equal(got, want), op
        del got, want
    # planes round trip
    planes = dense.to_planes(L, da, n, signed=True)
    assert torch.equal(dense.from_planes(L, planes), da)
