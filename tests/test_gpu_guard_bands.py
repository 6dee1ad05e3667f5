"""GPU: out-of-bounds writes would land in sentinel bands placed directly before and
after every output array (compute-sanitizer is closed on this pool, so this is the
memcheck): every kernel family, ragged sizes, aligned and misaligned starts."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

GUARD = 4096  # bytes on each side
SENT = 0xA5


class Guarded:
    def __init__(self, nbytes, offset_bytes=0):
        self.buf = torch.full((GUARD + offset_bytes + nbytes + GUARD,), SENT, dtype=torch.uint8, device="cuda")
        self.lo, self.hi = GUARD + offset_bytes, GUARD + offset_bytes + nbytes

    def view(self, dtype):
        return self.buf[self.lo:self.hi].view(dtype)

    def intact(self):
        return bool((self.buf[:self.lo] == SENT).all()) and bool((self.buf[self.hi:] == SENT).all())


@pytest.mark.parametrize("eb,bits", [(1, (1, 1, 1)), (2, (5, 6, 5)), (4, (8, 8, 8, 8)), (4, (7, 9, 3)),
                                     (8, (16, 16, 16, 16)), (8, (3, 7, 11, 13, 17, 13))])
def test_no_writes_outside_outputs(eb, bits):
    import pint_b200
    from pint_b200 import ops
    L = pint_b200.Layout(eb, *bits)
    udt = {1: torch.uint8, 2: torch.uint16, 4: torch.uint32, 8: torch.uint64}[eb]
    for n in (1, 3, 4, 17, 255, 256, 1023, 4097, 65537):
        for off_elems in (0, 1, 3):
            a = torch.empty(n + 8, dtype=udt, device="cuda")[off_elems:off_elems + n]
            b = torch.empty(n + 8, dtype=udt, device="cuda")[off_elems:off_elems + n]
            ops.fill_splitmix(a, 1)
            ops.fill_splitmix(b, 2)
            g = Guarded(n * eb, off_elems * eb)
            out = g.view(udt)
            for op in ("add_signed_saturate", "sub_unsigned_saturate", "bit_xor", "max_signed"):
                ops.binary(op, L, a, b, out=out)
                ops.fused_unpack_arith_repack(op, L, a, b, out=out)
            ops.shift_left(L, a, 1, out=out)
            ops.slice(L, 0, 1, a, out=out)
            # the extra caller shapes: constant operand (either side), interleaved operands
            pairs = torch.empty(2 * n + 16, dtype=udt, device="cuda")[2 * off_elems:2 * off_elems + 2 * n]
            ops.fill_splitmix(pairs, 4)
            for force in (0, 1):
                ops.set_option("force_generic", force)
                ops.binary_scalar("add_signed_saturate", L, a, 0x5A, out=out)
                ops.binary_scalar("sub_wrap", L, a, 0x33, out=out, scalar_first=True)
                ops.binary_pairs("add_unsigned_saturate", L, pairs, out=out)
            ops.set_option("force_generic", 0)
            fg = Guarded(n, off_elems)
            ops.equal(L, a, b, out=fg.view(torch.uint8))
            torch.cuda.synchronize()
            assert fg.intact(), f"write outside the equality flags: {L} n={n} off={off_elems}"
            ops.fill_splitmix(out, 9)
            pb = 8 if max(bits) > 32 else 4
            pg = [Guarded(n * pb, (off_elems % 2) * pb) for _ in bits]
            planes = [p.view(torch.int32 if pb == 4 else torch.int64) for p in pg]
            ops.unpack(L, a, signed=True, plane_bytes=pb, out=planes)
            ops.repack(L, planes, out=out)
            torch.cuda.synchronize()
            assert g.intact(), f"write outside `out`: {L} n={n} off={off_elems}"
            assert all(p.intact() for p in pg), f"write outside a plane: {L} n={n} off={off_elems}"


@pytest.mark.parametrize("eb,bits", [(4, (8, 8, 8)), (2, (12,)), (4, (7, 9, 3)), (8, (21, 21, 21)), (1, (2, 3, 2)),
                                     (8, (12, 12, 12)), (8, (16, 16, 16)), (4, (11, 11, 9))])
def test_dense_no_writes_outside_outputs(eb, bits):
    """Dense streams: binary ops, shifts, and the conversions in both kernel families (S compiled in:
    direct / staged / word-owner insert, group-owner extract; run-time geometry) -- whole tiles plus a
    remainder, outputs at 16-byte offsets, padded outputs and planes also at element offsets."""
    import pint_b200
    from pint_b200 import dense, ops
    L = pint_b200.Layout(eb, *bits)
    udt = {1: torch.uint8, 2: torch.uint16, 4: torch.uint32, 8: torch.uint64}[eb]
    try:
        for n in (1, 127, 4097, 70001):
            nb = dense.nbytes(L, n)
            x = torch.empty(n, dtype=udt, device="cuda")
            ops.fill_splitmix(x, 3)
            for windows in (0, 1):
                ops.set_option("dense_windows", windows)
                for off16 in (0, 16):
                    g = Guarded(nb, off16)
                    d = g.view(torch.uint8)
                    dense.pack(L, x, out=d)
                    g2 = Guarded(nb, off16)
                    d2 = g2.view(torch.uint8)
                    dense.binary("add_signed_saturate", L, d, d, n, out=d2)
                    dense.shift(1, L, d, 1, n, out=d2)
                    planes_g = [Guarded(n * 4, off16) for _ in bits]
                    planes = [p.view(torch.int32) for p in planes_g]
                    dense.to_planes(L, d, n, signed=True, out=planes)
                    dense.from_planes(L, planes, out=d2)
                    for off_elems in (0, 1):
                        gp = Guarded(n * eb, off_elems * eb)
                        dense.unpack(L, d, n, out=gp.view(udt))
                        torch.cuda.synchronize()
                        assert gp.intact(), f"write outside the padded array: {L} n={n} windows={windows} off={off_elems}"
                    torch.cuda.synchronize()
                    assert g.intact() and g2.intact(), f"write outside a dense stream: {L} n={n} windows={windows} off16={off16}"
                    assert all(p.intact() for p in planes_g), f"write outside a plane: {L} n={n} windows={windows}"
    finally:
        ops.set_option("dense_windows", 0)
