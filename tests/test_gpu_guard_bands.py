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
            ops.fill_splitmix(out, 9)
            pb = 8 if max(bits) > 32 else 4
            pg = [Guarded(n * pb, (off_elems % 2) * pb) for _ in bits]
            planes = [p.view(torch.int32 if pb == 4 else torch.int64) for p in pg]
            ops.unpack(L, a, signed=True, plane_bytes=pb, out=planes)
            ops.repack(L, planes, out=out)
            torch.cuda.synchronize()
            assert g.intact(), f"write outside `out`: {L} n={n} off={off_elems}"
            assert all(p.intact() for p in pg), f"write outside a plane: {L} n={n} off={off_elems}"
