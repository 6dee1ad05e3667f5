"""GPU: randomly drawn layouts (none of them in the registry, so the runtime-mask
kernels run) x all ops x SWAR and fused kernels, against the oracle.  Covers
saturation-mask types 0/1/2, 1-bit fields, unused high bits, up to 64 fields."""
import numpy as np
import pytest

from oracle import OPS

pytestmark = pytest.mark.gpu

from gpu_util import to_dev, to_np  # noqa: E402


def random_layouts(count, seed):
    rng = np.random.default_rng(seed)
    out = []
    while len(out) < count:
        eb = int(rng.choice([1, 2, 4, 8]))
        room = 8 * eb
        style = rng.integers(0, 4)
        bits = []
        if style == 0:      # many tiny fields
            while room > 0 and len(bits) < 64 and rng.random() < 0.95:
                w = int(rng.integers(1, 4)); w = min(w, room); bits.append(w); room -= w
        elif style == 1:    # equal widths, maybe with padding
            w = int(rng.integers(1, max(2, room // 2)))
            k = int(rng.integers(1, room // w + 1))
            bits = [w] * k
        else:               # irregular
            while room > 0 and rng.random() < 0.85:
                w = int(rng.integers(1, min(room, 40) + 1)); bits.append(w); room -= w
        if bits:
            out.append((eb, tuple(bits)))
    return out


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_random_layouts_vs_oracle(oracle, seed):
    import pint_b200
    from pint_b200 import ops
    types = set()
    for eb, bits in random_layouts(40, seed):
        L = pint_b200.Layout(eb, *bits)
        types.add(L.masks()["sat_type"])
        n = 20000 + 7 * len(bits)
        a_h, b_h = oracle.fill(eb, n, seed * 1000 + len(bits)), oracle.fill(eb, n, seed * 2000 + sum(bits))
        # mix in structured edges: all ones, H, L, ~H
        m = oracle.masks(eb, bits)[1]
        edge = np.array([0, m.word, m.H, m.L, m.word & ~m.H, m.A], dtype=np.uint64).astype(a_h.dtype)
        a_h[:36] = np.repeat(edge, 6); b_h[:36] = np.tile(edge, 6)
        a, b = to_dev(a_h), to_dev(b_h)
        for op in OPS:
            want = oracle.binary(op, eb, bits, a_h, b_h)
            np.testing.assert_array_equal(to_np(ops.binary(op, L, a, b)), want, err_msg=f"swar {L} {op}")
            np.testing.assert_array_equal(to_np(ops.fused_unpack_arith_repack(op, L, a, b)), want, err_msg=f"fused {L} {op}")
        pb = 8 if max(bits) > 32 else 4
        planes = ops.unpack(L, a, signed=True, plane_bytes=pb)
        want_p = oracle.unpack(eb, bits, a_h, True, pb)
        for f in range(len(bits)):
            np.testing.assert_array_equal(to_np(planes[f], want_p[f].dtype), want_p[f], err_msg=f"unpack {L} f{f}")
        np.testing.assert_array_equal(to_np(ops.repack(L, planes)), oracle.repack(eb, bits, want_p, pb))
        for amt in {0, 1, max(bits) // 2, max(bits) - 1, max(bits)}:
            if amt == 0 and len(bits) == 1 and bits[0] == 8 * eb and eb >= 4:
                continue  # reference UB, see oracle/pint_oracle.c
            np.testing.assert_array_equal(to_np(ops.shift_left(L, a, amt)), oracle.shift(1, eb, bits, amt, a_h), err_msg=f"shl {L} {amt}")
            np.testing.assert_array_equal(to_np(ops.shift_right_unsigned(L, a, amt)), oracle.shift(0, eb, bits, amt, a_h), err_msg=f"shr {L} {amt}")
        s, e = sorted(np.random.default_rng(seed + len(bits)).choice(len(bits) + 1, 2, replace=False)) if len(bits) > 1 else (0, 1)
        out, _ = ops.slice(L, int(s), int(e), a)
        np.testing.assert_array_equal(to_np(out), oracle.slice(eb, bits, int(s), int(e), a_h))
    assert types == {0, 1, 2}
