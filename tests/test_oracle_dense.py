"""CPU: the dense bit-stream restatement (oracle/pint_oracle.c, dense section) against an
independent big-integer model of the format spec, hand-computed byte vectors, and the
reference-pinned per-element ops (dense op == pack(op(unpack)))."""
import numpy as np
import pytest

from oracle import OPS

LAYOUTS = [(2, (12,)), (4, (8, 8, 8)), (4, (5, 6, 5, 3)), (8, (10, 10, 10, 10)), (8, (21, 21, 21)), (1, (3,)),
           (4, (10, 10, 10)), (8, (64,)), (8, (3, 7, 11, 13, 17, 12)), (4, (1,)), (2, (5, 6, 5)), (4, (7, 9, 3)),
           (8, (16, 16, 16)), (8, (31, 2)), (8, (1, 62))]


def model_pack(vals, S):
    """The spec, stated with Python integers: the stream is sum(v_i << (i*S)), little-endian."""
    big = 0
    for i, v in enumerate(vals):
        big |= (int(v) & ((1 << S) - 1)) << (i * S)
    nb = ((len(vals) * S + 127) // 128) * 16
    return np.frombuffer(big.to_bytes(nb, "little"), dtype=np.uint8)


def test_hand_vectors(oracle):
    # two 12-bit samples 0xABC, 0x123 -> bytes BC 3A 12 (the classic packed-12 layout)
    d = oracle.dense_pack(2, (12,), np.array([0xABC, 0x123], dtype=np.uint16))
    assert d[:3].tolist() == [0xBC, 0x3A, 0x12] and not d[3:].any() and d.size == 16
    # RGB24: <8,8,8> elements are 3 bytes each, field 0 first
    d = oracle.dense_pack(4, (8, 8, 8), np.array([0xFF332211, 0x00665544], dtype=np.uint32))
    assert d[:6].tolist() == [0x11, 0x22, 0x33, 0x44, 0x55, 0x66] and not d[6:].any()
    # 19-bit elements straddle bytes and words
    v = np.array([0x7FFFF, 0, 0x7FFFF], dtype=np.uint32)
    d = oracle.dense_pack(4, (7, 9, 3), v)
    assert int.from_bytes(d.tobytes(), "little") == 0x7FFFF | (0x7FFFF << 38)
    assert oracle.dense_bytes((7, 9, 3), 7) == 32 and oracle.dense_bytes((8, 8, 8), 0) == 0


@pytest.mark.parametrize("eb,bits", LAYOUTS)
def test_pack_unpack_vs_model(oracle, eb, bits):
    S = sum(bits)
    for n in (0, 1, 2, 7, 128, 129, 1000, 4097):
        x = oracle.fill(eb, n, 5)
        d = oracle.dense_pack(eb, bits, x)
        assert d.size == ((n * S + 127) // 128) * 16
        np.testing.assert_array_equal(d, model_pack(x, S))
        np.testing.assert_array_equal(oracle.dense_unpack(eb, bits, d, n), x & np.array((1 << S) - 1, dtype=x.dtype))


@pytest.mark.parametrize("eb,bits", LAYOUTS)
def test_dense_binary_is_pack_op_unpack(oracle, ref, eb, bits):
    S = sum(bits)
    n = 1531
    mask = np.array((1 << S) - 1, dtype={1: np.uint8, 2: np.uint16, 4: np.uint32, 8: np.uint64}[eb])
    x, y = oracle.fill(eb, n, 11) & mask, oracle.fill(eb, n, 12) & mask
    dx, dy = oracle.dense_pack(eb, bits, x), oracle.dense_pack(eb, bits, y)
    # garbage above n*S in the inputs must not reach the output
    dx[-1] |= 0x80 if (n * S) % 128 else 0
    for op in OPS:
        want = model_pack(oracle.binary(op, eb, bits, x, y), S)
        np.testing.assert_array_equal(oracle.dense_binary(op, eb, bits, dx, dy, n), want, err_msg=f"{bits} {op}")
        if ref is not None and ref.has(eb, bits):
            np.testing.assert_array_equal(model_pack(ref.binary(op, eb, bits, x, y), S), want)
