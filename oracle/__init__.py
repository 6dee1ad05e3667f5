"""ctypes loaders for the checker libraries.  TEST INFRASTRUCTURE ONLY.

`Oracle`  -> oracle/liboracle.so        (C restatement, pint_oracle.c; "port")
`Ref`     -> oracle/_ref/libpint_ref.so (unmodified reference header; "reference")

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
reference legs import this module.  pint_b200 never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(HERE, "liboracle.so")
REF_SO = os.path.join(HERE, "_ref", "libpint_ref.so")
REF_TEST_BIN = os.path.join(HERE, "_ref", "pint_test")

OPS = [
    "add_wrap", "add_unsigned_saturate", "add_signed_saturate",
    "sub_wrap", "sub_unsigned_saturate", "sub_signed_saturate",
    "bit_or", "bit_and", "bit_xor",
    "min_unsigned", "max_unsigned", "min_signed", "max_signed",
]
OP_CODE = {name: i for i, name in enumerate(OPS)}

_DT = {1: np.uint8, 2: np.uint16, 4: np.uint32, 8: np.uint64}


class CLayout(C.Structure):
    _fields_ = [("elem_bytes", C.c_uint32), ("nfields", C.c_uint32), ("bits", C.c_uint8 * 64)]


class CMasks(C.Structure):
    _fields_ = [("word", C.c_uint64), ("H", C.c_uint64), ("L", C.c_uint64), ("A", C.c_uint64),
                ("total", C.c_uint32), ("maxbits", C.c_uint32), ("sat_type", C.c_uint32),
                ("nuniq", C.c_uint32), ("off", C.c_uint8 * 64), ("uniq_w", C.c_uint8 * 64),
                ("uniq_lo", C.c_uint64 * 64)]


def clayout(elem_bytes, bits):
    l = CLayout()
    l.elem_bytes = elem_bytes
    l.nfields = len(bits)
    for i, b in enumerate(bits[:64]):
        l.bits[i] = b
    return l


def build(ref=True):
    """Build liboracle.so and, when /root/reference is present, oracle/_ref."""
    subprocess.check_call(["make", "-s", "-C", HERE, "liboracle.so"])
    if ref and os.path.exists("/root/reference/include/pint/pint.hpp"):
        subprocess.check_call(["make", "-s", "-C", HERE, "ref"])


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def _plane_ptrs(planes):
    arr = (C.c_void_p * len(planes))()
    for i, p in enumerate(planes):
        arr[i] = None if p is None else p.ctypes.data
    return arr


class Oracle:
    """The C restatement (kind "port")."""

    def __init__(self):
        if not os.path.exists(ORACLE_SO):
            build(ref=False)
        self.lib = C.CDLL(ORACLE_SO)
        L = self.lib
        L.pint_oracle_binary1.restype = C.c_uint64
        L.pint_oracle_binary1.argtypes = [C.c_int, C.c_void_p, C.c_uint64, C.c_uint64]
        L.pint_oracle_naive_binary1.restype = C.c_uint64
        L.pint_oracle_naive_binary1.argtypes = [C.c_int, C.c_void_p, C.c_uint64, C.c_uint64]
        L.pint_oracle_shift_left1.restype = C.c_uint64
        L.pint_oracle_shift_left1.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64]
        L.pint_oracle_shift_right_unsigned1.restype = C.c_uint64
        L.pint_oracle_shift_right_unsigned1.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64]
        L.pint_oracle_get1.restype = C.c_uint64
        L.pint_oracle_get1.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint64]
        L.pint_oracle_get_signed1.restype = C.c_int64
        L.pint_oracle_get_signed1.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint64]
        L.pint_oracle_slice1.restype = C.c_uint64
        L.pint_oracle_slice1.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint64]
        L.pint_oracle_make_truncate1.restype = C.c_uint64
        L.pint_oracle_make_truncate1.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.pint_oracle_splitmix64.restype = C.c_uint64
        L.pint_oracle_splitmix64.argtypes = [C.c_uint64]
        L.pint_oracle_storage_bytes.restype = C.c_uint32
        L.pint_oracle_storage_bytes.argtypes = [C.c_uint32]
        L.pint_oracle_binary.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]
        L.pint_oracle_shift.argtypes = [C.c_int, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_size_t]
        L.pint_oracle_slice.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p, C.c_size_t]
        L.pint_oracle_unpack.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_size_t]
        L.pint_oracle_repack.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_size_t]
        L.pint_oracle_fill.argtypes = [C.c_void_p, C.c_uint32, C.c_size_t, C.c_uint64, C.c_uint64]
        L.pint_oracle_checksum.argtypes = [C.c_void_p, C.c_uint32, C.c_size_t, C.c_void_p, C.c_void_p]
        L.pint_oracle_pairs_reduce.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
        L.pint_oracle_stream_checksum.argtypes = [C.c_int, C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint64, C.c_size_t,
                                                  C.c_void_p, C.c_void_p]
        L.pint_oracle_dense_bytes.restype = C.c_size_t
        L.pint_oracle_dense_bytes.argtypes = [C.c_uint32, C.c_size_t]
        L.pint_oracle_dense_pack.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]
        L.pint_oracle_dense_unpack.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]
        L.pint_oracle_dense_binary.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]

    # -- layout helpers
    def validate(self, elem_bytes, bits):
        l = clayout(elem_bytes, bits)
        if len(bits) > 64:
            return -3
        return self.lib.pint_oracle_validate(C.byref(l))

    def masks(self, elem_bytes, bits):
        l = clayout(elem_bytes, bits)
        m = CMasks()
        rc = self.lib.pint_oracle_make_masks(C.byref(l), C.byref(m))
        if rc:
            raise ValueError(f"invalid layout ({rc})")
        return l, m

    def storage_bytes(self, total):
        return self.lib.pint_oracle_storage_bytes(total)

    # -- single element
    def binary1(self, op, elem_bytes, bits, a, b):
        l, m = self.masks(elem_bytes, bits)
        return self.lib.pint_oracle_binary1(OP_CODE[op], C.byref(m), a, b)

    def naive_binary1(self, op, elem_bytes, bits, a, b):
        l = clayout(elem_bytes, bits)
        return self.lib.pint_oracle_naive_binary1(OP_CODE[op], C.byref(l), a, b)

    def shift1(self, left, elem_bytes, bits, v, amount):
        l, m = self.masks(elem_bytes, bits)
        f = self.lib.pint_oracle_shift_left1 if left else self.lib.pint_oracle_shift_right_unsigned1
        return f(C.byref(m), v, amount)

    def get1(self, elem_bytes, bits, index, v, signed=False):
        l, m = self.masks(elem_bytes, bits)
        f = self.lib.pint_oracle_get_signed1 if signed else self.lib.pint_oracle_get1
        return f(C.byref(l), C.byref(m), index, v)

    def slice1(self, elem_bytes, bits, s, e, v):
        l, m = self.masks(elem_bytes, bits)
        return self.lib.pint_oracle_slice1(C.byref(l), C.byref(m), s, e, v)

    def make_truncate1(self, elem_bytes, bits, values):
        l, m = self.masks(elem_bytes, bits)
        arr = (C.c_uint64 * len(values))(*[v & 0xFFFFFFFFFFFFFFFF for v in values])
        return self.lib.pint_oracle_make_truncate1(C.byref(l), C.byref(m), arr)

    # -- arrays (numpy)
    def binary(self, op, elem_bytes, bits, a, b):
        l = clayout(elem_bytes, bits)
        a = np.ascontiguousarray(a, dtype=_DT[elem_bytes])
        b = np.ascontiguousarray(b, dtype=_DT[elem_bytes])
        out = np.empty_like(a)
        rc = self.lib.pint_oracle_binary(OP_CODE[op] if isinstance(op, str) else op, C.byref(l),
                                         _ptr(a), _ptr(b), _ptr(out), a.size)
        if rc:
            raise ValueError(f"pint_oracle_binary rc={rc}")
        return out

    def shift(self, left, elem_bytes, bits, amount, x):
        l = clayout(elem_bytes, bits)
        x = np.ascontiguousarray(x, dtype=_DT[elem_bytes])
        out = np.empty_like(x)
        rc = self.lib.pint_oracle_shift(int(left), C.byref(l), amount, _ptr(x), _ptr(out), x.size)
        if rc:
            raise ValueError(f"pint_oracle_shift rc={rc}")
        return out

    def slice(self, elem_bytes, bits, s, e, x):
        l = clayout(elem_bytes, bits)
        x = np.ascontiguousarray(x, dtype=_DT[elem_bytes])
        out = np.empty_like(x)
        rc = self.lib.pint_oracle_slice(C.byref(l), s, e, _ptr(x), _ptr(out), x.size)
        if rc:
            raise ValueError(f"pint_oracle_slice rc={rc}")
        return out

    def unpack(self, elem_bytes, bits, x, signed, plane_bytes=4, fields=None):
        l = clayout(elem_bytes, bits)
        x = np.ascontiguousarray(x, dtype=_DT[elem_bytes])
        pdt = {(1, True): np.int8, (2, True): np.int16, (4, True): np.int32, (8, True): np.int64,
               (1, False): np.uint8, (2, False): np.uint16, (4, False): np.uint32, (8, False): np.uint64}[(plane_bytes, bool(signed))]
        planes = [np.empty(x.size, dtype=pdt) if (fields is None or f in fields) else None
                  for f in range(len(bits))]
        rc = self.lib.pint_oracle_unpack(C.byref(l), int(signed), _ptr(x), _plane_ptrs(planes), plane_bytes, x.size)
        if rc:
            raise ValueError(f"pint_oracle_unpack rc={rc}")
        return planes

    def repack(self, elem_bytes, bits, planes, plane_bytes=4):
        l = clayout(elem_bytes, bits)
        pdt = {1: np.int8, 2: np.int16, 4: np.int32, 8: np.uint64}[plane_bytes]
        planes = [np.ascontiguousarray(p).view(pdt) for p in planes]
        out = np.empty(planes[0].size, dtype=_DT[elem_bytes])
        rc = self.lib.pint_oracle_repack(C.byref(l), _plane_ptrs(planes), plane_bytes, _ptr(out), out.size)
        if rc:
            raise ValueError(f"pint_oracle_repack rc={rc}")
        return out

    def fill(self, elem_bytes, n, seed, start=0):
        out = np.empty(n, dtype=_DT[elem_bytes])
        self.lib.pint_oracle_fill(_ptr(out), elem_bytes, n, seed, start)
        return out

    def checksum(self, x):
        x = np.ascontiguousarray(x)
        s, r = C.c_uint64(), C.c_uint64()
        self.lib.pint_oracle_checksum(_ptr(x), x.dtype.itemsize, x.size, C.byref(s), C.byref(r))
        return s.value, r.value

    def stream_checksum(self, op, elem_bytes, bits, seed_a, seed_b, start, n, nthreads=1):
        """(sum, xor) of OP(a[i], b[i]) for i in [start, start + n) over the splitmix inputs, without the arrays."""
        l = clayout(elem_bytes, bits)
        code = OP_CODE[op] if isinstance(op, str) else op

        def part(lo, cnt):
            s, r = C.c_uint64(), C.c_uint64()
            rc = self.lib.pint_oracle_stream_checksum(code, C.byref(l), seed_a, seed_b, start + lo, cnt, C.byref(s), C.byref(r))
            if rc:
                raise ValueError(f"pint_oracle_stream_checksum rc={rc}")
            return s.value, r.value

        if nthreads <= 1 or n < (1 << 16):
            return part(0, n)
        from concurrent.futures import ThreadPoolExecutor  # ctypes releases the GIL during the call
        step = (n + nthreads - 1) // nthreads
        with ThreadPoolExecutor(nthreads) as ex:
            parts = list(ex.map(lambda lo: part(lo, min(step, n - lo)), range(0, n, step)))
        s = x = 0
        for ps, px in parts:
            s = (s + ps) & 0xFFFFFFFFFFFFFFFF
            x ^= px
        return s, x

    def pairs_reduce(self, op, elem_bytes, bits, pairs):
        l = clayout(elem_bytes, bits)
        pairs = np.ascontiguousarray(pairs, dtype=_DT[elem_bytes])
        s = C.c_uint64()
        rc = self.lib.pint_oracle_pairs_reduce(OP_CODE[op], C.byref(l), _ptr(pairs), pairs.size // 2, C.byref(s))
        if rc:
            raise ValueError(f"pint_oracle_pairs_reduce rc={rc}")
        return s.value

    # -- dense bit-stream storage (our format; element i at bits [i*S, (i+1)*S))
    def dense_bytes(self, bits, n):
        return self.lib.pint_oracle_dense_bytes(sum(bits), n)

    def dense_pack(self, elem_bytes, bits, x):
        l = clayout(elem_bytes, bits)
        x = np.ascontiguousarray(x, dtype=_DT[elem_bytes])
        out = np.empty(self.dense_bytes(bits, x.size), dtype=np.uint8)
        rc = self.lib.pint_oracle_dense_pack(C.byref(l), _ptr(x), _ptr(out), x.size)
        if rc:
            raise ValueError(f"pint_oracle_dense_pack rc={rc}")
        return out

    def dense_unpack(self, elem_bytes, bits, dense, n):
        l = clayout(elem_bytes, bits)
        dense = np.ascontiguousarray(dense, dtype=np.uint8)
        assert dense.size >= self.dense_bytes(bits, n)
        out = np.empty(n, dtype=_DT[elem_bytes])
        rc = self.lib.pint_oracle_dense_unpack(C.byref(l), _ptr(dense), _ptr(out), n)
        if rc:
            raise ValueError(f"pint_oracle_dense_unpack rc={rc}")
        return out

    def dense_binary(self, op, elem_bytes, bits, a, b, n):
        l = clayout(elem_bytes, bits)
        a = np.ascontiguousarray(a, dtype=np.uint8)
        b = np.ascontiguousarray(b, dtype=np.uint8)
        out = np.empty(self.dense_bytes(bits, n), dtype=np.uint8)
        rc = self.lib.pint_oracle_dense_binary(OP_CODE[op] if isinstance(op, str) else op, C.byref(l),
                                               _ptr(a), _ptr(b), _ptr(out), n)
        if rc:
            raise ValueError(f"pint_oracle_dense_binary rc={rc}")
        return out


class Ref:
    """The unmodified reference header compiled here (kind "reference")."""

    @staticmethod
    def available():
        return os.path.exists(REF_SO)

    def __init__(self):
        if not os.path.exists(REF_SO):
            raise FileNotFoundError(REF_SO + " (build with `make -C oracle ref` where /root/reference exists)")
        self.lib = C.CDLL(REF_SO)
        L = self.lib
        u32, vp, sz = C.c_uint32, C.c_void_p, C.c_size_t
        L.pint_ref_layout_at.argtypes = [C.c_int, vp, vp, vp]
        L.pint_ref_has_layout.argtypes = [u32, u32, vp]
        L.pint_ref_constants.argtypes = [u32, u32, vp, vp]
        L.pint_ref_binary.argtypes = [C.c_int, u32, u32, vp, vp, vp, vp, sz, C.c_int]
        L.pint_ref_pairs_reduce.argtypes = [C.c_int, u32, u32, vp, vp, sz, vp]
        L.pint_ref_shift.argtypes = [C.c_int, u32, u32, vp, C.c_uint64, vp, vp, sz, C.c_int]
        L.pint_ref_stream_checksum.argtypes = [C.c_int, u32, u32, vp, C.c_uint64, C.c_uint64, C.c_uint64, sz, C.c_int, vp]
        L.pint_ref_unpack.argtypes = [u32, u32, vp, C.c_int, vp, vp, C.c_int, sz]
        L.pint_ref_repack.argtypes = [u32, u32, vp, vp, C.c_int, vp, sz]
        L.pint_ref_slice.argtypes = [u32, u32, vp, u32, u32, vp, vp, sz]
        L.pint_ref_storage_bytes.restype = u32
        L.pint_ref_storage_bytes.argtypes = [u32]

    @staticmethod
    def _bits(bits):
        return (C.c_uint8 * 64)(*bits)

    def layouts(self):
        out = []
        for i in range(self.lib.pint_ref_num_layouts()):
            eb, nf, bits = C.c_uint32(), C.c_uint32(), (C.c_uint8 * 64)()
            self.lib.pint_ref_layout_at(i, C.byref(eb), C.byref(nf), bits)
            out.append((eb.value, tuple(bits[:nf.value])))
        return out

    def has(self, elem_bytes, bits):
        return bool(self.lib.pint_ref_has_layout(elem_bytes, len(bits), self._bits(bits)))

    def constants(self, elem_bytes, bits):
        out = (C.c_uint64 * 4)()
        rc = self.lib.pint_ref_constants(elem_bytes, len(bits), self._bits(bits), out)
        if rc:
            raise KeyError((elem_bytes, bits))
        return {"H": out[0], "L": out[1], "A": out[2], "sat_type": out[3]}

    def storage_bytes(self, total):
        return self.lib.pint_ref_storage_bytes(total)

    def hardware_threads(self):
        return self.lib.pint_ref_hardware_threads()

    def binary(self, op, elem_bytes, bits, a, b, nthreads=1, out=None):
        a = np.ascontiguousarray(a, dtype=_DT[elem_bytes])
        b = np.ascontiguousarray(b, dtype=_DT[elem_bytes])
        if out is None:
            out = np.empty_like(a)
        rc = self.lib.pint_ref_binary(OP_CODE[op] if isinstance(op, str) else op, elem_bytes, len(bits),
                                      self._bits(bits), _ptr(a), _ptr(b), _ptr(out), a.size, nthreads)
        if rc:
            raise KeyError(f"pint_ref_binary rc={rc} for {(elem_bytes, bits)}")
        return out

    def stream_checksum(self, op, elem_bytes, bits, seed_a, seed_b, start, n, nthreads=1):
        """(sum, xor) of pint::OP(P(a[i]), P(b[i])).value() for i in [start, start + n) over the splitmix inputs."""
        out = (C.c_uint64 * 2)()
        rc = self.lib.pint_ref_stream_checksum(OP_CODE[op] if isinstance(op, str) else op, elem_bytes, len(bits),
                                               self._bits(bits), seed_a, seed_b, start, n, nthreads, out)
        if rc:
            raise KeyError(f"pint_ref_stream_checksum rc={rc} for {(elem_bytes, bits)}")
        return out[0], out[1]

    def pairs_reduce(self, op, elem_bytes, bits, pairs):
        pairs = np.ascontiguousarray(pairs, dtype=_DT[elem_bytes])
        s = C.c_uint64()
        rc = self.lib.pint_ref_pairs_reduce(OP_CODE[op], elem_bytes, len(bits), self._bits(bits),
                                            _ptr(pairs), pairs.size // 2, C.byref(s))
        if rc:
            raise KeyError(f"pint_ref_pairs_reduce rc={rc}")
        return s.value

    def shift(self, left, elem_bytes, bits, amount, x, nthreads=1):
        x = np.ascontiguousarray(x, dtype=_DT[elem_bytes])
        out = np.empty_like(x)
        rc = self.lib.pint_ref_shift(int(left), elem_bytes, len(bits), self._bits(bits), amount,
                                     _ptr(x), _ptr(out), x.size, nthreads)
        if rc:
            raise KeyError(f"pint_ref_shift rc={rc}")
        return out

    def slice(self, elem_bytes, bits, s, e, x):
        x = np.ascontiguousarray(x, dtype=_DT[elem_bytes])
        out = np.empty_like(x)
        rc = self.lib.pint_ref_slice(elem_bytes, len(bits), self._bits(bits), s, e, _ptr(x), _ptr(out), x.size)
        if rc:
            raise KeyError(f"pint_ref_slice rc={rc}")
        return out

    def unpack(self, elem_bytes, bits, x, signed, plane_bytes=4):
        x = np.ascontiguousarray(x, dtype=_DT[elem_bytes])
        if signed:
            pdt = np.int32 if plane_bytes == 4 else np.int64
        else:
            pdt = np.uint32 if plane_bytes == 4 else np.uint64
        planes = [np.empty(x.size, dtype=pdt) for _ in bits]
        rc = self.lib.pint_ref_unpack(elem_bytes, len(bits), self._bits(bits), int(signed), _ptr(x),
                                      _plane_ptrs(planes), plane_bytes, x.size)
        if rc:
            raise KeyError(f"pint_ref_unpack rc={rc}")
        return planes

    def repack(self, elem_bytes, bits, planes, plane_bytes=4):
        pdt = np.int32 if plane_bytes == 4 else np.uint64
        planes = [np.ascontiguousarray(p).view(pdt) for p in planes]
        out = np.empty(planes[0].size, dtype=_DT[elem_bytes])
        rc = self.lib.pint_ref_repack(elem_bytes, len(bits), self._bits(bits), _plane_ptrs(planes),
                                      plane_bytes, _ptr(out), out.size)
        if rc:
            raise KeyError(f"pint_ref_repack rc={rc}")
        return out
