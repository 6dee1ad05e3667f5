"""Layout: the Python mirror of `packed_int<Integer, Bits...>` /
`make_packed_int<Bits...>` (/root/reference/include/pint/pint.hpp:758-795).

The reference rejects bad layouts at compile time (static_asserts at
pint.hpp:761-764, 727-728); here the same conditions raise ValueError at
construction.  Mask values come from the library's mask generator
(include/pint/cuda_layout.hpp) so Python, C++ and the kernels agree by
construction.
"""
import ctypes as C

from . import _lib

_INTEGERS = {"uint8_t": 1, "uint16_t": 2, "uint32_t": 4, "uint64_t": 8, "u8": 1, "u16": 2, "u32": 4, "u64": 8}


class Layout:
    def __init__(self, integer, *bits):
        """Layout("uint32_t", 8, 8, 8, 8) or Layout(4, 8, 8, 8, 8) == packed_int<uint32_t,8,8,8,8>."""
        elem_bytes = _INTEGERS.get(integer, integer)
        if len(bits) == 1 and isinstance(bits[0], (tuple, list)):
            bits = tuple(bits[0])
        self.elem_bytes = elem_bytes
        self.bits = tuple(int(b) for b in bits)
        if elem_bytes not in (1, 2, 4, 8):
            raise ValueError("Integer must be unsigned integer (uint8_t, uint16_t, uint32_t or uint64_t)")
        if not (1 <= len(self.bits) <= _lib.MAX_FIELDS):
            raise ValueError("layout needs 1..64 fields")
        if any(b < 1 for b in self.bits):
            raise ValueError("field of 0 bits")
        if sum(self.bits) > 8 * elem_bytes:
            raise ValueError("Integer won't fit given number of bits")
        self.c = _lib.PintLayout()
        self.c.elem_bytes = elem_bytes
        self.c.nfields = len(self.bits)
        for i, b in enumerate(self.bits):
            self.c.bits[i] = b

    @classmethod
    def make(cls, *bits):
        """make_packed_int<Bits...>: smallest of uint8/16/32/64_t that fits (pint.hpp:789-795)."""
        total = sum(bits)
        eb = storage_bytes(total)
        if eb == 0:
            raise ValueError("Requested amount of bits is not supported")
        return cls(eb, *bits)

    @property
    def nfields(self):
        return len(self.bits)

    @property
    def total_bits(self):
        return sum(self.bits)

    @property
    def ref(self):
        return C.byref(self.c)

    def masks(self):
        H, L, A, t = C.c_uint64(), C.c_uint64(), C.c_uint64(), C.c_uint32()
        _lib.check(_lib.lib().pint_cuda_layout_masks(self.ref, C.byref(H), C.byref(L), C.byref(A), C.byref(t)),
                   "pint_cuda_layout_masks")
        return {"H": H.value, "L": L.value, "A": A.value, "sat_type": t.value}

    def is_specialized(self):
        return bool(_lib.lib().pint_cuda_layout_is_specialized(self.ref))

    def sliced(self, start, end):
        """Result type of slice<start, end> (sliced_int, pint.hpp:746-754)."""
        if not (start < end <= self.nfields):
            raise ValueError("Incorrect slice bounds")
        return Layout(self.elem_bytes, *self.bits[start:end])

    def __eq__(self, other):
        return isinstance(other, Layout) and (self.elem_bytes, self.bits) == (other.elem_bytes, other.bits)

    def __hash__(self):
        return hash((self.elem_bytes, self.bits))

    def __repr__(self):
        return f"packed_int<uint{8 * self.elem_bytes}_t,{','.join(map(str, self.bits))}>"


def storage_bytes(total_bits):
    return _lib.lib().pint_cuda_storage_bytes(total_bits)
