"""pint_b200.dense -- the dense bit-stream storage format (include/pint_cuda.h, "dense" section).

A dense stream is a torch.uint8 CUDA tensor of `nbytes(layout, n)` bytes in which element i of
`layout` occupies bits [i*S, (i+1)*S), S = sum(Bits).  Per-element semantics are the reference's
(pint.hpp): `binary(op, L, a, b, n)` equals `pack(op(unpack(a), unpack(b)))`.  Every function is one
call into libpint_cuda.so on the current torch stream; there is no CPU path.
"""
import torch

from . import _lib
from .ops import OP_CODE, _PLANE_DT, _check, _plane_array, _stream, dtype_of


def nbytes(layout, n):
    return _lib.lib().pint_cuda_dense_bytes(layout.ref, n)


def empty(layout, n, device="cuda"):
    return torch.empty(nbytes(layout, n), dtype=torch.uint8, device=device)


def _check_dense(layout, n, *streams):
    need = nbytes(layout, n)
    for t in streams:
        if not isinstance(t, torch.Tensor) or not t.is_cuda:
            raise TypeError("pint_b200 operates on CUDA tensors only (there is no CPU fallback)")
        if t.dtype != torch.uint8 or not t.is_contiguous() or t.numel() < need:
            raise ValueError(f"a dense stream of {n} elements is a contiguous uint8 tensor of >= {need} bytes")


def binary(op, layout, a, b, n, out=None, stream=None):
    """out = op(a, b) on dense streams of n elements; `out` may be `a` or `b`."""
    code = OP_CODE[op] if isinstance(op, str) else int(op)
    if out is None:
        out = torch.empty_like(a)
    _check_dense(layout, n, a, b, out)
    with torch.cuda.device(a.device):
        _lib.check(_lib.lib().pint_cuda_dense_binary(code, layout.ref, a.data_ptr(), b.data_ptr(), out.data_ptr(), n,
                                                     _stream(stream)), "pint_cuda_dense_binary")
    return out


def binary_multi(op_list, layout, a, b, n, outs=None, stream=None):
    """outs[j] = op_list[j](a, b) on dense streams; op pairs with a one-pass kernel read a and b once."""
    import ctypes as C
    if outs is None:
        outs = [torch.empty_like(a) for _ in op_list]
    _check_dense(layout, n, a, b, *outs)
    codes = (C.c_int * len(op_list))(*[OP_CODE[o] if isinstance(o, str) else int(o) for o in op_list])
    ptrs = (C.c_void_p * len(op_list))(*[o.data_ptr() for o in outs])
    with torch.cuda.device(a.device):
        _lib.check(_lib.lib().pint_cuda_dense_binary_multi(len(op_list), codes, layout.ref, a.data_ptr(), b.data_ptr(), ptrs,
                                                           n, _stream(stream)), "pint_cuda_dense_binary_multi")
    return outs


def shift(left, layout, dense_in, amount, n, out=None, stream=None):
    """out = shift_left (left=True) / shift_right_unsigned of every field by a uniform amount."""
    if out is None:
        out = torch.empty_like(dense_in)
    _check_dense(layout, n, dense_in, out)
    with torch.cuda.device(dense_in.device):
        _lib.check(_lib.lib().pint_cuda_dense_shift(int(bool(left)), layout.ref, int(amount), dense_in.data_ptr(), out.data_ptr(),
                                                    n, _stream(stream)), "pint_cuda_dense_shift")
    return out


def pack(layout, padded, out=None, stream=None):
    """Padded packed_int array -> dense stream (drops the bits above sum(Bits))."""
    n = _check(layout, padded)
    if out is None:
        out = empty(layout, n, padded.device)
    _check_dense(layout, n, out)
    with torch.cuda.device(padded.device):
        _lib.check(_lib.lib().pint_cuda_dense_pack(layout.ref, padded.data_ptr(), out.data_ptr(), n, _stream(stream)),
                   "pint_cuda_dense_pack")
    return out


def unpack(layout, dense, n, out=None, stream=None):
    """Dense stream -> padded packed_int array (zero-extended to Integer)."""
    if out is None:
        out = torch.empty(n, dtype=dtype_of(layout), device=dense.device)
    _check(layout, out)
    _check_dense(layout, n, dense)
    with torch.cuda.device(dense.device):
        _lib.check(_lib.lib().pint_cuda_dense_unpack(layout.ref, dense.data_ptr(), out.data_ptr(), n, _stream(stream)),
                   "pint_cuda_dense_unpack")
    return out


def to_planes(layout, dense, n, signed=False, plane_bytes=4, fields=None, out=None, stream=None):
    """planes[f][i] = get<f> / get_signed<f> of element i of the dense stream (pint.hpp:799-822)."""
    _check_dense(layout, n, dense)
    pdt = _PLANE_DT[(plane_bytes, bool(signed))]
    if out is None:
        out = [torch.empty(n, dtype=pdt, device=dense.device) if (fields is None or f in fields) else None
               for f in range(layout.nfields)]
    ptrs = _plane_array([None if p is None else p.data_ptr() for p in out])
    with torch.cuda.device(dense.device):
        _lib.check(_lib.lib().pint_cuda_dense_to_planes(layout.ref, int(bool(signed)), dense.data_ptr(), ptrs, plane_bytes,
                                                        n, _stream(stream)), "pint_cuda_dense_to_planes")
    return out


def from_planes(layout, planes, out=None, stream=None):
    """dense[i] = packed_int(planes[0][i], ...) -- truncating field constructor (pint.hpp:770-774)."""
    if len(planes) != layout.nfields:
        raise ValueError("need one plane per field")
    pb, n = planes[0].element_size(), planes[0].numel()
    for p in planes:
        if not p.is_cuda or p.element_size() != pb or p.numel() != n or not p.is_contiguous():
            raise ValueError("bad plane tensor")
    if out is None:
        out = empty(layout, n, planes[0].device)
    _check_dense(layout, n, out)
    ptrs = _plane_array([p.data_ptr() for p in planes])
    with torch.cuda.device(out.device):
        _lib.check(_lib.lib().pint_cuda_dense_from_planes(layout.ref, ptrs, pb, out.data_ptr(), n, _stream(stream)),
                   "pint_cuda_dense_from_planes")
    return out
