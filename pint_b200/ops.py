"""pint_b200.ops -- batched pint operations on device arrays.

Same names and argument meaning as the reference's scalar API
(/root/reference/include/pint/pint.hpp:799-1029, README.md:88-279), applied
elementwise: `add_signed_saturate(layout, a, b)` is
`out[i] = pint::add_signed_saturate(P(a[i]), P(b[i])).value()`.

Arrays are torch CUDA tensors whose element size equals sizeof(Integer)
(torch.uint8/16/32/64 or the signed dtypes of the same width; only the bits
matter).  torch is used for device memory and streams only: every function is
one call into libpint_cuda.so on the current torch stream.  There is no CPU
path: CPU tensors are rejected.
"""
import ctypes as C

import torch

from . import _lib
from .layout import Layout

OPS = [
    "add_wrap", "add_unsigned_saturate", "add_signed_saturate",
    "sub_wrap", "sub_unsigned_saturate", "sub_signed_saturate",
    "bit_or", "bit_and", "bit_xor",
    "min_unsigned", "max_unsigned", "min_signed", "max_signed",
]
OP_CODE = {name: i for i, name in enumerate(OPS)}

_UDT = {1: torch.uint8, 2: torch.uint16, 4: torch.uint32, 8: torch.uint64}
_PLANE_DT = {(1, True): torch.int8, (1, False): torch.uint8, (2, True): torch.int16, (2, False): torch.uint16,
             (4, True): torch.int32, (4, False): torch.uint32, (8, True): torch.int64, (8, False): torch.uint64}


def dtype_of(layout):
    return _UDT[layout.elem_bytes]


def _stream(stream):
    if stream is None:
        return torch.cuda.current_stream().cuda_stream
    return stream.cuda_stream if hasattr(stream, "cuda_stream") else int(stream)


def _check(layout, *tensors):
    n = None
    for t in tensors:
        if not isinstance(t, torch.Tensor) or not t.is_cuda:
            raise TypeError("pint_b200 operates on CUDA tensors only (there is no CPU fallback)")
        if t.element_size() != layout.elem_bytes:
            raise TypeError(f"tensor element size {t.element_size()} != sizeof(Integer) {layout.elem_bytes} of {layout}")
        if not t.is_contiguous():
            raise ValueError("arrays must be contiguous")
        if n is None:
            n = t.numel()
        elif t.numel() != n:
            raise ValueError("arrays must have the same number of elements")
    return n


def binary(op, layout, a, b, out=None, stream=None, fused=False):
    """out[i] = op(a[i], b[i]); `out` may be `a` or `b` (in place)."""
    code = OP_CODE[op] if isinstance(op, str) else int(op)
    if out is None:
        out = torch.empty_like(a)
    n = _check(layout, a, b, out)
    f = _lib.lib().pint_cuda_fused_unpack_arith_repack if fused else _lib.lib().pint_cuda_binary
    with torch.cuda.device(a.device):
        _lib.check(f(code, layout.ref, a.data_ptr(), b.data_ptr(), out.data_ptr(), n, _stream(stream)),
                   "pint_cuda_fused_unpack_arith_repack" if fused else "pint_cuda_binary")
    return out


def binary_multi(op_list, layout, a, b, outs=None, stream=None):
    """outs[j] = op_list[j](a, b) for up to 4 ops in ONE pass over a and b (pint_cuda_binary_multi)."""
    if outs is None:
        outs = [torch.empty_like(a) for _ in op_list]
    n = _check(layout, a, b, *outs)
    if len(outs) != len(op_list):
        raise ValueError("need one output per op")
    codes = (C.c_int * len(op_list))(*[OP_CODE[o] if isinstance(o, str) else int(o) for o in op_list])
    ptrs = _plane_array([o.data_ptr() for o in outs])
    with torch.cuda.device(a.device):
        _lib.check(_lib.lib().pint_cuda_binary_multi(len(op_list), codes, layout.ref, a.data_ptr(), b.data_ptr(), ptrs, n,
                                                     _stream(stream)), "pint_cuda_binary_multi")
    return outs


def _make_binary(name):
    def fn(layout, a, b, out=None, stream=None):
        return binary(name, layout, a, b, out=out, stream=stream)
    fn.__name__ = name
    fn.__doc__ = f"Batched pint::{name} (see include/pint_cuda.h for the pint.hpp lines)."
    return fn


for _name in OPS:
    globals()[_name] = _make_binary(_name)


def fused_unpack_arith_repack(op, layout, a, b, out=None, stream=None):
    """Same result as `binary`, computed as unpack -> per-field arithmetic -> repack in one kernel."""
    return binary(op, layout, a, b, out=out, stream=stream, fused=True)


def _shift(left, layout, x, amount, out, stream):
    if amount < 0:
        raise ValueError("shift amount must be non-negative")
    if out is None:
        out = torch.empty_like(x)
    n = _check(layout, x, out)
    with torch.cuda.device(x.device):
        _lib.check(_lib.lib().pint_cuda_shift(int(left), layout.ref, int(amount), x.data_ptr(), out.data_ptr(), n,
                                              _stream(stream)), "pint_cuda_shift")
    return out


def shift_left(layout, x, amount, out=None, stream=None):
    """Batched pint::shift_left with a uniform amount (pint.hpp:1006-1017)."""
    return _shift(True, layout, x, amount, out, stream)


def shift_right_unsigned(layout, x, amount, out=None, stream=None):
    """Batched pint::shift_right_unsigned (pint.hpp:1019-1029)."""
    return _shift(False, layout, x, amount, out, stream)


def slice(layout, start, end, x, out=None, stream=None):  # noqa: A001 - mirrors pint::slice
    """Batched pint::slice<start, end> (pint.hpp:923-932).  Returns (array, result layout)."""
    result_layout = layout.sliced(start, end)
    if out is None:
        out = torch.empty_like(x)
    n = _check(layout, x, out)
    with torch.cuda.device(x.device):
        _lib.check(_lib.lib().pint_cuda_slice(layout.ref, start, end, x.data_ptr(), out.data_ptr(), n, _stream(stream)),
                   "pint_cuda_slice")
    return out, result_layout


def _plane_array(ptrs):
    arr = (C.c_void_p * len(ptrs))()
    for i, p in enumerate(ptrs):
        arr[i] = p
    return arr


def unpack(layout, packed, signed=False, plane_bytes=4, fields=None, out=None, stream=None):
    """planes[f][i] = get<f>(packed[i]) (pint.hpp:799-807) or get_signed<f> (:809-822).

    Returns a list with one tensor per field (None for fields not in `fields`)."""
    n = _check(layout, packed)
    if plane_bytes not in (1, 2, 4, 8):
        raise ValueError("plane_bytes must be 1, 2, 4 or 8")
    pdt = _PLANE_DT[(plane_bytes, bool(signed))]
    if out is None:
        out = [torch.empty(n, dtype=pdt, device=packed.device) if (fields is None or f in fields) else None
               for f in range(layout.nfields)]
    if len(out) != layout.nfields:
        raise ValueError("need one plane (or None) per field")
    for p in out:
        if p is not None and (not p.is_cuda or p.element_size() != plane_bytes or p.numel() != n or not p.is_contiguous()):
            raise ValueError("bad plane tensor")
    ptrs = _plane_array([None if p is None else p.data_ptr() for p in out])
    with torch.cuda.device(packed.device):
        _lib.check(_lib.lib().pint_cuda_unpack(layout.ref, int(bool(signed)), packed.data_ptr(), ptrs, plane_bytes, n,
                                               _stream(stream)), "pint_cuda_unpack")
    return out


def unpack_signed(layout, packed, **kw):
    return unpack(layout, packed, signed=True, **kw)


def repack(layout, planes, out=None, stream=None):
    """packed[i] = packed_int(planes[0][i], planes[1][i], ...) -- truncating field constructor (pint.hpp:770-774)."""
    if len(planes) != layout.nfields:
        raise ValueError("need one plane per field")
    pb = planes[0].element_size()
    n = planes[0].numel()
    for p in planes:
        if not p.is_cuda or p.element_size() != pb or p.numel() != n or not p.is_contiguous():
            raise ValueError("bad plane tensor")
    if out is None:
        out = torch.empty(n, dtype=dtype_of(layout), device=planes[0].device)
    _check(layout, out)
    ptrs = _plane_array([p.data_ptr() for p in planes])
    with torch.cuda.device(out.device):
        _lib.check(_lib.lib().pint_cuda_repack(layout.ref, ptrs, pb, out.data_ptr(), n, _stream(stream)),
                   "pint_cuda_repack")
    return out


def pairs_reduce(op, layout, pairs, stream=None):
    """sum_i op(pairs[2i], pairs[2i+1]) mod 2^64 -- the loop of tests/pint_bench.cpp:90-94.  Returns an int."""
    _check(layout, pairs)
    res = torch.zeros(1, dtype=torch.int64, device=pairs.device)
    with torch.cuda.device(pairs.device):
        _lib.check(_lib.lib().pint_cuda_pairs_reduce(OP_CODE[op], layout.ref, pairs.data_ptr(), pairs.numel() // 2,
                                                     res.data_ptr(), _stream(stream)), "pint_cuda_pairs_reduce")
    return int(res.item()) & 0xFFFFFFFFFFFFFFFF


def binary_pairs(op, layout, pairs, out=None, stream=None):
    """out[i] = op(pairs[2i], pairs[2i+1]): interleaved operands (the benchmark's std::pair array,
    tests/pint_bench.cpp:22) in, plain array out."""
    _check(layout, pairs)
    n = pairs.numel() // 2
    if out is None:
        out = torch.empty(n, dtype=pairs.dtype, device=pairs.device)
    _check(layout, out)
    if out.numel() != n:
        raise ValueError("out must hold one element per pair")
    with torch.cuda.device(pairs.device):
        _lib.check(_lib.lib().pint_cuda_binary_pairs(OP_CODE[op] if isinstance(op, str) else int(op), layout.ref,
                                                     pairs.data_ptr(), out.data_ptr(), n, _stream(stream)),
                   "pint_cuda_binary_pairs")
    return out


def binary_scalar(op, layout, a, scalar, out=None, stream=None, scalar_first=False):
    """out[i] = op(a[i], P(scalar)) -- or op(P(scalar), a[i]) with scalar_first -- for one packed_int
    value `scalar` (an int, packed_int::value()); `out` may be `a`."""
    if out is None:
        out = torch.empty_like(a)
    n = _check(layout, a, out)
    with torch.cuda.device(a.device):
        _lib.check(_lib.lib().pint_cuda_binary_scalar(OP_CODE[op] if isinstance(op, str) else int(op), layout.ref,
                                                      a.data_ptr(), int(scalar), 1 if scalar_first else 0,
                                                      out.data_ptr(), n, _stream(stream)), "pint_cuda_binary_scalar")
    return out


def equal(layout, a, b, flags=True, count=True, out=None, stream=None):
    """packed_int::operator== over arrays: (flags, not_equal) with flags[i] = (a[i] == b[i]) as a uint8
    tensor (None unless `flags`) and not_equal = number of differing elements (None unless `count`)."""
    n = _check(layout, a, b)
    f = None
    if flags:
        f = out if out is not None else torch.empty(n, dtype=torch.uint8, device=a.device)
        if f.numel() != n or f.element_size() != 1 or not f.is_cuda or not f.is_contiguous():
            raise ValueError("flags must be n contiguous bytes on the device")
    c = torch.zeros(1, dtype=torch.int64, device=a.device) if count else None
    with torch.cuda.device(a.device):
        _lib.check(_lib.lib().pint_cuda_equal(layout.ref, a.data_ptr(), b.data_ptr(), n, f.data_ptr() if flags else None,
                                              c.data_ptr() if count else None, _stream(stream)), "pint_cuda_equal")
    return f, (int(c.item()) if count else None)


def fill_splitmix(out, seed, start=0, stream=None):
    """out[i] = (Integer) splitmix64(seed + start + i): the synthetic input of SURVEY.md 8(d)."""
    if not out.is_cuda or not out.is_contiguous():
        raise TypeError("CUDA contiguous tensor required")
    with torch.cuda.device(out.device):
        _lib.check(_lib.lib().pint_cuda_fill_splitmix(out.data_ptr(), out.element_size(), out.numel(),
                                                      seed & 0xFFFFFFFFFFFFFFFF, start, _stream(stream)),
                   "pint_cuda_fill_splitmix")
    return out


def checksum(x, stream=None):
    """(sum, xor) of the elements zero-extended to 64 bit."""
    if not x.is_cuda or not x.is_contiguous():
        raise TypeError("CUDA contiguous tensor required")
    res = torch.zeros(2, dtype=torch.int64, device=x.device)
    with torch.cuda.device(x.device):
        _lib.check(_lib.lib().pint_cuda_checksum(x.data_ptr(), x.element_size(), x.numel(), res.data_ptr(),
                                                 _stream(stream)), "pint_cuda_checksum")
    s, r = res.tolist()
    return s & 0xFFFFFFFFFFFFFFFF, r & 0xFFFFFFFFFFFFFFFF


def binary_host(op, layout, a, b, out):
    """End-to-end call with HOST buffers (numpy arrays or pinned CPU tensors): H2D, kernel, D2H
    pipelined inside the library (pint_cuda_binary_host).  Synchronous."""
    def ptr_n(x):
        if isinstance(x, torch.Tensor):
            if x.is_cuda:
                raise TypeError("binary_host takes host buffers")
            return x.data_ptr(), x.numel(), x.element_size()
        return x.ctypes.data, x.size, x.itemsize
    (pa, n, es), (pb, nb, _), (po, no, _) = ptr_n(a), ptr_n(b), ptr_n(out)
    if not (n == nb == no) or es != layout.elem_bytes:
        raise ValueError("host arrays must match the layout and each other")
    code = OP_CODE[op] if isinstance(op, str) else int(op)
    _lib.check(_lib.lib().pint_cuda_binary_host(code, layout.ref, pa, pb, po, n), "pint_cuda_binary_host")
    return out


def binary_multi_host(op_list, layout, a, b, outs):
    """Several ops on the same HOST inputs with one upload (pint_cuda_binary_multi_host)."""
    def ptr_n(x):
        if isinstance(x, torch.Tensor):
            return x.data_ptr(), x.numel(), x.element_size()
        return x.ctypes.data, x.size, x.itemsize
    (pa, n, es), (pb, nb, _) = ptr_n(a), ptr_n(b)
    if n != nb or es != layout.elem_bytes or len(outs) != len(op_list):
        raise ValueError("host arrays must match the layout and each other")
    codes = (C.c_int * len(op_list))(*[OP_CODE[o] if isinstance(o, str) else int(o) for o in op_list])
    po = (C.c_void_p * len(outs))()
    for j, o in enumerate(outs):
        pj, nj, _ = ptr_n(o)
        if nj != n:
            raise ValueError("output size mismatch")
        po[j] = pj
    _lib.check(_lib.lib().pint_cuda_binary_multi_host(len(op_list), codes, layout.ref, pa, pb, po, n),
               "pint_cuda_binary_multi_host")
    return outs


def launch_count():
    return _lib.lib().pint_cuda_launch_count()


def set_option(key, value):
    _lib.check(_lib.lib().pint_cuda_set_option(key.encode(), int(value)), f"pint_cuda_set_option({key})")


def get_option(key):
    return _lib.lib().pint_cuda_get_option(key.encode())
