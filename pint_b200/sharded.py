"""Multi-GPU partitioner: independent contiguous shards, no collective.

`out[i]` depends only on `a[i]`, `b[i]`, so an array is split into contiguous
128-byte-aligned ranges, one per GPU; nothing is exchanged (BASELINE.json
north_star (4)).  Two ways to drive it:
  * one process per GPU (torchrun): `shard_range(n, elem_bytes, world, rank)`;
  * one process, several devices: `sharded_binary(...)`.
"""
import ctypes as C

import torch

from . import _lib
from .ops import OP_CODE, _check


def plan_shards(n, elem_bytes, nshards):
    """[(begin, count)] for each shard; starts are multiples of 128 bytes, remainder to the last."""
    if nshards < 1:
        raise ValueError("nshards must be >= 1")
    begin = (C.c_size_t * nshards)()
    count = (C.c_size_t * nshards)()
    _lib.check(_lib.lib().pint_cuda_plan_shards(n, elem_bytes, nshards, begin, count), "pint_cuda_plan_shards")
    return [(begin[k], count[k]) for k in range(nshards)]


def shard_range(n, elem_bytes, world_size, rank):
    return plan_shards(n, elem_bytes, world_size)[rank]


def sharded_binary(op, layout, a_shards, b_shards, out_shards, streams=None, fused=False):
    """Launch `op` on every device's shard (tensors already resident on their devices).
    Asynchronous; synchronise the devices (or streams) afterwards.  fused=True runs the
    fused unpack->arith->repack kernel (config C4)."""
    k = len(a_shards)
    devices = (C.c_int * k)(*[t.device.index for t in a_shards])
    counts = (C.c_size_t * k)()
    pa, pb, po = (C.c_void_p * k)(), (C.c_void_p * k)(), (C.c_void_p * k)()
    ps = (C.c_void_p * k)()
    for i in range(k):
        counts[i] = _check(layout, a_shards[i], b_shards[i], out_shards[i])
        pa[i], pb[i], po[i] = a_shards[i].data_ptr(), b_shards[i].data_ptr(), out_shards[i].data_ptr()
        if streams is not None:
            ps[i] = streams[i].cuda_stream
        else:
            ps[i] = torch.cuda.current_stream(a_shards[i].device).cuda_stream
    f = _lib.lib().pint_cuda_sharded_fused_unpack_arith_repack if fused else _lib.lib().pint_cuda_sharded_binary
    _lib.check(f(OP_CODE[op], layout.ref, k, devices, pa, pb, po, counts, ps), "pint_cuda_sharded_binary")
    return out_shards


def _stream_array(tensors, streams):
    ps = (C.c_void_p * len(tensors))()
    for i, t in enumerate(tensors):
        ps[i] = streams[i].cuda_stream if streams is not None else torch.cuda.current_stream(t.device).cuda_stream
    return ps


def sharded_unpack(layout, packed_shards, plane_shards, signed=True, streams=None):
    """plane_shards[k][f] (on packed_shards[k]'s device) = field f of shard k; same element ranges."""
    k, nf = len(packed_shards), layout.nfields
    pb = plane_shards[0][0].element_size()
    devices = (C.c_int * k)(*[t.device.index for t in packed_shards])
    counts = (C.c_size_t * k)(*[_check(layout, t) for t in packed_shards])
    pp = (C.c_void_p * k)(*[t.data_ptr() for t in packed_shards])
    planes = (C.c_void_p * (k * nf))()
    for i in range(k):
        for f in range(nf):
            p = plane_shards[i][f]
            planes[i * nf + f] = None if p is None else p.data_ptr()
    _lib.check(_lib.lib().pint_cuda_sharded_unpack(layout.ref, int(bool(signed)), k, devices, pp, planes, pb, counts,
                                                   _stream_array(packed_shards, streams)), "pint_cuda_sharded_unpack")
    return plane_shards


def sharded_repack(layout, plane_shards, packed_shards, streams=None):
    k, nf = len(packed_shards), layout.nfields
    pb = plane_shards[0][0].element_size()
    devices = (C.c_int * k)(*[t.device.index for t in packed_shards])
    counts = (C.c_size_t * k)(*[_check(layout, t) for t in packed_shards])
    pp = (C.c_void_p * k)(*[t.data_ptr() for t in packed_shards])
    planes = (C.c_void_p * (k * nf))()
    for i in range(k):
        for f in range(nf):
            planes[i * nf + f] = plane_shards[i][f].data_ptr()
    _lib.check(_lib.lib().pint_cuda_sharded_repack(layout.ref, k, devices, planes, pb, pp, counts,
                                                   _stream_array(packed_shards, streams)), "pint_cuda_sharded_repack")
    return packed_shards


# ---- dense bit streams (pint_b200.dense): shard starts are multiples of 128 elements ----

def dense_plan_shards(n, nshards):
    """[(begin, count)] in ELEMENTS; every begin is a multiple of 128 elements, so shard k's stream
    starts at byte begin * S / 8 of the global stream (a 16-byte boundary) and is a valid dense
    stream of count elements on its own."""
    if nshards < 1:
        raise ValueError("nshards must be >= 1")
    begin = (C.c_size_t * nshards)()
    count = (C.c_size_t * nshards)()
    _lib.check(_lib.lib().pint_cuda_dense_plan_shards(n, nshards, begin, count), "pint_cuda_dense_plan_shards")
    return [(begin[k], count[k]) for k in range(nshards)]


def dense_shard_range(n, world_size, rank):
    return dense_plan_shards(n, world_size)[rank]


def sharded_dense_binary(op, layout, a_shards, b_shards, out_shards, counts, streams=None):
    """Launch the dense op on every device's shard (uint8 streams resident on their devices)."""
    k = len(a_shards)
    devices = (C.c_int * k)(*[t.device.index for t in a_shards])
    cnt = (C.c_size_t * k)(*counts)
    pa = (C.c_void_p * k)(*[t.data_ptr() for t in a_shards])
    pb = (C.c_void_p * k)(*[t.data_ptr() for t in b_shards])
    po = (C.c_void_p * k)(*[t.data_ptr() for t in out_shards])
    _lib.check(_lib.lib().pint_cuda_sharded_dense_binary(OP_CODE[op], layout.ref, k, devices, pa, pb, po, cnt,
                                                         _stream_array(a_shards, streams)),
               "pint_cuda_sharded_dense_binary")
    return out_shards


# ---- arrays that start on the host or on one device ----

def sharded_binary_multi_host(op_list, layout, devices, a, b, outs):
    """outs[j][i] = op_list[j](a[i], b[i]) for HOST arrays (numpy or CPU tensors, ideally pinned) over
    `devices` of this process: pint_cuda_sharded_binary_multi_host splits [0, n) into contiguous shards,
    one per device, and streams every shard through its device's own staging ring.  Synchronous."""
    def ptr_n(x):
        if isinstance(x, torch.Tensor):
            if x.is_cuda:
                raise TypeError("host arrays required")
            return x.data_ptr(), x.numel(), x.element_size()
        return x.ctypes.data, x.size, x.itemsize
    (pa, n, es), (pb_, nb, _) = ptr_n(a), ptr_n(b)
    if n != nb or es != layout.elem_bytes or len(outs) != len(op_list):
        raise ValueError("host arrays must match the layout and each other")
    codes = (C.c_int * len(op_list))(*[OP_CODE[o] if isinstance(o, str) else int(o) for o in op_list])
    po = (C.c_void_p * len(outs))()
    for j, o in enumerate(outs):
        pj, nj, _ = ptr_n(o)
        if nj != n:
            raise ValueError("output size mismatch")
        po[j] = pj
    devs = (C.c_int * len(devices))(*devices)
    _lib.check(_lib.lib().pint_cuda_sharded_binary_multi_host(len(op_list), codes, layout.ref, len(devices), devs,
                                                              pa, pb_, po, n), "pint_cuda_sharded_binary_multi_host")
    return outs


def enable_peer_access(devices):
    devs = (C.c_int * len(devices))(*devices)
    _lib.check(_lib.lib().pint_cuda_enable_peer_access(len(devices), devs), "pint_cuda_enable_peer_access")


def _move(fn_name, whole, shards, streams):
    k = len(shards)
    devs = (C.c_int * k)(*[t.device.index for t in shards])
    ps = (C.c_void_p * k)(*[t.data_ptr() for t in shards])
    f = getattr(_lib.lib(), fn_name)
    _lib.check(f(whole.data_ptr(), whole.device.index, whole.element_size(), whole.numel(), k, devs, ps,
                 _stream_array(shards, streams)), fn_name)


def scatter_shards(src, devices, streams=None):
    """An array resident on ONE device -> the shards of plan_shards on `devices` (one peer copy per
    shard, asynchronous on each destination's current stream).  Returns the shard tensors."""
    plan = plan_shards(src.numel(), src.element_size(), len(devices))
    shards = [torch.empty(cnt, dtype=src.dtype, device=torch.device("cuda", d)) for (_, cnt), d in zip(plan, devices)]
    _move("pint_cuda_scatter_shards", src, shards, streams)
    return shards


def gather_shards(shards, dst, streams=None):
    """The inverse: shard tensors (one per device) -> the contiguous array `dst` on its device."""
    _move("pint_cuda_gather_shards", dst, shards, streams)
    return dst
