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


def sharded_binary(op, layout, a_shards, b_shards, out_shards, streams=None):
    """Launch `op` on every device's shard (tensors already resident on their devices).
    Asynchronous; synchronise the devices (or streams) afterwards."""
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
    _lib.check(_lib.lib().pint_cuda_sharded_binary(OP_CODE[op], layout.ref, k, devices, pa, pb, po, counts, ps),
               "pint_cuda_sharded_binary")
    return out_shards
