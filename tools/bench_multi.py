#!/usr/bin/env python
"""tools/bench_multi.py -- BASELINE.json configs C2 and C4 on N GPUs (one process per GPU, torchrun).
  C2: sub_signed_saturate / add_signed_saturate on 2^31 packed_int<uint64_t,16,16,16,16> (16 GiB per
      array) SHARDED over the ranks (strong scaling: the total is fixed).
  C4: fused unpack -> add_signed_saturate -> repack on packed_int<uint64_t,3,7,11,13,17,13>, 2^30
      elements (8 GiB per array) PER GPU (weak scaling: 64 GiB of operand a at 8 GPUs).
No collective on the data path; NCCL only for the barrier and the max-over-ranks of the CUDA-event time.
One JSON line per measurement from rank 0."""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pint_b200  # noqa: E402
from pint_b200 import ops, sharded  # noqa: E402

rank = int(os.environ.get("RANK", "0"))
local = int(os.environ.get("LOCAL_RANK", "0"))
world = int(os.environ.get("WORLD_SIZE", "1"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)


def barrier():
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def timed(fn, reps=10, warm=3):
    for _ in range(warm):
        fn()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1) / reps
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    return ms


def run(config, L, total, per_rank_begin_count, op, fused, scaling):
    begin, count = per_rank_begin_count
    a = torch.empty(count, dtype=torch.uint64, device=dev)
    b = torch.empty_like(a)
    out = torch.empty_like(a)
    ops.fill_splitmix(a, 0x5EED0001, start=begin)
    ops.fill_splitmix(b, 0x5EED0002, start=begin)
    ms = timed(lambda: ops.binary(op, L, a, b, out=out, fused=fused))
    s, x = ops.checksum(out)
    if rank == 0:
        print(json.dumps({"config": config, "n_gpus": world, "op": op, "kernel": "fused" if fused else "swar", "scaling": scaling,
                          "elements_total": total, "ms": round(ms, 4), "gelem_s_whole_job": round(total / ms / 1e6, 2),
                          "gbs_whole_job": round(24 * total / ms / 1e6, 1), "gbs_per_gpu": round(24 * total / ms / 1e6 / world, 1),
                          "rank0_checksum": [s, x]}), flush=True)
    del a, b, out
    torch.cuda.empty_cache()


which = sys.argv[1:] or ["C2", "C4"]
if "C2" in which:
    L = pint_b200.Layout(8, 16, 16, 16, 16)
    total = 1 << 31
    rng = sharded.shard_range(total, 8, world, rank)
    for op in ("sub_signed_saturate", "add_signed_saturate"):
        run("C2", L, total, rng, op, False, "strong")
if "C4" in which:
    L = pint_b200.Layout(8, 3, 7, 11, 13, 17, 13)
    per = 1 << 30
    rng = sharded.shard_range(per * world, 8, world, rank)
    run("C4", L, per * world, rng, "add_signed_saturate", True, "weak")
if world > 1:
    dist.destroy_process_group()
