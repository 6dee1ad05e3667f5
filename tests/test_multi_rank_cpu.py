"""world_size-2 gloo test (CPU) of the N>1 host logic bench.py and a torchrun user
rely on: every rank derives its contiguous shard from the same planner, generates
its slice of the synthetic input from (seed, global start), and the per-shard
results stitch into exactly the single-process result.  The per-element
arithmetic is done by the oracle here (this is a host-logic test; the CUDA
kernels are covered by the -m gpu tests)."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from pint_b200 import sharded

BITS = (3, 7, 11, 13, 17, 13)
N = 1_000_003


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import Oracle
    o = Oracle()
    begin, count = sharded.shard_range(N, 8, world, rank)
    a = o.fill(8, count, 0x5EED0001, start=begin)
    b = o.fill(8, count, 0x5EED0002, start=begin)
    out = o.binary("add_signed_saturate", 8, BITS, a, b)
    s, x = o.checksum(out)
    # sum mod 2^64 travels as two 32-bit halves (gloo has no uint64 reduction); xor is reduced with BXOR
    t = torch.tensor([s & 0xFFFFFFFF, s >> 32, count], dtype=torch.int64)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    xr = torch.tensor([x & 0x7FFFFFFFFFFFFFFF, x >> 63], dtype=torch.int64)
    dist.all_reduce(xr, op=dist.ReduceOp.BXOR)
    # the time a bench would report is the max over ranks
    tm = torch.tensor([1.0 + rank], dtype=torch.float64)
    dist.all_reduce(tm, op=dist.ReduceOp.MAX)
    if rank == 0:
        total = (int(t[0]) + (int(t[1]) << 32)) & 0xFFFFFFFFFFFFFFFF
        q.put((total, int(xr[0]) | (int(xr[1]) << 63), int(t[2]), float(tm[0])))
    dist.barrier()
    dist.destroy_process_group()


def test_two_ranks_stitch_to_single_process_result():
    from oracle import Oracle
    o = Oracle()
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    total, xr, count, tmax = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    a, b = o.fill(8, N, 0x5EED0001), o.fill(8, N, 0x5EED0002)
    want_s, want_x = o.checksum(o.binary("add_signed_saturate", 8, BITS, a, b))
    assert count == N
    assert (total, xr) == (want_s, want_x)
    assert tmax == 2.0


def test_shards_cover_array_without_overlap():
    for world in (1, 2, 4, 8):
        plan = sharded.plan_shards(N, 8, world)
        covered = np.zeros(N, dtype=np.uint8)
        for beg, cnt in plan:
            covered[beg:beg + cnt] += 1
        assert (covered == 1).all()


# ---- dense bit streams: shards are cut on 128-element boundaries and stitch byte for byte ----

DENSE_BITS = (7, 9, 3)
DENSE_N = 300_007


def _dense_worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import Oracle
    o = Oracle()
    S = sum(DENSE_BITS)
    begin, count = sharded.dense_shard_range(DENSE_N, world, rank)
    a = o.dense_pack(4, DENSE_BITS, o.fill(4, count, 0x5EED0001, start=begin))
    b = o.dense_pack(4, DENSE_BITS, o.fill(4, count, 0x5EED0002, start=begin))
    out = o.dense_binary("sub_signed_saturate", 4, DENSE_BITS, a, b, count)
    # every rank contributes its bytes at byte offset begin * S / 8 of the global stream
    total_bytes = o.dense_bytes(DENSE_BITS, DENSE_N)
    buf = torch.zeros(total_bytes, dtype=torch.int32)
    off = begin * S // 8
    assert (begin * S) % 128 == 0
    valid = (count * S + 7) // 8
    buf[off:off + valid] = torch.from_numpy(out[:valid].astype(np.int32))
    dist.all_reduce(buf, op=dist.ReduceOp.SUM)
    if rank == 0:
        q.put(buf.numpy().astype(np.uint8))
    dist.barrier()
    dist.destroy_process_group()


def test_dense_shards_stitch_byte_for_byte():
    from oracle import Oracle
    o = Oracle()
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_dense_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    a = o.dense_pack(4, DENSE_BITS, o.fill(4, DENSE_N, 0x5EED0001))
    b = o.dense_pack(4, DENSE_BITS, o.fill(4, DENSE_N, 0x5EED0002))
    want = o.dense_binary("sub_signed_saturate", 4, DENSE_BITS, a, b, DENSE_N)
    np.testing.assert_array_equal(got, want)
    for w in (1, 2, 3, 8):
        plan = sharded.dense_plan_shards(DENSE_N, w)
        assert sum(c for _, c in plan) == DENSE_N and all(bg % 128 == 0 for bg, _ in plan)
        assert all(plan[k][0] + plan[k][1] == plan[k + 1][0] for k in range(w - 1))
