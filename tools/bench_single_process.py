#!/usr/bin/env python
"""tools/bench_single_process.py -- the ONE-process partitioner (north_star (4)): one host thread drives
G = 1, 2, 4, 8 devices through pint_cuda_sharded_binary / pint_cuda_sharded_fused_unpack_arith_repack,
one stream per device, no collective.

For every G and shard size: CUDA-event time per device (events recorded on each device's stream around
its `reps` launches) and the wall clock of the whole loop (launch loop + synchronise), so that the cost
of issuing from one thread shows: efficiency = per-device event time of G = 1 / wall time at G.  At
2^30 elements per GPU the launch loop is free; the smaller shard sizes find where it starts to count.
The whole-job result is checked against the reference's streamed checksum (oracle/_ref), shard by shard.
One JSON line per measurement."""
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pint_b200  # noqa: E402
from pint_b200 import ops, sharded  # noqa: E402
from oracle import Oracle, Ref  # noqa: E402  (checker only)

M64 = 0xFFFFFFFFFFFFFFFF
ndev = torch.cuda.device_count()
ref = Ref() if Ref.available() else None
port = Oracle()
NT = os.cpu_count() or 1


def ref_checksum(op, eb, bits, sa, sb, start, n):
    if ref is not None and ref.has(eb, bits):
        return ref.stream_checksum(op, eb, bits, sa, sb, start, n, nthreads=NT)
    return port.stream_checksum(op, eb, bits, sa, sb, start, n, nthreads=NT)


def run(name, L, bits, op, fused, per_gpu_log2, G, reps, check):
    eb = L.elem_bytes
    dt = {4: torch.uint32, 8: torch.uint64}[eb]
    n = (1 << per_gpu_log2) * G
    plan = sharded.plan_shards(n, eb, G)
    a_s, b_s, o_s, streams = [], [], [], []
    for k, (beg, cnt) in enumerate(plan):
        with torch.cuda.device(k):
            a = torch.empty(cnt, dtype=dt, device=f"cuda:{k}")
            b = torch.empty_like(a)
            ops.fill_splitmix(a, 0x5EED0001, start=beg)
            ops.fill_splitmix(b, 0x5EED0002, start=beg)
            a_s.append(a); b_s.append(b); o_s.append(torch.empty_like(a))
            streams.append(torch.cuda.Stream(device=k))
    for k in range(G):
        torch.cuda.synchronize(k)

    def launch():
        sharded.sharded_binary(op, L, a_s, b_s, o_s, streams=streams, fused=fused)

    for _ in range(3):
        launch()
    for k in range(G):
        torch.cuda.synchronize(k)
    ev = []
    for k in range(G):
        with torch.cuda.device(k):
            ev.append((torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)))
    t0 = time.perf_counter()
    for k in range(G):
        with torch.cuda.device(k):
            ev[k][0].record(streams[k])
    t_issue0 = time.perf_counter()
    for _ in range(reps):
        launch()
    t_issue = time.perf_counter() - t_issue0
    for k in range(G):
        with torch.cuda.device(k):
            ev[k][1].record(streams[k])
    for k in range(G):
        torch.cuda.synchronize(k)
    wall = time.perf_counter() - t0
    dev_ms = [ev[k][0].elapsed_time(ev[k][1]) / reps for k in range(G)]
    line = {"config": name, "devices": G, "elements_per_gpu": 1 << per_gpu_log2, "op": op, "kernel": "fused" if fused else "swar",
            "reps": reps, "wall_ms_per_pass": round(1e3 * wall / reps, 4), "issue_us_per_pass": round(1e6 * t_issue / reps, 1),
            "device_ms_per_pass_max": round(max(dev_ms), 4), "device_ms_per_pass_min": round(min(dev_ms), 4),
            "gbs_whole_job_wall": round(3 * eb * n / (wall / reps) / 1e9, 1),
            "gbs_per_gpu_device_time": round(3 * eb * (n / G) / (max(dev_ms) * 1e-3) / 1e9, 1)}
    if check:
        s = x = rs = rx = 0
        for k, (beg, cnt) in enumerate(plan):
            cs, cx = ops.checksum(o_s[k])
            ws, wx = ref_checksum(op, eb, bits, 0x5EED0001, 0x5EED0002, beg, cnt)
            s, x, rs, rx = (s + cs) & M64, x ^ cx, (rs + ws) & M64, rx ^ wx
        line["whole_job_checksum"] = [s, x]
        line["reference_checksum"] = [rs, rx]
        line["parity_ok"] = (s, x) == (rs, rx)
    print(json.dumps(line), flush=True)
    del a_s, b_s, o_s
    torch.cuda.empty_cache()
    return line


def main():
    big = int(sys.argv[1]) if len(sys.argv) > 1 else 30
    L1, bits1 = pint_b200.Layout(4, 8, 8, 8, 8), (8, 8, 8, 8)
    L4, bits4 = pint_b200.Layout(8, 3, 7, 11, 13, 17, 13), (3, 7, 11, 13, 17, 13)
    base = {}
    Gs = [g for g in (1, 2, 4, 8) if g <= ndev]
    for G in Gs:
        for name, L, bits, op, fused in (("C1 single op", L1, bits1, "add_signed_saturate", False),
                                         ("C4 fused", L4, bits4, "add_signed_saturate", True)):
            r = run(name, L, bits, op, fused, big, G, 10, check=True)
            base.setdefault(name, r["device_ms_per_pass_max"])
            print(json.dumps({"config": name, "devices": G, "efficiency_wall_vs_1gpu_device_time":
                              round(base[name] / r["wall_ms_per_pass"], 4)}), flush=True)
    # where one issuing thread starts to matter: small shards (C0-sized and below)
    for G in Gs:
        for lg in (26, 24, 22, 20, 18):
            run("C1 single op, small shards", L1, bits1, "add_signed_saturate", False, lg, G, 200, check=False)
    ok = True
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
