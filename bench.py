#!/usr/bin/env python
"""bench.py -- BASELINE.json's metric on BASELINE.json's config, on B200.

Workload (config C1, the one BASELINE.json's target is quoted on for 1 GPU):
  add_unsigned_saturate + add_signed_saturate over 2^30
  packed_int<uint32_t,8,8,8,8> (RGBA8) per GPU.  One "step" = one pass of each
  op over the batch, 2 * 2^30 packed elements processed.  The step is ONE launch of
  pint_cuda_binary_multi: a and b are read once and both results are written
  (16 algorithmic bytes per element instead of 2 * 12); the two-launch form is
  timed beside it and reported as "separate_passes".
  N > 1: every rank owns an independent contiguous shard of the same size (weak
  scaling, no collective on the data path -- nothing reduces).

  value     packed elements/s, whole job, inputs resident in HBM (CUDA events, max over ranks).
            Counted per op RESULT: a step yields 2 * 2^30 result elements per GPU (both arms count
            the same way); "packed_elements_per_s" beside it counts every input element once.
  configs   the other BASELINE.json configs on the same ranks: C0 (2^26 u16 <5,6,5> add_wrap, also as a
            CUDA graph of back-to-back launches), C2 (2^31 u64 <16,16,16,16> STRONG-scaled over the
            ranks), C3 (irregular u64 unpack -> 6 int32 planes, repack), C4 (fused, 2^30 per GPU)
  parity    every rank checksums its shard of the C1 / C2 / C4 outputs on the device and compares it
            with the reference's streamed checksum of the same index range (oracle/_ref: the
            unmodified header over the same splitmix inputs; the checker, never the thing timed);
            the per-rank (sum, xor) pairs are gathered into whole_job_checksum / parity_ok
  e2e       same metric through the C ABI's host-buffer entry point
            (pint_cuda_binary_host): pinned host arrays in, host array out, the
            H2D/D2H copies inside the timed region
  roofline  the streaming kernel against the measured HBM copy bandwidth
  cpu_baseline  the reference's own scalar functions in the tests/pint_bench.cpp
            loop shape on this box's host cores (oracle/_ref), rank 0, N=1 only

A parity mismatch prints the line with "parity_ok": false and exits 1.

`--impl reference` times only the CPU reference (all host threads) on a bounded
sample of the same workload.
"""
import argparse
import ctypes as C
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

BITS = (8, 8, 8, 8)
ELEM_BYTES = 4
OPS = ("add_unsigned_saturate", "add_signed_saturate")
BYTES_PER_ELEM = 3 * ELEM_BYTES  # binary op on W-byte packs moves 3W bytes
METRIC = "packed elements/sec"
FALLBACK_HBM_GBS = 6650.0        # /opt/skills/guides/B200_PROFILING.md fallback


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return FALLBACK_HBM_GBS, "fallback (B200_PROFILING.md)"


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons of one GPU while the timed region runs."""

    def __init__(self, index, period=0.1):
        super().__init__(daemon=True)
        self.index, self.period = index, period
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._halt = threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {
            nv.nvmlClocksEventReasonHwSlowdown if hasattr(nv, "nvmlClocksEventReasonHwSlowdown") else 0x8: "hw_slowdown",
            0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap",
        }
        while not self._halt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if bit and (r & bit):
                        self.reasons.add(name)
            except Exception:
                pass
            self._halt.wait(self.period)

    def stop(self):
        self._halt.set()
        self.join(timeout=2)
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(s)}


def cpu_reference_run(n_sample, reps, threads=None):
    """The reference's scalar functions (oracle/_ref, else the oracle port) over a bounded sample.
    Returns (elements/s over both ops, kind, cores, description)."""
    import numpy as np
    from oracle import Oracle, Ref
    o = Oracle()
    a = o.fill(ELEM_BYTES, n_sample, 0x5EED0001)
    b = o.fill(ELEM_BYTES, n_sample, 0x5EED0002)
    out = np.empty_like(a)
    cores = threads or os.cpu_count() or 1
    if Ref.available():
        r = Ref()
        kind = "reference"

        def run(op):
            r.binary(op, ELEM_BYTES, BITS, a, b, nthreads=cores, out=out)
    else:
        kind, cores = "port", 1

        def run(op):
            o.binary(op, ELEM_BYTES, BITS, a, b)
    for op in OPS:
        run(op)  # warm: page in, spin up
    best = float("inf")
    for _ in range(reps):
        t0 = time.perf_counter()
        for op in OPS:
            run(op)
        best = min(best, time.perf_counter() - t0)
    return len(OPS) * n_sample / best, kind, cores, best


def run_reference_arm(args, rank):
    if rank != 0:
        return 0
    n_sample = 1 << args.cpu_log2
    # each "step" is one pass of both ops over the sample
    import numpy as np  # noqa: F401
    from oracle import Oracle, Ref
    o = Oracle()
    a = o.fill(ELEM_BYTES, n_sample, 0x5EED0001)
    b = o.fill(ELEM_BYTES, n_sample, 0x5EED0002)
    out = a.copy()
    cores = os.cpu_count() or 1
    if Ref.available():
        r, kind = Ref(), "reference"
        step = lambda: [r.binary(op, ELEM_BYTES, BITS, a, b, nthreads=cores, out=out) for op in OPS]  # noqa: E731
    else:
        kind, cores = "port", 1
        step = lambda: [o.binary(op, ELEM_BYTES, BITS, a, b) for op in OPS]  # noqa: E731
    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = time.perf_counter() - t0
    value = args.steps * len(OPS) * n_sample / dt
    sample = f"{'+'.join(OPS)} over 2^{args.cpu_log2} elements per step, {cores} threads, out[i]=op(a[i],b[i])"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "elements/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u32", "data": "synthetic",
        # the same config object as the GPU arm prints (the CPU loop is DRAM-bound, so its per-element
        # rate on the bounded sample holds for the full arrays; the sample is named in cpu_baseline)
        "config": workload_config(args, max(1, args.gpus)),
        "cpu_baseline": {"value": value, "unit": "elements/s", "cores": cores, "kind": kind, "sample": sample,
                         "sample_elements": n_sample},
        "e2e": {"value": value, "unit": "elements/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


def workload_config(args, world):
    return {
        "workload": "C1: add_unsigned_saturate + add_signed_saturate over 2^%d packed_int<uint32_t,8,8,8,8> per GPU, "
                    "both ops per step (one pint_cuda_binary_multi launch: a,b read once, two outputs)" % args.log2,
        "layout": "packed_int<uint32_t,8,8,8,8>", "elements_per_gpu": 1 << args.log2,
        "elements_per_step": len(OPS) * (1 << args.log2) * world,
        "value_counts": "op results: %d ops x elements; packed_elements_per_s counts each input element once" % len(OPS),
        "bytes_per_element": (2 + len(OPS)) * ELEM_BYTES, "bytes_per_element_single_op_pass": BYTES_PER_ELEM,
        "parallelism": "%d independent contiguous shard(s), no collective" % world,
        "l2": "inputs (2 x %.1f GiB per GPU) far larger than the 126 MB L2; no flush needed" % ((1 << args.log2) * 4 / 2 ** 30),
    }


M64 = 0xFFFFFFFFFFFFFFFF


class Checker:
    """The reference's streamed checksum of OP over an index range of the synthetic inputs
    (oracle/_ref = the unmodified header when built, else the oracle port).  Checker only."""

    def __init__(self, threads):
        from oracle import Oracle, Ref
        self.threads = max(1, threads)
        self.ref = Ref() if Ref.available() else None
        self.port = Oracle()
        self.kind = "reference (oracle/_ref: unmodified pint.hpp)" if self.ref else "port (oracle/pint_oracle.c)"

    def stream_checksum(self, op, elem_bytes, bits, seed_a, seed_b, start, n):
        if self.ref is not None and self.ref.has(elem_bytes, bits):
            return self.ref.stream_checksum(op, elem_bytes, bits, seed_a, seed_b, start, n, nthreads=self.threads)
        return self.port.stream_checksum(op, elem_bytes, bits, seed_a, seed_b, start, n, nthreads=self.threads)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--log2", type=int, default=30, help="log2 of packed elements per GPU")
    ap.add_argument("--e2e-log2", type=int, default=None,
                    help="log2 of elements per GPU in the host-buffer (e2e) leg (default: --log2 at N=1, 28 at N>1: "
                         "N x 16 GiB of pinned host memory is more than the box should be asked for)")
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--cpu-log2", type=int, default=27, help="log2 of the CPU baseline sample")
    ap.add_argument("--c2-log2", type=int, default=31, help="log2 of the WHOLE C2 array (strong-scaled over the ranks)")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="skip the C0/C2/C3/C4 legs")
    ap.add_argument("--no-parity", action="store_true", help="skip the streamed reference checksums")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    if args.impl == "reference":
        return run_reference_arm(args, rank)

    import torch
    import torch.distributed as dist
    import pint_b200
    from pint_b200 import _lib, ops, sharded

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: libpint_cuda.so has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def gather_u64(values):
        """every rank's list of uint64 -> [rank][i] on every rank (NCCL has no xor reduction: gather, combine on the host)"""
        if world == 1:
            return [list(values)]
        t = torch.tensor([v - (1 << 64) if v >= (1 << 63) else v for v in values], dtype=torch.int64, device=dev)
        out = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(out, t)
        return [[int(x) & M64 for x in o.tolist()] for o in out]

    def combine(pairs):
        s = x = 0
        for ps, px in pairs:
            s = (s + ps) & M64
            x ^= px
        return [s, x]

    cpu_threads = max(1, (os.cpu_count() or 1) // world)
    checker = None if args.no_parity else Checker(cpu_threads)
    parity = {"checker": checker.kind if checker else None, "cpu_threads_per_rank": cpu_threads, "checks": []}

    def parity_check(name, op, elem_bytes, bits, seed_a, seed_b, begin, count, out_tensor):
        """device checksum of this rank's output shard vs the reference's streamed checksum of [begin, begin+count);
        gathered over the ranks: whole-job checksums and one flag"""
        got = ops.checksum(out_tensor)
        t0 = time.perf_counter()
        want = checker.stream_checksum(op, elem_bytes, bits, seed_a, seed_b, begin, count) if checker else got
        cpu_s = time.perf_counter() - t0
        rows = gather_u64([got[0], got[1], want[0], want[1]])
        whole_gpu = combine([(r[0], r[1]) for r in rows])
        whole_ref = combine([(r[2], r[3]) for r in rows])
        ok = all(r[0] == r[2] and r[1] == r[3] for r in rows)
        entry = {"config": name, "op": op, "elements_checked_per_rank": count, "ranks": world,
                 "whole_job_checksum": whole_gpu, "reference_checksum": whole_ref if checker else None,
                 "parity_ok": bool(ok) if checker else None, "cpu_seconds": round(cpu_s, 2)}
        parity["checks"].append(entry)
        return entry

    peak, peak_src = measured_peak()

    def timed_launches(fn, launches, warm=3):
        """ms per launch (max over ranks): `launches` calls queued back to back between two events"""
        for _ in range(warm):
            fn(0)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(launches):
            fn(i)
        e1.record()
        barrier()
        return max_over_ranks(e0.elapsed_time(e1)) / launches

    def config_line(name, workload, scaling, elems_rank, elems_job, bytes_per_elem, ms, **extra):
        gbs = elems_rank * bytes_per_elem / (ms * 1e-3) / 1e9   # this rank's shard (the largest, by the max-over-ranks time)
        return dict(config=name, workload=workload, scaling=scaling, elements_per_gpu=elems_rank, elements_whole_job=elems_job,
                    bytes_per_element=bytes_per_elem, ms=round(ms, 4), gbs_per_gpu=round(gbs, 1),
                    frac_measured_peak=round(gbs / peak, 4), frac_spec_8000=round(gbs / 8000.0, 4),
                    gelem_s_whole_job=round(elems_job / (ms * 1e-3) / 1e9, 2), **extra)

    L = pint_b200.Layout(ELEM_BYTES, *BITS)
    assert L.is_specialized(), "RGBA8 must run the compiled-in-mask kernels"
    n = 1 << args.log2
    # this rank's shard of the global array of world * n elements
    begin, count = sharded.shard_range(n * world, ELEM_BYTES, world, rank)
    a = torch.empty(count, dtype=torch.uint32, device=dev)
    b = torch.empty(count, dtype=torch.uint32, device=dev)
    outs = [torch.empty(count, dtype=torch.uint32, device=dev) for _ in OPS]
    ops.fill_splitmix(a, 0x5EED0001, start=begin)
    ops.fill_splitmix(b, 0x5EED0002, start=begin)

    def step_separate():
        for op, o in zip(OPS, outs):
            ops.binary(op, L, a, b, out=o)

    def step():
        # both ops of the step in ONE pass over a and b (pint_cuda_binary_multi): 2 reads + 2 writes per element
        ops.binary_multi(OPS, L, a, b, outs=outs)

    def timed_region(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        return max_over_ranks(e0.elapsed_time(e1))

    # reference point: the same step as two separate kernel passes (3W bytes per element each)
    for _ in range(args.warmup):
        step_separate()
    ms_separate = timed_region(step_separate, args.steps)

    for _ in range(args.warmup):
        step()
    barrier()
    sampler = ClockSampler(local_rank, period=0.01)
    sampler.start()
    launches0 = ops.launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(args.steps):
        step()
    ev1.record()
    barrier()
    launches = ops.launch_count() - launches0
    clocks = sampler.stop()
    ms = max_over_ranks(ev0.elapsed_time(ev1))
    total_elems = len(OPS) * args.steps * n * world
    value = total_elems / (ms * 1e-3)
    # dominant (only) kernel in the timed region: binary_pair_stream256_kernel, one launch per step;
    # algorithmic bytes per launch: read a and b once, write one result per op
    per_launch_s = (ms * 1e-3) / args.steps
    multi_bytes_per_elem = (2 + len(OPS)) * ELEM_BYTES
    achieved = multi_bytes_per_elem * count / per_launch_s / 1e9
    separate_gbs = len(OPS) * BYTES_PER_ELEM * count / (ms_separate * 1e-3 / args.steps) / 1e9
    traffic, traffic_src = None, None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            tj = json.load(f)
        traffic = tj.get("binary_pair_stream_kernel_dram_bytes_per_launch")
        traffic_src = "NOT measured in this run: " + tj.get("source", "profiles/traffic.json") + \
                      ", per-element DRAM bytes of that capture scaled to this launch"
        if traffic is not None and tj.get("dram_bytes_per_element"):
            traffic = int(tj["dram_bytes_per_element"] * count)
    except Exception:
        pass

    # the timed outputs against the reference: every rank checks its whole shard, both ops
    for op, o in zip(OPS, outs):
        parity_check("C1", op, ELEM_BYTES, BITS, 0x5EED0001, 0x5EED0002, begin, count, o)
    chk = ops.checksum(outs[-1])
    del a, b, outs
    torch.cuda.empty_cache()

    # ---- the other BASELINE.json configs (same ranks, same timing rules) ----
    configs = []
    if not args.no_configs:
        UDT = {2: torch.uint16, 4: torch.uint32, 8: torch.uint64}
        # C0: the reference's CPU-sized case, 2^26 u16 <5,6,5> add_wrap per GPU (0.4 GB per pass).  The three
        # arrays (384 MiB) are near the L2's size, so the pass rotates over 8 buffer sets (3 GiB).
        L0 = pint_b200.Layout(2, 5, 6, 5)
        n0, sets = 1 << 26, 8
        b0, c0 = sharded.shard_range(n0 * world, 2, world, rank)
        bufs = []
        for k in range(sets):
            x, y = torch.empty(c0, dtype=torch.uint16, device=dev), torch.empty(c0, dtype=torch.uint16, device=dev)
            ops.fill_splitmix(x, 0xC0A + 2 * k, start=b0)
            ops.fill_splitmix(y, 0xC0B + 2 * k, start=b0)
            bufs.append((x, y, torch.empty_like(x)))
        ms0 = timed_launches(lambda i: ops.binary("add_wrap", L0, *bufs[i % sets][:2], out=bufs[i % sets][2]), 5 * sets)
        configs.append(config_line("C0", "add_wrap over 2^26 packed_int<uint16_t,5,6,5> per GPU, one launch per pass (programmatic dependent launch, the default), "
                                   "8 rotating buffer sets (L2 cannot hold a pass)", "weak", c0, n0 * world, 6, ms0,
                                   kernel="binary_stream256_kernel<swar_fn<add_wrap, static <5,6,5>>>"))
        # the same passes as plain launches (option pdl = 0): the default launches the stream kernels with programmatic
        # stream serialization, so the next pass's CTAs come onto the SMs while this one drains
        try:
            ops.set_option("pdl", 0)
            msp = timed_launches(lambda i: ops.binary("add_wrap", L0, *bufs[i % sets][:2], out=bufs[i % sets][2]), 5 * sets)
            configs.append(config_line("C0-no-pdl", "the same passes as plain launches (option pdl = 0)", "weak",
                                       c0, n0 * world, 6, msp, kernel="binary_stream256_kernel<swar_fn<add_wrap, static <5,6,5>>>"))
        finally:
            ops.set_option("pdl", 1)
        # the same 40 passes as ONE CUDA graph launch (launch latency and the ramp between passes amortised)
        try:
            g = torch.cuda.CUDAGraph()
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                with torch.cuda.graph(g, stream=side):
                    for i in range(5 * sets):
                        ops.binary("add_wrap", L0, *bufs[i % sets][:2], out=bufs[i % sets][2])
            torch.cuda.current_stream().wait_stream(side)
            msg = timed_launches(lambda i: g.replay(), 3, warm=2) / (5 * sets)
            configs.append(config_line("C0-graph", "the same 40 passes replayed as one CUDA graph", "weak", c0, n0 * world, 6, msg,
                                       kernel="binary_stream256_kernel<swar_fn<add_wrap, static <5,6,5>>>"))
            del g
        except Exception as e:  # noqa: BLE001 - a capture problem must not lose the rest of the line
            configs.append({"config": "C0-graph", "error": repr(e)[:200]})
        parity_check("C0", "add_wrap", 2, (5, 6, 5), 0xC0A, 0xC0B, b0, c0, bufs[0][2])
        del bufs
        torch.cuda.empty_cache()

        # C2: 2^31 u64 <16,16,16,16> (16 GiB per array) STRONG-scaled: the whole array is split over the ranks
        L2 = pint_b200.Layout(8, 16, 16, 16, 16)
        n2 = 1 << args.c2_log2
        free = torch.cuda.mem_get_info()[0]
        while world == 1 and 3 * 8 * n2 + (2 << 30) > free and n2 > (1 << 26):
            n2 >>= 1  # a smaller card than B200: say so in the line
        b2, c2 = sharded.shard_range(n2, 8, world, rank)
        x, y = torch.empty(c2, dtype=torch.uint64, device=dev), torch.empty(c2, dtype=torch.uint64, device=dev)
        o2 = torch.empty_like(x)
        ops.fill_splitmix(x, 0x5EED0001, start=b2)
        ops.fill_splitmix(y, 0x5EED0002, start=b2)
        for op in ("sub_signed_saturate", "add_signed_saturate"):
            ms2 = timed_launches(lambda i: ops.binary(op, L2, x, y, out=o2), 10)
            e = parity_check("C2", op, 8, (16, 16, 16, 16), 0x5EED0001, 0x5EED0002, b2, c2, o2)
            configs.append(config_line("C2", "%s over 2^%d packed_int<uint64_t,16,16,16,16> (%.0f GiB per array) split over %d GPU(s)"
                                       % (op, n2.bit_length() - 1, 8 * n2 / 2 ** 30, world), "strong", c2, n2, 24, ms2, op=op,
                                       parity_ok=e["parity_ok"], kernel="binary_stream_kernel<swar_fn<op, static 16x4>> (u64)"))
        del x, y, o2
        torch.cuda.empty_cache()

        # C3: irregular u64 <3,7,11,13,17,13>: get_signed of all six fields -> int32 planes, and repack
        bits3 = (3, 7, 11, 13, 17, 13)
        L3 = pint_b200.Layout(8, *bits3)
        n3 = 1 << 28
        b3, c3 = sharded.shard_range(n3 * world, 8, world, rank)
        x = torch.empty(c3, dtype=torch.uint64, device=dev)
        ops.fill_splitmix(x, 0xC3, start=b3)
        planes = [torch.empty(c3, dtype=torch.int32, device=dev) for _ in bits3]
        back = torch.empty_like(x)
        msu = timed_launches(lambda i: ops.unpack(L3, x, signed=True, out=planes), 10)
        msr = timed_launches(lambda i: ops.repack(L3, planes, out=back), 10)
        rt_ok = bool(torch.equal(back.view(torch.int64), x.view(torch.int64)))   # sum(Bits) = 64: repack(unpack(x)) == x
        rows = gather_u64([int(rt_ok)])
        rt_all = all(r[0] == 1 for r in rows)
        configs.append(config_line("C3-unpack", "get_signed of 6 fields of 2^28 packed_int<uint64_t,3,7,11,13,17,13> per GPU -> int32 planes",
                                   "weak", c3, n3 * world, 32, msu, roundtrip_ok=rt_all, kernel="unpack kernel (int32 planes)"))
        configs.append(config_line("C3-repack", "6 int32 planes -> packed (truncating field constructor)", "weak", c3, n3 * world, 32, msr,
                                   roundtrip_ok=rt_all, kernel="repack kernel (int32 planes)"))
        parity["checks"].append({"config": "C3", "op": "repack(unpack_signed(x)) == x on every rank", "parity_ok": rt_all})
        del x, planes, back
        torch.cuda.empty_cache()

        # C4: fused unpack -> add_signed_saturate -> repack, 2^30 irregular u64 per GPU (64 GiB of operand a at N = 8)
        n4 = 1 << 30
        b4, c4 = sharded.shard_range(n4 * world, 8, world, rank)
        x, y = torch.empty(c4, dtype=torch.uint64, device=dev), torch.empty(c4, dtype=torch.uint64, device=dev)
        o4 = torch.empty_like(x)
        ops.fill_splitmix(x, 0xC4A, start=b4)
        ops.fill_splitmix(y, 0xC4B, start=b4)
        ms4 = timed_launches(lambda i: ops.fused_unpack_arith_repack("add_signed_saturate", L3, x, y, out=o4), 10)
        e = parity_check("C4", "add_signed_saturate", 8, bits3, 0xC4A, 0xC4B, b4, c4, o4)
        configs.append(config_line("C4", "fused unpack -> add_signed_saturate -> repack over 2^30 packed_int<uint64_t,3,7,11,13,17,13> per GPU "
                                   "(%d GiB per operand whole job)" % (8 * world), "weak", c4, n4 * world, 24, ms4,
                                   parity_ok=e["parity_ok"], kernel="binary_stream_kernel<fused_fn<add_signed_saturate, static fields>>"))
        del x, y, o4
        torch.cuda.empty_cache()

    # ---- end-to-end leg: host buffers through pint_cuda_binary_multi_host ----
    e2e = None
    if not args.no_e2e:
        e2e_log2 = args.e2e_log2 if args.e2e_log2 is not None else (args.log2 if world == 1 else min(args.log2, 28))
        ne = 1 << e2e_log2
        nbytes = ne * ELEM_BYTES
        lib = _lib.lib()
        hp = [C.c_void_p() for _ in range(2 + len(OPS))]
        for p in hp:
            _lib.check(lib.pint_cuda_host_alloc(C.byref(p), nbytes), "pint_cuda_host_alloc")
        # fill the pinned inputs once (untimed) from the device generator
        tmp = torch.empty(ne, dtype=torch.uint32, device=dev)
        for p, seed in ((hp[0], 0xE2E0001), (hp[1], 0xE2E0002)):
            ops.fill_splitmix(tmp, seed, start=rank * ne)
            host_view = torch.frombuffer((C.c_uint8 * nbytes).from_address(p.value), dtype=torch.uint8)
            host_view.copy_(tmp.view(torch.uint8))
        torch.cuda.synchronize()
        del tmp

        codes = (C.c_int * len(OPS))(*[ops.OP_CODE[op] for op in OPS])
        houts = (C.c_void_p * len(OPS))(*[p.value for p in hp[2:]])

        def e2e_step():
            # both ops of the step on the same host inputs: one upload, two downloads
            _lib.check(lib.pint_cuda_binary_multi_host(len(OPS), codes, L.ref, hp[0], hp[1], houts, ne), "binary_multi_host")

        e2e_step()  # warm: allocates the staging ring
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.e2e_steps):
            e2e_step()  # synchronous: result is in host memory on return
        dt = max_over_ranks(time.perf_counter() - t0)
        first = list((C.c_uint32 * 4).from_address(hp[2].value))   # copied now: the ceiling leg below reuses the buffers
        # parity of what came back: host-side checksum of both result arrays against the reference's
        import numpy as np
        e2e_ok = None
        if checker:
            e2e_ok = True
            for j, op in enumerate(OPS):
                res = np.frombuffer((C.c_uint8 * nbytes).from_address(hp[2 + j].value), dtype=np.uint32)
                got = checker.port.checksum(res)
                want = checker.stream_checksum(op, ELEM_BYTES, BITS, 0xE2E0001, 0xE2E0002, rank * ne, ne)
                e2e_ok = e2e_ok and tuple(got) == tuple(want)
            e2e_ok = all(r[0] == 1 for r in gather_u64([int(e2e_ok)]))
            parity["checks"].append({"config": "e2e (host buffers)", "op": "+".join(OPS), "parity_ok": e2e_ok,
                                     "elements_checked_per_rank": ne})
        # the raw ceiling of this box for the same traffic: plain concurrent H2D (2 arrays) + D2H (2 arrays)
        # copies of the same pinned buffers on two streams, no kernel (cudaMemcpyAsync through torch's streams)
        ceiling = None
        try:
            from cuda.bindings import runtime as rt
            dbuf = [torch.empty(nbytes, dtype=torch.uint8, device=dev) for _ in range(4)]
            s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()
            kind_h2d, kind_d2h = rt.cudaMemcpyKind.cudaMemcpyHostToDevice, rt.cudaMemcpyKind.cudaMemcpyDeviceToHost

            def raw_step():
                for j in range(2):
                    err, = rt.cudaMemcpyAsync(dbuf[j].data_ptr(), hp[j].value, nbytes, kind_h2d, s_in.cuda_stream)
                    assert int(err) == 0, err
                for j in range(2):
                    err, = rt.cudaMemcpyAsync(hp[2 + j].value, dbuf[2 + j].data_ptr(), nbytes, kind_d2h, s_out.cuda_stream)
                    assert int(err) == 0, err
                s_in.synchronize()
                s_out.synchronize()
            raw_step()
            barrier()
            t1 = time.perf_counter()
            for _ in range(args.e2e_steps):
                raw_step()
            dtr = max_over_ranks(time.perf_counter() - t1)
            ceiling = 2 * nbytes * args.e2e_steps / dtr / 1e9   # GB/s per direction per GPU, both directions busy
            del dbuf
        except Exception as ex:  # noqa: BLE001
            ceiling = None
            print("e2e ceiling leg failed: %r" % (ex,), file=sys.stderr)
        achieved_dir = 2 * nbytes * args.e2e_steps / dt / 1e9
        e2e = {"value": len(OPS) * args.e2e_steps * ne * world / dt, "unit": "elements/s",
               "h2d_bytes_per_step": 2 * nbytes * world, "d2h_bytes_per_step": len(OPS) * nbytes * world,
               "elements_per_step": len(OPS) * ne * world, "elements_per_gpu": ne, "steps": args.e2e_steps,
               "api": "pint_cuda_binary_multi_host (pinned host in/out; a,b uploaded once per step, both results downloaded; chunked H2D/kernel/D2H ring)",
               "roofline": {"bound": "pcie (host<->device copies)", "achieved_gbs_per_direction_per_gpu": round(achieved_dir, 2),
                            "ceiling_gbs_per_direction_per_gpu": round(ceiling, 2) if ceiling else None,
                            "frac": round(achieved_dir / ceiling, 4) if ceiling else None,
                            "ceiling": "same pinned buffers, same bytes, plain cudaMemcpyAsync H2D and D2H on two streams "
                                       "concurrently on all %d rank(s), no kernel; measured in this run" % world},
               "parity_ok": e2e_ok, "result_head": [int(x) for x in first]}
        for p in hp:
            lib.pint_cuda_host_free(p)
        lib.pint_cuda_host_release()

    cpu, cpu1 = None, None
    if rank == 0 and not args.no_cpu:
        v1, kind1, _, best1 = cpu_reference_run(1 << max(args.cpu_log2 - 2, 20), reps=3, threads=1)
        cpu1 = {"value": v1, "unit": "elements/s", "cores": 1, "kind": kind1,
                "sample": f"{'+'.join(OPS)} over 2^{max(args.cpu_log2 - 2, 20)} elements, 1 thread, best of 3"}
        if world == 1:
            v, kind, cores, best = cpu_reference_run(1 << args.cpu_log2, reps=5)
            cpu = {"value": v, "unit": "elements/s", "cores": cores, "kind": kind, "single_thread_value": v1,
                   "sample": f"{'+'.join(OPS)} over 2^{args.cpu_log2} elements (best of 5, {best:.3f} s per pass pair), "
                             f"out[i]=op(a[i],b[i]) loop of tests/pint_bench.cpp:90-94 over SoA arrays"}
    barrier()

    checks = [c["parity_ok"] for c in parity["checks"] if c.get("parity_ok") is not None]
    parity_ok = all(checks) if checks else None
    rc = 0 if parity_ok in (True, None) else 1
    if rank == 0:
        c1 = [c for c in parity["checks"] if c["config"] == "C1"]
        line = {
            "metric": METRIC, "value": value, "unit": "elements/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u32", "data": "synthetic",
            "config": workload_config(args, world),
            "packed_elements_per_s": args.steps * n * world / (ms * 1e-3),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "traffic_source": traffic_src,
                         "kernel": "binary_pair_stream256_kernel<pair_fn<add_unsigned_saturate, add_signed_saturate, static RGBA8 masks>> (both ops, one pass, 256-bit accesses)",
                         "peak_source": peak_src, "frac_of_spec_8000": achieved / 8000.0,
                         "algorithmic_bytes_per_launch": multi_bytes_per_elem * count,
                         "algorithmic_bytes_per_element": multi_bytes_per_elem},
            "separate_passes": {"value": len(OPS) * args.steps * n * world / (ms_separate * 1e-3), "unit": "elements/s",
                                "ms_per_step": ms_separate / args.steps, "achieved_gbs": separate_gbs,
                                "frac": separate_gbs / peak,
                                "note": "same step as two pint_cuda_binary launches (12 B/element each)"},
            "configs": configs,
            "parity_ok": parity_ok, "parity": parity,
            "whole_job_checksum": c1[-1]["whole_job_checksum"] if c1 else None,
            "cpu_baseline": cpu, "cpu_single_thread": cpu1, "e2e": e2e, "gpu_launches": int(launches) * world, "clocks": clocks,
            "hbm_gbs_whole_job": achieved * world, "out_checksum": list(chk),
        }
        print(json.dumps(line), flush=True)
        if rc:
            print("PARITY MISMATCH: " + json.dumps([c for c in parity["checks"] if c.get("parity_ok") is False]), file=sys.stderr)
    if world > 1:
        dist.destroy_process_group()
    return rc


if __name__ == "__main__":
    sys.exit(main())
