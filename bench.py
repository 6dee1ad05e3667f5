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

  value     packed elements/s, whole job, inputs resident in HBM (CUDA events, max over ranks)
  e2e       same metric through the C ABI's host-buffer entry point
            (pint_cuda_binary_host): pinned host arrays in, host array out, the
            H2D/D2H copies inside the timed region
  roofline  the streaming kernel against the measured HBM copy bandwidth
  cpu_baseline  the reference's own scalar functions in the tests/pint_bench.cpp
            loop shape on this box's host cores (oracle/_ref), rank 0, N=1 only

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
    times = []
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
        "config": workload_config(args, 1, sample_log2=args.cpu_log2),
        "cpu_baseline": {"value": value, "unit": "elements/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": "elements/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


def workload_config(args, world, sample_log2=None):
    cfg = {
        "workload": "C1: add_unsigned_saturate + add_signed_saturate over 2^%d packed_int<uint32_t,8,8,8,8> per GPU, "
                    "both ops per step (one pint_cuda_binary_multi launch: a,b read once, two outputs)" % args.log2,
        "layout": "packed_int<uint32_t,8,8,8,8>", "elements_per_gpu": 1 << args.log2,
        "elements_per_step": len(OPS) * (1 << args.log2) * world,
        "bytes_per_element": (2 + len(OPS)) * ELEM_BYTES, "bytes_per_element_single_op_pass": BYTES_PER_ELEM,
        "parallelism": "%d independent contiguous shard(s), no collective" % world,
        "l2": "inputs (2 x %.1f GiB per GPU) far larger than the 126 MB L2; no flush needed" % ((1 << args.log2) * 4 / 2 ** 30),
    }
    if sample_log2 is not None:
        cfg["cpu_sample_elements"] = 1 << sample_log2
    return cfg


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--log2", type=int, default=30, help="log2 of packed elements per GPU")
    ap.add_argument("--e2e-log2", type=int, default=28, help="log2 of elements per GPU in the host-buffer (e2e) leg")
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--cpu-log2", type=int, default=27, help="log2 of the CPU baseline sample")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
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

    L = pint_b200.Layout(ELEM_BYTES, *BITS)
    assert L.is_specialized(), "RGBA8 must run the compiled-in-mask kernels"
    n = 1 << args.log2
    # this rank's shard of the global array of world * n elements
    begin, count = sharded.shard_range(n * world, ELEM_BYTES, world, rank)
    a = torch.empty(count, dtype=torch.uint32, device=dev)
    b = torch.empty(count, dtype=torch.uint32, device=dev)
    outs = [torch.empty(count, dtype=torch.uint32, device=dev) for _ in OPS]
    out = outs[-1]
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
    peak, peak_src = measured_peak()
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            traffic = json.load(f).get("binary_pair_stream_kernel_dram_bytes_per_launch")
    except Exception:
        pass

    # sanity: the output of the last pass is the signed-saturated sum (checked against a
    # device-side identity, not the oracle: a + b with both saturations agree where no field overflows)
    chk = ops.checksum(out)

    # ---- end-to-end leg: host buffers through pint_cuda_binary_host ----
    e2e = None
    if not args.no_e2e:
        ne = 1 << args.e2e_log2
        nbytes = ne * ELEM_BYTES
        hp = [C.c_void_p() for _ in range(2 + len(OPS))]
        for p in hp:
            _lib.check(_lib.lib().pint_cuda_host_alloc(C.byref(p), nbytes), "pint_cuda_host_alloc")
        # fill the pinned inputs once (untimed) from the device generator
        tmp = torch.empty(ne, dtype=torch.uint32, device=dev)
        for p, seed in ((hp[0], 0xE2E0001), (hp[1], 0xE2E0002)):
            ops.fill_splitmix(tmp, seed)
            host_view = torch.frombuffer((C.c_uint8 * nbytes).from_address(p.value), dtype=torch.uint8)
            host_view.copy_(tmp.view(torch.uint8))
        torch.cuda.synchronize()
        del tmp
        lib = _lib.lib()

        codes = (C.c_int * len(OPS))(*[ops.OP_CODE[op] for op in OPS])
        outs = (C.c_void_p * len(OPS))(*[p.value for p in hp[2:]])

        def e2e_step():
            # both ops of the step on the same host inputs: one upload, two downloads
            _lib.check(lib.pint_cuda_binary_multi_host(len(OPS), codes, L.ref, hp[0], hp[1], outs, ne), "binary_multi_host")

        e2e_step()  # warm: allocates the staging ring
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.e2e_steps):
            e2e_step()  # synchronous: result is in host memory on return
        dt = max_over_ranks(time.perf_counter() - t0)
        first = (C.c_uint32 * 4).from_address(hp[2].value)
        e2e = {"value": len(OPS) * args.e2e_steps * ne * world / dt, "unit": "elements/s",
               "h2d_bytes_per_step": 2 * nbytes * world, "d2h_bytes_per_step": len(OPS) * nbytes * world,
               "elements_per_step": len(OPS) * ne * world, "steps": args.e2e_steps,
               "api": "pint_cuda_binary_multi_host (pinned host in/out; a,b uploaded once per step, both results downloaded; chunked H2D/kernel/D2H ring)",
               "result_head": [int(x) for x in first]}
        for p in hp:
            lib.pint_cuda_host_free(p)

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        v, kind, cores, best = cpu_reference_run(1 << args.cpu_log2, reps=5)
        v1, _, _, best1 = cpu_reference_run(1 << max(args.cpu_log2 - 2, 20), reps=3, threads=1)
        cpu = {"value": v, "unit": "elements/s", "cores": cores, "kind": kind, "single_thread_value": v1,
               "sample": f"{'+'.join(OPS)} over 2^{args.cpu_log2} elements (best of 5, {best:.3f} s per pass pair), "
                         f"out[i]=op(a[i],b[i]) loop of tests/pint_bench.cpp:90-94 over SoA arrays"}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": "elements/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u32", "data": "synthetic",
            "config": workload_config(args, world),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "kernel": "binary_pair_stream256_kernel<pair_fn<add_unsigned_saturate, add_signed_saturate, static RGBA8 masks>> (both ops, one pass, 256-bit accesses)",
                         "peak_source": peak_src, "frac_of_spec_8000": achieved / 8000.0,
                         "algorithmic_bytes_per_launch": multi_bytes_per_elem * count,
                         "algorithmic_bytes_per_element": multi_bytes_per_elem},
            "separate_passes": {"value": len(OPS) * args.steps * n * world / (ms_separate * 1e-3), "unit": "elements/s",
                                "ms_per_step": ms_separate / args.steps, "achieved_gbs": separate_gbs,
                                "frac": separate_gbs / peak,
                                "note": "same step as two pint_cuda_binary launches (12 B/element each)"},
            "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches) * world, "clocks": clocks,
            "hbm_gbs_whole_job": achieved * world, "out_checksum": list(chk),
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
