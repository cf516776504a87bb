#!/usr/bin/env python
"""bench.py — fpemu cast throughput on B200 (BASELINE.json metric: cast GB/s & % of HBM peak).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--elems E] [--impl ours|reference]

A *step* is one pass of the hot path over one batch of synthetic input: the format x rounding sweep of
BASELINE.json configs[1] — {E5M2, E4M3, E3M4, BFLOAT16} x {RNE, STOCHASTIC, RTZ} out-of-place QDQ casts of one
2^30-element fp32 tensor per GPU (scale 1), plus one per-tensor amax-scaled E4M3_RNE cast (amax kernel ->
[NCCL all-reduce MAX of one float when N > 1] -> device-side scale -> cast).  Each rank owns
its own 2^30-element row slab (weak scaling); the only communication is that 4-byte all-reduce.

value = algorithmic bytes of all ranks / max-over-ranks device time (8 B/elem per plain cast: read 1 + write 1
fp32; 12 B/elem for the amax-scaled cast: amax read + cast read + write).  Inputs (4 GiB per tensor) are far
larger than the 126 MB L2, so no explicit flush is needed between iterations.

Outside the timed headline region the same run measures, under `detail` (device events, a few repetitions each):
  bf16      the same 12-mode sweep on a 2^30-element bfloat16 tensor (4 B/elem);
  inplace   the fp32 sweep in place;
  c5        BASELINE configs[4]: the Llama-7B weight set (291 tensors, 6.74 G elements) row-sharded over the N ranks,
            per-tensor amax-scaled E4M3_RNE through fpemu_b200.dist.ShardedPlan — STRONG scaling: total work is fixed,
            bf16 and fp32.  At N > 1 both exchanges of the maxima are timed alternately and must agree bit for bit: one
            NCCL all-reduce(MAX) between the amax and the cast launch, and the kernels' own exchange over NVLink peer
            memory (pushed behind the amax launch, awaited inside the cast launch); the faster one is `ms_per_sweep`.

e2e = the sweep through the host-buffer entry point fpemu_cast_host_multi: one pinned host tensor of 2^28 elements
in, thirteen pinned host tensors out (12 modes + the amax-scaled cast), copies inside the timed region.

`--impl reference` times the reference's own CPU op (oracle/_ref, built unmodified from
/root/reference/mpemu/pytquant/cpp/fpemu_impl.cpp) on the host cores on the same 2^28-element tensor: the modes of
the sweep its dispatcher has (9 of 12: upstream has no truncate mode for E4M3 / E3M4 / BFLOAT16) plus the amax-scaled
E4M3_RNE leg as the reference performs it (torch.max(abs(x)) + the op, mpemu/qutils.py:282-287).
"""
import argparse
import json
import os
import statistics
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "fp8-emulation-toolkit_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)

FORMATS = ["E5M2", "E4M3", "E3M4", "BFLOAT16"]
ROUNDINGS = ["RNE", "STOCHASTIC", "RTZ"]
SWEEP = ["%s_%s" % (f, r) for f in FORMATS for r in ROUNDINGS]
# the reference's dispatcher has no truncate mode for E4M3 / E3M4 / BFLOAT16 (fpemu_impl.cpp:1866-2009)
REF_SWEEP = [m for m in SWEEP if not (m.endswith("_RTZ") and not m.startswith("E5M2"))]
METRIC = "fpemu_cast_throughput"
UNIT = "GB/s"


def workload_config(n):
    """`config` of the JSON line — identical for both arms (the reference arm times a bounded sample of it)."""
    return {"workload": "format x rounding QDQ sweep {E5M2,E4M3,E3M4,BFLOAT16}x{RNE,STOCHASTIC,RTZ} + amax-scaled "
                        "E4M3_RNE on one 2^%d-element fp32 tensor per GPU, out-of-place (BASELINE configs[1])"
                        % (n.bit_length() - 1),
            "elems_per_gpu": n, "modes": SWEEP + ["E4M3_RNE_amax_scaled"],
            "l2": "inputs (%.1f GiB per tensor) larger than L2, no flush" % (4.0 * n / 2 ** 30),
            "parallelism": "row slabs, one rank per GPU, one 4-byte NCCL all-reduce(MAX) per step",
            "e2e_elems_per_tensor": 1 << 28}


def measured_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def committed_traffic(kernel_key):
    """dram__bytes_read+write per launch of the dominant kernel from the committed ncu --set full capture."""
    try:
        with open(os.path.join(ROOT, "profiles", "roofline_traffic.json")) as f:
            return json.load(f).get(kernel_key)
    except Exception:
        return None


class ClockSampler(object):
    """SM clock and throttle reasons sampled through NVML every 20 ms during the timed region (a thread of this
    process; the nvidia-smi CLI takes longer to start than a short timed region lasts)."""

    def __init__(self, index):
        self.index = index
        self.samples = []
        self.reasons = set()
        self.max_mhz = None
        self._stop = False
        self._thread = None
        self._h = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self._nv = pynvml
            uuid = None
            try:
                import torch
                uuid = "GPU-" + str(torch.cuda.get_device_properties(index).uuid)
            except Exception:
                pass
            self._h = None
            if uuid is not None:
                try:
                    self._h = pynvml.nvmlDeviceGetHandleByUUID(uuid)
                except Exception:
                    self._h = None
            if self._h is None:
                self._h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self._h = None

    def _loop(self):
        nv = self._nv
        names = {"hw_slowdown": getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8),
                 "hw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40),
                 "sw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20),
                 "sw_power_cap": getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4)}
        get_reasons = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or \
            getattr(nv, "nvmlDeviceGetCurrentClocksThrottleReasons", None)
        while not self._stop:
            try:
                self.samples.append(float(nv.nvmlDeviceGetClockInfo(self._h, nv.NVML_CLOCK_SM)))
                if get_reasons is not None:
                    mask = int(get_reasons(self._h))
                    for k, bit in names.items():
                        if mask & bit:
                            self.reasons.add(k)
            except Exception:
                pass
            time.sleep(0.02)

    def start(self):
        if self._h is None:
            return
        import threading
        self._thread = threading.Thread(target=self._loop, daemon=True)
        self._thread.start()

    def stop(self):
        if self._thread is None:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "samples": 0}
        self._stop = True
        self._thread.join(timeout=2)
        sm = self.samples
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(sm)}


# ---------------------------------------------------------------------------------------------- reference arm
def bind_thread_near_gpu(index):
    """Restrict the calling thread to the CPUs NVML reports as local to GPU `index` (its NUMA node).  Returns
    (previous affinity or None when nothing was changed, number of CPUs now allowed)."""
    if os.environ.get("FPEMU_BENCH_NO_AFFINITY"):
        return None, len(os.sched_getaffinity(0))
    try:
        import pynvml
        import torch
        pynvml.nvmlInit()
        try:
            h = pynvml.nvmlDeviceGetHandleByUUID("GPU-" + str(torch.cuda.get_device_properties(index).uuid))
        except Exception:
            h = pynvml.nvmlDeviceGetHandleByIndex(index)
        saved = os.sched_getaffinity(0)
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (max(saved) // 64) + 1)
        cpus = {64 * w + b for w, word in enumerate(words) for b in range(64) if (int(word) >> b) & 1} & saved
        if not cpus or cpus == saved:
            return None, len(saved)
        os.sched_setaffinity(0, cpus)
        return saved, len(cpus)
    except Exception:
        return None, len(os.sched_getaffinity(0))


def time_reference_cpu(n_elems, steps, warmup, modes, amax_leg=True):
    """Seconds per sweep of the reference CPU op over `modes` on an n_elems fp32 tensor, all host threads, plus
    (amax_leg) the per-tensor amax-scaled E4M3_RNE cast as the reference performs it on a CPU tensor:
    torch.max(abs(x)), scale = flt_max / vmax with its clamp, then the op (mpemu/qutils.py:282-287)."""
    import torch
    from oracle import ref_loader
    mod, label = ref_loader.best_cpu()
    if mod is None:
        return None, label, 0
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    g = torch.Generator().manual_seed(0)
    x = torch.randn(n_elems, dtype=torch.float32, generator=g)

    def sweep():
        for m in modes:
            mod.forward(x, m, n_elems, False, 1.0, False, 1)
        if amax_leg:
            vmax = torch.max(abs(torch.flatten(x.detach())))
            scale = 448.0 / vmax
            scale = 6.55e+04 if scale > 3.275e+04 else scale
            mod.forward(x, "E4M3_RNE", n_elems, False, float(scale), False, 1)
    for _ in range(warmup):
        sweep()
    t0 = time.perf_counter()
    for _ in range(steps):
        sweep()
    dt = (time.perf_counter() - t0) / steps
    return dt, label, cores


def ref_bytes(n):
    return 8.0 * n * len(REF_SWEEP) + 12.0 * n


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    n = args.ref_elems
    dt, label, cores = time_reference_cpu(n, args.steps, args.warmup, REF_SWEEP)
    if dt is None:
        print(json.dumps({"impl": "reference", "unavailable": label}))
        return 0
    gbs = ref_bytes(n) / dt / 1e9
    sample = ("reference CPU op (%s build of mpemu/pytquant/cpp/fpemu_impl.cpp), per step: the %d modes of the sweep its "
              "dispatcher has (%s) + the amax-scaled E4M3_RNE cast (torch.max(abs(x)) + op) of one 2^%d-element fp32 "
              "tensor, host memory only" % (label, len(REF_SWEEP), ",".join(REF_SWEEP), n.bit_length() - 1))
    line = {"impl": "reference", "metric": METRIC, "value": gbs, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args.elems),
            "cpu_baseline": {"value": gbs, "unit": UNIT, "cores": cores, "kind": "reference", "sample": sample},
            "e2e": {"value": gbs, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0,
                    "elems_per_tensor": n},
            "gpu_launches": 0}
    print(json.dumps(line))
    return 0


# ---------------------------------------------------------------------------------------------- our arm
def llama7b_shapes():
    """BASELINE configs[4] / SURVEY §8(d) C5: embed 32000x4096; 32 x {q,k,v,o 4096x4096; gate,up 11008x4096; down
    4096x11008; 2 norms 4096}; final norm; lm_head 32000x4096 = 291 tensors, 6.74 G elements."""
    shapes = [("embed_tokens", 32000, 4096)]
    for i in range(32):
        for nm in ("q_proj", "k_proj", "v_proj", "o_proj"):
            shapes.append(("layers.%d.%s" % (i, nm), 4096, 4096))
        shapes.append(("layers.%d.gate_proj" % i, 11008, 4096))
        shapes.append(("layers.%d.up_proj" % i, 11008, 4096))
        shapes.append(("layers.%d.down_proj" % i, 4096, 11008))
        shapes.append(("layers.%d.input_layernorm" % i, 4096, 1))
        shapes.append(("layers.%d.post_attention_layernorm" % i, 4096, 1))
    shapes.append(("norm", 4096, 1))
    shapes.append(("lm_head", 32000, 4096))
    return shapes


def c5_sweep(dev, world, rank, dtype_name, reps, pipeline, mode="E4M3_RNE", packed=False, multi=True, exchange="auto"):
    """One measurement of the C5 workload on this rank's row slabs; returns the result dict (same on every rank).
    exchange: how the float[291] vector of maxima is exchanged at N > 1 — "nccl" (one all-reduce between the two
    launches), "peer" (pushed over NVLink peer memory behind the amax launch, awaited inside the cast launch), "auto" = both are timed, checked to give
    identical scales and outputs, and the faster one is reported as `ms_per_sweep`."""
    import torch
    import torch.distributed as dist
    from fpemu_b200 import dist as fd
    dtype = {"bf16": torch.bfloat16, "fp32": torch.float32, "fp16": torch.float16}[dtype_name]
    esz = torch.empty(0, dtype=dtype).element_size()
    shapes = llama7b_shapes()
    slabs, offsets, cols, total = [], [], [], 0
    for t, (_, r, c) in enumerate(shapes):
        r0, r1 = fd.slab_bounds(r, world, rank)
        g = torch.Generator(device=dev).manual_seed(1000003 * t + rank)
        w = torch.randn((r1 - r0) * c, device=dev, dtype=torch.float32, generator=g) * 0.02
        slabs.append(w.to(dtype).view(r1 - r0, c))
        del w
        offsets.append(r0)
        cols.append(c)
        total += r * c
    outs = [torch.empty(x.shape, dtype=torch.uint8 if packed else dtype, device=dev) for x in slabs]
    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def measure(plan):
        def run():
            if packed:
                return plan.amax_scaled_encode(mode, seed=7)
            if plan is not None:
                return plan.amax_scaled_cast(mode, seed=7)
            return fd.sharded_amax_scaled_cast(slabs, mode, row_offsets=offsets, cols=cols, outs=outs, seed=7, pipeline=pipeline)
        for _ in range(3):
            res = run()
        barrier()
        a.record()
        for _ in range(reps):
            run()
        b.record()
        barrier()
        t = torch.tensor([a.elapsed_time(b) / reps], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        digest = torch.stack([o.view(torch.uint8)[:: max(1, o.numel() // 4096)].sum(dtype=torch.int64) for o in outs if o.numel()]).sum()
        return float(t.item()), res[1].clone(), int(digest.item())

    timings, plan = {}, None
    if multi or packed:
        kinds = ["nccl"] if (world == 1 or exchange == "nccl") else (["peer"] if exchange == "peer" else ["nccl", "peer"])
        plans, ref = {}, None
        for kind in kinds:
            try:
                plans[kind] = fd.ShardedPlan(slabs, row_offsets=offsets, cols=cols, outs=outs, pipeline=pipeline, exchange=kind)
            except Exception as e:  # noqa: BLE001  (symmetric memory unavailable on this box)
                timings[kind] = "unavailable: %s" % str(e)[:120]
        # the exchanges alternate, round after round, and each reports its best round: whichever ran first would
        # otherwise be measured on the cooler GPU (a B200 streaming at 7 TB/s loses ~5 % in its first second)
        for rnd in range(3 if len(plans) > 1 else 1):
            for kind in kinds:
                if kind not in plans:
                    continue
                ms_k, scales_k, digest_k = measure(plans[kind])
                timings[kind] = min(ms_k, timings.get(kind, ms_k))
                if ref is None:
                    ref = (scales_k, digest_k)
                else:
                    assert torch.equal(ref[0].view(torch.int32), scales_k.view(torch.int32)) and ref[1] == digest_k, \
                        "peer-memory exchange and NCCL all-reduce disagree"
        plan = plans.get(kinds[0])
        plans = None
        best = min((k for k in timings if isinstance(timings[k], float)), key=lambda k: timings[k])
        ms = timings[best]
    else:
        ms, _, _ = measure(None)
        best = "nccl"
    ar_us = 0.0
    if world > 1:                                     # the exchange step alone (float32[291], back to back)
        vec = torch.zeros(len(slabs), dtype=torch.float32, device=dev)
        for _ in range(5):
            dist.all_reduce(vec, op=dist.ReduceOp.MAX)
        torch.cuda.synchronize()
        a.record()
        for _ in range(50):
            dist.all_reduce(vec, op=dist.ReduceOp.MAX)
        b.record()
        torch.cuda.synchronize()
        ar_us = a.elapsed_time(b) / 50 * 1e3
    runs = 1 if world == 1 else pipeline
    gbs = ((2.0 * esz + 1.0) if packed else 3.0 * esz) * total / (ms * 1e-3) / 1e9
    res = {"workload": "Llama-7B weight set, per-tensor amax-scaled %s%s, row-sharded" % (mode, " -> 8-bit codes" if packed else ""),
           "dtype": dtype_name, "n_gpus": world, "tensors": len(shapes), "elements": total,
           "elements_per_rank": sum(x.numel() for x in slabs), "ms_per_sweep": ms,
           "aggregate_GBps": gbs, "per_gpu_GBps": gbs / world, "bytes_per_element": (2 * esz + 1) if packed else 3 * esz,
           "allreduce_us": ar_us, "pipeline_runs": runs,
           "exchange": best if world > 1 else "none (one rank)",
           "ms_per_sweep_by_exchange": timings if world > 1 else None,
           "launches_per_rank_per_sweep": (3 if (world > 1 and best == "peer") else 2 * runs + 1) if (multi or packed)
           else 3 * len(shapes) + 1,
           "scaling": "strong"}
    del slabs, outs
    torch.cuda.empty_cache()
    return res


def check_non_current_device(F, local, ndev):
    """Entry points on a GPU that is not the caller's current device (only possible on a multi-GPU box, so the bench
    does it where the driver's scaling run can see it): same bits as on the current device, current device unchanged."""
    import torch
    other = (local + 1) % ndev
    if other == local:
        return "skipped (one GPU)"
    g = torch.Generator().manual_seed(5)
    h = torch.randn(300005, generator=g)
    h2 = torch.randn(96, 1024, generator=g)
    res = []
    for d in (local, other):
        x, x2 = h.to("cuda:%d" % d), h2.to("cuda:%d" % d)
        r = [F.cast(x, "E4M3_RNE"), F.cast(x, "E5M2_STOCHASTIC", seed=9), F.cast_amax_scaled(x, "E4M3_RNE"),
             F.cast_amax_scaled(x, "E4M3_RNE", fused=False), F.cast_amax_scaled_noalloc(x, "E3M4_RNE", torch.empty_like(x)),
             F.cast_rowwise(x2, "E4M3_RNE"), F.decode(F.encode(x, "E5M2_RNE"), "E5M2"), F.amax(x)]
        plan = F.MultiTensorPlan([x, x2.reshape(-1)], [torch.empty_like(x), torch.empty_like(x2.reshape(-1))])
        r += list(plan.amax_scaled_cast("E4M3_RNE")[0])
        assert all(t.device.index == d for t in r)
        torch.cuda.synchronize(d)
        res.append([t.cpu() for t in r])
        del x, x2, r, plan
    assert torch.cuda.current_device() == local, "an entry point changed the caller's current device"
    for a, b in zip(*res):
        assert torch.equal(a.view(torch.uint8), b.view(torch.uint8)), "results differ on a non-current device"
    return "ok (cuda:%d while cuda:%d is current: %d results bit-identical)" % (other, local, len(res[0]))


def host_mem_available_gib():
    try:
        with open("/proc/meminfo") as f:
            for line in f:
                if line.startswith("MemAvailable"):
                    return int(line.split()[1]) / 2 ** 20
    except Exception:
        pass
    return None


def run_ours(args):
    import torch
    import torch.distributed as dist
    import fpemu_b200 as F
    import fpemu_cuda  # noqa: F401  (the reference-facing module must import)

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    # keep stdout to the ONE JSON line: libraries (NCCL prints its version banner with printf) write to fd 1, so
    # fd 1 is pointed at stderr for the whole run and the JSON line goes to a private copy of the real stdout
    sys.stdout.flush()
    real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    F._lib.lib()                                                   # fail loudly if the CUDA library is missing

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    other_device = None
    if world > 1 and not args.no_detail:
        other_device = check_non_current_device(F, local, torch.cuda.device_count())
        barrier()

    n = args.elems
    g = torch.Generator(device=dev).manual_seed(1234 + rank)
    x = torch.randn(n, dtype=torch.float32, device=dev, generator=g)
    y = torch.empty_like(x)
    d_amax = torch.zeros(1, dtype=torch.float32, device=dev)
    slab_offset = rank * n                                         # this rank's row slab of the world x n tensor

    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(len(SWEEP) + 3)] for _ in range(args.steps)]

    def step(k=None):
        # per-tensor amax for the scaled cast first; its 4-byte all-reduce (N > 1) then overlaps the plain casts
        if k is not None:
            ev[k][0].record()
        d_amax.zero_()
        F.amax(x, out=d_amax)
        work = dist.all_reduce(d_amax, op=dist.ReduceOp.MAX, async_op=True) if world > 1 else None
        for i, mode in enumerate(SWEEP):
            if k is not None:
                ev[k][i + 1].record()
            F.cast(x, mode, out=y, seed=1000 + i, elem_offset=slab_offset)
        if k is not None:
            ev[k][len(SWEEP) + 1].record()
        if work is not None:
            work.wait()
        F.cast_amax_scaled(x, "E4M3_RNE", out=y, d_amax=d_amax)
        if k is not None:
            ev[k][len(SWEEP) + 2].record()

    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    t_start = torch.cuda.Event(enable_timing=True)
    t_stop = torch.cuda.Event(enable_timing=True)
    t_start.record()
    for k in range(args.steps):
        step(k)
    t_stop.record()
    barrier()
    clk = clocks.stop() if rank == 0 else None
    elapsed_ms = t_start.elapsed_time(t_stop)
    t = torch.tensor([elapsed_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    elapsed_ms = float(t.item())

    bytes_per_step_rank = 8.0 * n * len(SWEEP) + 12.0 * n
    value = bytes_per_step_rank * world * args.steps / (elapsed_ms * 1e-3) / 1e9

    # per-launch durations (device events inside the timed region, this rank)
    per_mode = {}
    for i, mode in enumerate(SWEEP):
        ms = [ev[k][i + 1].elapsed_time(ev[k][i + 2]) for k in range(args.steps)]
        per_mode[mode] = sum(ms) / len(ms)
    amax_ms = sum(ev[k][0].elapsed_time(ev[k][1]) for k in range(args.steps)) / args.steps
    scaled_cast_ms = sum(ev[k][len(SWEEP) + 1].elapsed_time(ev[k][len(SWEEP) + 2]) for k in range(args.steps)) / args.steps
    amax_scaled_ms = amax_ms + scaled_cast_ms

    # ---- outside the headline region: bf16 and in-place sweeps of configs[1], and configs[4] (C5, strong scaling)
    extra = {}
    if not args.no_detail:
        reps = max(2, min(args.steps, 5))

        def timed_sweep(src, dst, esz):
            evs = [[torch.cuda.Event(enable_timing=True) for _ in range(len(SWEEP) + 1)] for _ in range(reps)]
            for r in range(-2, reps):
                for i, mode in enumerate(SWEEP):
                    if r >= 0:
                        evs[r][i].record()
                    F.cast(src, mode, out=dst, seed=1000 + i, elem_offset=slab_offset)
                if r >= 0:
                    evs[r][len(SWEEP)].record()
            torch.cuda.synchronize()
            out = {}
            for i, mode in enumerate(SWEEP):
                ms = sum(evs[r][i].elapsed_time(evs[r][i + 1]) for r in range(reps)) / reps
                out[mode] = {"ms": round(ms, 4), "GB/s": round(2.0 * esz * src.numel() / (ms * 1e-3) / 1e9, 1)}
            return out
        y.copy_(x)
        extra["inplace_f32"] = timed_sweep(y, y, 4)
        xb = x.to(torch.bfloat16)
        yb = torch.empty_like(xb)
        extra["bf16"] = timed_sweep(xb, yb, 2)
        extra["bf16_elems_per_gpu"] = xb.numel()
        del xb, yb
    del x, y
    torch.cuda.empty_cache()
    if not args.no_detail:
        barrier()
        extra["c5"] = {dt: c5_sweep(dev, world, rank, dt, max(3, min(args.steps, 10)), args.pipeline) for dt in ("bf16", "fp32")}
        barrier()

    # ---- end to end: host buffers in, host buffers out, copies inside the timed region
    e2e = None
    if not args.no_e2e:
        ne = args.e2e_elems
        # pinned buffers are placed by first touch: run this thread on the CPUs next to this rank's GPU while they
        # are allocated and used, so that on a multi-socket box the DMA traffic does not cross the socket link
        saved_affinity, near = bind_thread_near_gpu(local)
        njobs = len(SWEEP) + 1
        avail = host_mem_available_gib()
        need = 4.0 * ne * (njobs + 1) / 2 ** 30 * world
        distinct = njobs
        if avail is not None and need > 0.5 * avail:              # small host: reuse output buffers cyclically, and say so
            distinct = max(1, int(0.5 * avail / (4.0 * ne / 2 ** 30 * world)) - 1)
        h_in = torch.randn(ne, dtype=torch.float32).pin_memory()
        h_outs = [torch.empty(ne, dtype=torch.float32).pin_memory() for _ in range(min(distinct, njobs))]
        caster = F.HostCaster(torch.float32, chunk_elems=1 << 24, device=local)
        jobs = [(h_outs[i % len(h_outs)], mode, 1.0, 1000 + i) for i, mode in enumerate(SWEEP)]
        jobs.append((h_outs[len(SWEEP) % len(h_outs)], "E4M3_RNE", "amax", 0))

        def e2e_step():
            # one call: the tensor is uploaded once, cast 13 times (the amax-scaled cast in a second pass over the
            # device-resident copy) and every result is copied back to its own host buffer; returns synchronised.
            # N > 1: each rank casts its own slab with its own maximum (the global-maximum exchange is timed in `value`).
            caster.run_multi(h_in, jobs)
        e2e_step()
        barrier()
        e2e_steps = max(1, min(args.steps, args.e2e_steps))
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            e2e_step()
        barrier()
        dt = time.perf_counter() - t0
        tt = torch.tensor([dt], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dt = float(tt.item())
        e2e_bytes = (8.0 * ne * len(SWEEP) + 12.0 * ne) * world * e2e_steps
        e2e = {"value": e2e_bytes / dt / 1e9, "unit": UNIT,
               "h2d_bytes_per_step": int(4 * ne), "d2h_bytes_per_step": int(4 * ne * njobs),
               "elems_per_tensor": ne, "steps": e2e_steps, "distinct_host_outputs": len(h_outs),
               "api": "fpemu_cast_host_multi (C ABI: one pinned host tensor in, 13 pinned host tensors out; one upload into "
                      "a device-resident copy, 2 x 64 MiB output staging, 3 streams)",
               "host_cpus_near_gpu": near}
        del h_in, h_outs, caster, jobs
        if saved_affinity is not None:
            os.sched_setaffinity(0, saved_affinity)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    peak, peak_src = measured_peak()
    dominant = max(per_mode, key=lambda m: per_mode[m])
    dom_ms = per_mode[dominant]
    achieved = 8.0 * n / (dom_ms * 1e-3) / 1e9
    detail = {m: {"ms": round(per_mode[m], 4), "GB/s": round(8.0 * n / (per_mode[m] * 1e-3) / 1e9, 1),
                  "Gelem/s": round(n / (per_mode[m] * 1e-3) / 1e9, 1),
                  "frac_of_peak": round(8.0 * n / (per_mode[m] * 1e-3) / 1e9 / peak, 4)} for m in SWEEP}
    detail["E4M3_RNE_amax_scaled"] = {"ms": round(amax_scaled_ms, 4),
                                      "GB/s": round(12.0 * n / (amax_scaled_ms * 1e-3) / 1e9, 1),
                                      "Gelem/s": round(n / (amax_scaled_ms * 1e-3) / 1e9, 1),
                                      "frac_of_peak": round(12.0 * n / (amax_scaled_ms * 1e-3) / 1e9 / peak, 4)}
    detail.update(extra)
    if other_device is not None:
        detail["non_current_device_check"] = other_device

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": elapsed_ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(n),
            "roofline": {"bound": "hbm", "kernel": "cast_vec_kernel<%s>" % dominant, "achieved": achieved, "peak": peak,
                         "unit": "GB/s", "frac": achieved / peak, "traffic": committed_traffic(dominant),
                         "peak_source": peak_src, "algorithmic_bytes_per_launch": 8 * n,
                         "avg_launch_ms": dom_ms, "frac_of_nominal_8000": achieved / 8000.0},
            "detail": detail,
            "clocks": clk,
            "gpu_launches": args.steps * (len(SWEEP) + 3)}
    if e2e is not None:
        line["e2e"] = e2e

    if world == 1 and not args.no_cpu:
        dt, label, cores = time_reference_cpu(args.ref_elems, 2, 1, REF_SWEEP)
        if dt is not None:
            line["cpu_baseline"] = {
                "value": ref_bytes(args.ref_elems) / dt / 1e9, "unit": UNIT, "cores": cores, "kind": "reference",
                "sample": "reference CPU op (%s build), %d-mode sweep + amax-scaled E4M3_RNE of one 2^%d-element fp32 "
                          "tensor, 2 timed passes" % (label, len(REF_SWEEP), args.ref_elems.bit_length() - 1)}
        else:
            line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": 0, "kind": "reference", "sample": label}
    real_stdout.write(json.dumps(line) + "\n")
    real_stdout.flush()
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--elems", type=int, default=1 << 30, help="elements per GPU per tensor")
    ap.add_argument("--e2e-elems", type=int, default=1 << 28)
    ap.add_argument("--e2e-steps", type=int, default=2)
    ap.add_argument("--ref-elems", type=int, default=1 << 28, help="CPU tensor size (reference arm / cpu_baseline) = the e2e size")
    ap.add_argument("--pipeline", type=int, default=1, help="runs of tensors per C5 sweep (1 measured fastest at 8 GPUs: profiles/llama7b_r2_n8.jsonl)")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-detail", action="store_true", help="skip the bf16 / in-place / C5 legs")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
