#!/usr/bin/env python
"""Format x rounding x dtype x size sweep of the cast on one B200 (BASELINE.json configs[0..1]) and, next to
it, the reference's own CUDA kernels (oracle/_ref/fpemu_cuda_ref.so, unmodified, compiled for sm_100a) and CPU op.

    python tools/sweep.py [--quick] [--out gpurun_out/sweep.json]

Timing: median of `reps` CUDA-event-timed launches after 3 warm-ups.  Tensors smaller than 2x L2 are rotated
through enough distinct buffers that the working set exceeds 256 MiB, so every number is an HBM number."""
import argparse
import json
import os
import statistics
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "fp8-emulation-toolkit_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

import fpemu_b200 as F  # noqa: E402

FORMATS = ["E5M2", "E4M3", "E3M4", "BFLOAT16"]
ROUNDINGS = ["RNE", "STOCHASTIC", "RTZ"]
DTYPES = {"fp32": torch.float32, "bf16": torch.bfloat16, "fp16": torch.float16}
L2_BYTES = 126 << 20


def time_launches(fn, nbuf, reps):
    for i in range(3):
        fn(i % nbuf)
    torch.cuda.synchronize()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    for i, (a, b) in enumerate(evs):
        a.record()
        fn(i % nbuf)
        b.record()
    torch.cuda.synchronize()
    return statistics.median(a.elapsed_time(b) for a, b in evs)


def buffers(n, dtype, inplace):
    esz = torch.empty(0, dtype=dtype).element_size()
    per = n * esz * (1 if inplace else 2)
    nbuf = max(1, min(64, -(-2 * L2_BYTES // per)))
    g = torch.Generator(device="cuda").manual_seed(0)
    xs = [torch.randn(n, device="cuda", dtype=torch.float32, generator=g).to(dtype) for _ in range(nbuf)]
    ys = xs if inplace else [torch.empty_like(x) for x in xs]
    return xs, ys, nbuf


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "sweep.json"))
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--odd", action="store_true", help="only the odd-layout rows (misaligned views, odd element offsets)")
    args = ap.parse_args()
    sizes = [20, 24, 28, 30] if args.quick else [20, 22, 24, 26, 28, 30, 31, 32]
    rows = []
    peak = 6545.0
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        pass

    if args.odd:
        # odd layouts only: slabs whose element offset is not a multiple of 16 (stochastic), views off a 32-byte
        # boundary with input and output off by the same / by different amounts, a tensor list with odd offsets
        n = 1 << 28
        for dname, dtype in (("fp32", torch.float32), ("bf16", torch.bfloat16)):
            esz = torch.empty(0, dtype=dtype).element_size()
            x = torch.randn(n + 64, device="cuda").to(dtype)
            y = torch.empty(n + 64, dtype=dtype, device="cuda")
            cases = [("aligned, elem_offset 0", 0, 0, 0), ("aligned, elem_offset 3", 0, 0, 3), ("aligned, elem_offset 8", 0, 0, 8),
                     ("in and out off by 1 element", 1, 1, 0), ("in off by 1, out aligned", 1, 0, 0),
                     ("in aligned, out off by 3", 0, 3, 0), ("in off by 5, out off by 2, elem_offset 3", 5, 2, 3)]
            for mode in ("E4M3_RNE", "E4M3_STOCHASTIC", "BFLOAT16_STOCHASTIC"):
                for label, io, oo, eo in cases:
                    xi, yo = x[io:io + n], y[oo:oo + n]
                    ms = time_launches(lambda i: F.cast(xi, mode, out=yo, seed=7, elem_offset=eo), 1, args.reps)
                    rows.append({"impl": "b200", "mode": mode, "dtype": dname, "log2n": 28, "layout": label, "ms": ms,
                                 "GB/s": 2.0 * n * esz / (ms * 1e-3) / 1e9})
            parts = [x[k * (n // 8):(k + 1) * (n // 8)] for k in range(8)]
            outs = [y[k * (n // 8):(k + 1) * (n // 8)] for k in range(8)]
            for label, offs in (("list of 8 slabs, offsets multiple of 16", [16 * k for k in range(8)]),
                                ("list of 8 slabs, odd offsets", [16 * k + 3 for k in range(8)])):
                plan = F.MultiTensorPlan(parts, outs, elem_offsets=offs)
                ms = time_launches(lambda i: plan.cast("E4M3_STOCHASTIC", seed=7), 1, args.reps)
                rows.append({"impl": "b200", "mode": "E4M3_STOCHASTIC (fpemu_cast_multi)", "dtype": dname, "log2n": 28,
                             "layout": label, "ms": ms, "GB/s": 2.0 * n * esz / (ms * 1e-3) / 1e9})
            del x, y, parts, outs
            torch.cuda.empty_cache()
        os.makedirs(os.path.dirname(args.out), exist_ok=True)
        with open(args.out, "w") as f:
            json.dump({"gpu": torch.cuda.get_device_name(0), "peak_measured_GBs": peak, "rows": rows}, f, indent=0)
        for r in rows:
            print("%-36s %-5s %-44s %8.1f GB/s" % (r["mode"], r["dtype"], r["layout"], r["GB/s"]))
        return

    # ---- ours
    for dname, dtype in DTYPES.items():
        esz = torch.empty(0, dtype=dtype).element_size()
        for lg in sizes:
            n = 1 << lg
            if n * esz * 2 > 72 << 30:
                continue
            for inplace in (False, True):
                if inplace and lg not in (24, 30):
                    continue
                xs, ys, nbuf = buffers(n, dtype, inplace)
                for fmt in FORMATS:
                    for rnd in ROUNDINGS:
                        for scale in ((1.0, 3.7) if (fmt != "BFLOAT16" and rnd == "RNE" and lg in (24, 30) and not inplace) else (1.0,)):
                            mode = "%s_%s" % (fmt, rnd)
                            ms = time_launches(lambda i: F.cast(xs[i], mode, out=ys[i], seed=7, scale=scale), nbuf, args.reps)
                            gbs = 2.0 * n * esz / (ms * 1e-3) / 1e9
                            rows.append({"impl": "b200", "mode": mode, "dtype": dname, "log2n": lg, "inplace": inplace,
                                         "scale": scale, "ms": ms, "GB/s": gbs, "Gelem/s": n / (ms * 1e-3) / 1e9,
                                         "frac_measured_peak": gbs / peak, "frac_nominal_8000": gbs / 8000.0})
                del xs, ys
                torch.cuda.empty_cache()
        # amax-scaled per-tensor cast (12 B / 6 B per element)
        for lg in (24, 30):
            n = 1 << lg
            xs, ys, nbuf = buffers(n, dtype, False)
            ms = time_launches(lambda i: F.cast_amax_scaled(xs[i], "E4M3_RNE", out=ys[i]), nbuf, args.reps)
            rows.append({"impl": "b200", "mode": "E4M3_RNE amax-scaled", "dtype": dname, "log2n": lg, "inplace": False, "scale": "amax",
                         "ms": ms, "GB/s": 3.0 * n * esz / (ms * 1e-3) / 1e9, "Gelem/s": n / (ms * 1e-3) / 1e9,
                         "frac_measured_peak": 3.0 * n * esz / (ms * 1e-3) / 1e9 / peak})
            ms = time_launches(lambda i: F.amax(xs[i]), nbuf, args.reps)
            rows.append({"impl": "b200", "mode": "amax only", "dtype": dname, "log2n": lg, "inplace": False, "scale": "-",
                         "ms": ms, "GB/s": 1.0 * n * esz / (ms * 1e-3) / 1e9, "Gelem/s": n / (ms * 1e-3) / 1e9,
                         "frac_measured_peak": 1.0 * n * esz / (ms * 1e-3) / 1e9 / peak})
            del xs, ys
            torch.cuda.empty_cache()

    # ---- row-wise and blocked variants (fp32)
    for rows_, cols in ((4096, 4096), (11008, 4096), (32000, 4096), (64, 576), (2048, 4608)):
        x = torch.randn(rows_, cols, device="cuda")
        y = torch.empty_like(x)
        ms = time_launches(lambda i: F.cast_rowwise(x, "E4M3_RNE", out=y), 1, args.reps)
        rows.append({"impl": "b200", "mode": "E4M3_RNE row-wise %dx%d" % (rows_, cols), "dtype": "fp32", "log2n": None, "inplace": False,
                     "scale": "per-row amax", "ms": ms, "GB/s": 8.0 * rows_ * cols / (ms * 1e-3) / 1e9})
    x = torch.randn(1 << 28, device="cuda")
    y = torch.empty_like(x)
    for bs in (32, 64, 256, 768, 1024):
        nb = ((1 << 28) // bs) * bs
        ms = time_launches(lambda i: F.cast_blocked(x, "E4M3_RNE", bs, out=y, n=nb), 1, args.reps)
        rows.append({"impl": "b200", "mode": "E4M3_RNE blocked bs=%d" % bs, "dtype": "fp32", "log2n": 28, "inplace": False,
                     "scale": "per-block", "ms": ms, "GB/s": 8.0 * nb / (ms * 1e-3) / 1e9})
    del x, y
    torch.cuda.empty_cache()

    # ---- the reference's own CUDA kernels on this GPU (the kernel to beat)
    try:
        from oracle import ref_loader
        ref = ref_loader.load("fpemu_cuda_ref") if ref_loader.available("fpemu_cuda_ref") else None
    except Exception:
        ref = None
    if ref is not None:
        for dname, dtype in (("fp32", torch.float32), ("fp16", torch.float16)):
            esz = 4 if dtype == torch.float32 else 2
            for lg in (24, 28, 30):
                n = 1 << lg
                x = torch.randn(n, device="cuda").to(dtype)
                for mode in ("E5M2_RNE", "E5M2_STOCHASTIC", "E4M3_RNE", "E4M3_STOCHASTIC", "E3M4_RNE", "BFLOAT16_RNE"):
                    for inplace in (False, True):
                        xx = x.clone() if inplace else x
                        torch.cuda.synchronize()
                        ts = []
                        for _ in range(2 + min(args.reps, 7)):
                            t0 = time.perf_counter()
                            ref.forward(xx, mode, n, inplace, 1.0, False, 1)      # synchronises the device itself
                            ts.append((time.perf_counter() - t0) * 1e3)
                        ms = statistics.median(ts[2:])
                        rows.append({"impl": "reference_cuda", "mode": mode, "dtype": dname, "log2n": lg, "inplace": inplace,
                                     "scale": 1.0, "ms": ms, "GB/s": 2.0 * n * esz / (ms * 1e-3) / 1e9,
                                     "Gelem/s": n / (ms * 1e-3) / 1e9})
                del x
                torch.cuda.empty_cache()

    # ---- reference CPU op on this host
    try:
        from oracle import ref_loader
        cpu, label = ref_loader.best_cpu()
    except Exception:
        cpu, label = None, "unavailable"
    if cpu is not None:
        torch.set_num_threads(os.cpu_count() or 1)
        n = 1 << 24
        x = torch.randn(n, generator=torch.Generator().manual_seed(0))
        for mode in ("E5M2_RNE", "E5M2_STOCHASTIC", "E5M2_RTZ", "E4M3_RNE", "E4M3_STOCHASTIC", "E3M4_RNE", "E3M4_STOCHASTIC",
                     "BFLOAT16_RNE", "BFLOAT16_STOCHASTIC"):
            ts = []
            for _ in range(5):
                t0 = time.perf_counter()
                cpu.forward(x, mode, n, False, 1.0, False, 1)
                ts.append((time.perf_counter() - t0) * 1e3)
            ms = min(ts[1:])
            rows.append({"impl": "reference_cpu(%s,%d threads)" % (label, os.cpu_count() or 1), "mode": mode, "dtype": "fp32", "log2n": 24,
                         "inplace": False, "scale": 1.0, "ms": ms, "GB/s": 8.0 * n / (ms * 1e-3) / 1e9, "Gelem/s": n / (ms * 1e-3) / 1e9})

    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    with open(args.out, "w") as f:
        json.dump({"gpu": torch.cuda.get_device_name(0), "peak_measured_GBs": peak, "rows": rows}, f, indent=0)
    print("wrote %d rows to %s" % (len(rows), args.out))


if __name__ == "__main__":
    main()
