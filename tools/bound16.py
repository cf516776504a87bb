#!/usr/bin/env python
"""The instruction-bound corner of the path on 16-bit tensors, one table (one B200): stochastic 8-bit encode, FP4,
the 8-bit casts on bf16 / fp16, short bf16 rows.  Each row: median of `reps` launches timed with CUDA events on a
tensor larger than L2 (2^29 elements), algorithmic GB/s and the fraction of the measured copy peak.

    python tools/bound16.py [--out gpurun_out/bound16.json] [--peak 6545]"""
import argparse
import json
import os
import statistics
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "fp8-emulation-toolkit_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

import fpemu_b200 as F  # noqa: E402


def timed(fn, reps):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    for a, b in evs:
        a.record()
        fn()
        b.record()
    torch.cuda.synchronize()
    return statistics.median(a.elapsed_time(b) for a, b in evs)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "bound16.json"))
    ap.add_argument("--log2n", type=int, default=29)
    ap.add_argument("--reps", type=int, default=15)
    ap.add_argument("--peak", type=float, default=0.0, help="copy peak in GB/s (default: MEASURED_PEAKS.json, else 6545)")
    ap.add_argument("--only", default="", help="substring filter on the op name")
    args = ap.parse_args()
    peak = args.peak
    if not peak:
        try:
            mp = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
            peak = float(mp.get("hbm_gbs") or 6545.0)
        except Exception:  # noqa: BLE001
            peak = 6545.0
    n = 1 << args.log2n
    rows = []

    def add(op, mode, dname, nbytes, fn):
        if args.only and args.only not in op:
            return
        ms = timed(fn, args.reps)
        rows.append({"op": op, "mode": mode, "dtype": dname, "ms": round(ms, 4), "GB/s": round(nbytes / ms / 1e6, 1),
                     "frac_peak": round(nbytes / ms / 1e6 / peak, 3)})
        print(rows[-1], flush=True)

    d_scale = torch.tensor([3.7], device="cuda")
    for dname, dtype in (("bf16", torch.bfloat16), ("fp16", torch.float16), ("fp32", torch.float32)):
        esz = 4 if dtype == torch.float32 else 2
        m = n if esz == 2 else n // 2
        x = (torch.randn(m, device="cuda") * 0.3).to(dtype)
        y = torch.empty_like(x)
        c = torch.empty(m, dtype=torch.uint8, device="cuda")
        for mode in ("FP4_NEAREST",):
            add("cast", mode, dname, 2 * esz * m, lambda: F.cast(x, mode, out=y))
            add("cast scaled", mode, dname, 2 * esz * m, lambda: F.cast(x, mode, out=y, d_scale=d_scale))
        for mode in ("E4M3_RNE", "E4M3_STOCHASTIC", "E5M2_STOCHASTIC", "E3M4_STOCHASTIC", "BFLOAT16_STOCHASTIC",
                     "FLOAT16_STOCHASTIC", "E4M3_IEEE_RNE"):
            add("cast", mode, dname, 2 * esz * m, lambda: F.cast(x, mode, out=y, seed=3))
        add("cast scaled", "E4M3_RNE", dname, 2 * esz * m, lambda: F.cast(x, "E4M3_RNE", out=y, d_scale=d_scale))
        add("cast scaled", "E4M3_STOCHASTIC", dname, 2 * esz * m, lambda: F.cast(x, "E4M3_STOCHASTIC", out=y, d_scale=d_scale, seed=3))
        for mode in ("E4M3_RNE", "E4M3_STOCHASTIC", "E5M2_STOCHASTIC", "E3M4_STOCHASTIC"):
            add("encode", mode, dname, (esz + 1) * m, lambda: F.encode(x, mode, out=c, seed=3))
        add("encode scaled", "E4M3_STOCHASTIC", dname, (esz + 1) * m, lambda: F.encode(x, "E4M3_STOCHASTIC", out=c, seed=3, d_scale=d_scale))
        del x, y, c
        torch.cuda.empty_cache()
    for cols in (576, 768, 1024, 4608, 11008):
        r = (1 << 28) // cols
        x = (torch.randn(r, cols, device="cuda") * 0.02).to(torch.bfloat16)
        y = torch.empty_like(x)
        add("rows %d" % cols, "E4M3_RNE", "bf16", 4 * x.numel(), lambda: F.cast_rowwise(x, "E4M3_RNE", out=y))
        del x, y
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    json.dump({"peak_GBs": peak, "log2n": args.log2n, "rows": rows}, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
