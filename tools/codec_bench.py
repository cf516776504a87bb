#!/usr/bin/env python
"""Throughput of fpemu_encode / fpemu_decode next to the value-to-value cast (one B200, 2^30 elements fp32 /
2^30 elements 16-bit).  Algorithmic bytes: encode/decode 5 B (fp32) or 3 B (16-bit) per element, cast 8 B / 4 B.

    python tools/codec_bench.py [--out gpurun_out/codec.json]"""
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


def timed(fn, reps=20):
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
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "codec.json"))
    ap.add_argument("--log2n", type=int, default=30)
    args = ap.parse_args()
    n = 1 << args.log2n
    rows = []
    for dname, dtype in (("fp32", torch.float32), ("bf16", torch.bfloat16), ("fp16", torch.float16)):
        esz = 4 if dtype == torch.float32 else 2
        x = torch.randn(n, device="cuda", dtype=torch.float32).to(dtype)
        y = torch.empty_like(x)
        c = torch.empty(n, dtype=torch.uint8, device="cuda")
        d_scale = torch.tensor([3.7], device="cuda")
        for fmt in ("E5M2", "E4M3", "E3M4"):
            for rnd in ("RNE", "STOCHASTIC", "RTZ"):
                mode = "%s_%s" % (fmt, rnd)
                for scaled in (False, True):
                    if scaled and rnd != "RNE":
                        continue
                    kw = {"d_scale": d_scale} if scaled else {}
                    ms = timed(lambda: F.encode(x, mode, out=c, seed=7, **kw))
                    rows.append({"op": "encode", "mode": mode, "dtype": dname, "scaled": scaled, "ms": ms,
                                 "GB/s": n * (esz + 1) / ms / 1e6, "Gelem/s": n / ms / 1e6})
                    print(rows[-1], flush=True)
            for scaled in (False, True):
                kw = {"d_scale": d_scale} if scaled else {}
                ms = timed(lambda: F.decode(c, fmt, out=y, **kw))
                rows.append({"op": "decode", "mode": fmt, "dtype": dname, "scaled": scaled, "ms": ms,
                             "GB/s": n * (esz + 1) / ms / 1e6, "Gelem/s": n / ms / 1e6})
                print(rows[-1], flush=True)
            ms = timed(lambda: F.cast(x, fmt + "_RNE", out=y))
            rows.append({"op": "cast", "mode": fmt + "_RNE", "dtype": dname, "scaled": False, "ms": ms,
                         "GB/s": n * 2 * esz / ms / 1e6, "Gelem/s": n / ms / 1e6})
            print(rows[-1], flush=True)
        del x, y, c
        torch.cuda.empty_cache()
    # per-channel wire form on weight-matrix shapes
    for dname, dtype in (("fp32", torch.float32), ("bf16", torch.bfloat16)):
        esz = 4 if dtype == torch.float32 else 2
        for shape in ((32000, 4096), (11008, 4096), (4096, 11008), (4096, 4096)):
            x = (torch.randn(shape, device="cuda") * 0.02).to(dtype)
            c = torch.empty(shape, dtype=torch.uint8, device="cuda")
            sc = torch.empty(shape[0], dtype=torch.float32, device="cuda")
            y = torch.empty_like(x)
            nn = x.numel()
            ms = timed(lambda: F.encode_rowwise(x, "E4M3_RNE", codes=c, scales=sc))
            rows.append({"op": "encode_rowwise", "mode": "E4M3_RNE", "dtype": dname, "shape": list(shape), "ms": ms,
                         "GB/s": nn * (esz + 1) / ms / 1e6, "Gelem/s": nn / ms / 1e6})
            print(rows[-1], flush=True)
            ms = timed(lambda: F.decode_rowwise(c, sc, "E4M3", out=y))
            rows.append({"op": "decode_rowwise", "mode": "E4M3", "dtype": dname, "shape": list(shape), "ms": ms,
                         "GB/s": nn * (esz + 1) / ms / 1e6, "Gelem/s": nn / ms / 1e6})
            print(rows[-1], flush=True)
            ms = timed(lambda: F.cast_rowwise(x, "E4M3_RNE", out=y))
            rows.append({"op": "cast_rowwise", "mode": "E4M3_RNE", "dtype": dname, "shape": list(shape), "ms": ms,
                         "GB/s": nn * 2 * esz / ms / 1e6, "Gelem/s": nn / ms / 1e6})
            print(rows[-1], flush=True)
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    json.dump({"gpu": torch.cuda.get_device_name(0), "log2n": args.log2n, "rows": rows}, open(args.out, "w"), indent=0)


if __name__ == "__main__":
    main()
