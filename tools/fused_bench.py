#!/usr/bin/env python
"""Per-tensor amax-scaled E4M3 cast, single cooperative launch vs memset + amax + cast, over tensor sizes
(median of CUDA-event-timed calls; buffers rotate so the working set exceeds L2 for the larger sizes)."""
import json
import os
import statistics
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "fp8-emulation-toolkit_b200")):
    sys.path.insert(0, p)
import fpemu_b200 as F  # noqa: E402


def timed(fn, nbuf, reps=40):
    for i in range(5):
        fn(i % nbuf)
    torch.cuda.synchronize()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    for i, (a, b) in enumerate(evs):
        a.record()
        fn(i % nbuf)
        b.record()
    torch.cuda.synchronize()
    return statistics.median(a.elapsed_time(b) for a, b in evs) * 1e3


rows = []
for dname, dtype in (("fp32", torch.float32), ("bf16", torch.bfloat16)):
    for n in (1 << 12, 1 << 16, 768 * 768, 1 << 20, 768 * 3072, 12 * 384 * 768, 1 << 22, 4849664, 1 << 23):
        nbuf = max(1, min(32, (256 << 20) // (n * 8)))
        xs = [torch.randn(n, device="cuda").to(dtype) for _ in range(nbuf)]
        ys = [torch.empty_like(x) for x in xs]
        r = {"dtype": dname, "n": n}
        for name, fused in (("fused_us", True), ("three_launch_us", False)):
            r[name] = round(timed(lambda i: F.cast_amax_scaled(xs[i], "E4M3_RNE", out=ys[i], fused=fused), nbuf), 2)
        # the same inside a CUDA graph of 20 calls: device time without host launch cost
        for name, fused in (("fused_graph_us", True), ("three_launch_graph_us", False)):
            s = torch.cuda.Stream()
            with torch.cuda.stream(s):
                F.cast_amax_scaled(xs[0], "E4M3_RNE", out=ys[0], fused=fused)
                s.synchronize()
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=s):
                    for i in range(20):
                        F.cast_amax_scaled(xs[i % nbuf], "E4M3_RNE", out=ys[i % nbuf], fused=fused)
            r[name] = round(timed(lambda i: g.replay(), 1, reps=20) / 20, 2)
        rows.append(r)
        print(r, flush=True)
        del xs, ys
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(rows, open(os.path.join(ROOT, "gpurun_out", "fused.json"), "w"), indent=0)
