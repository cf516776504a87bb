#!/usr/bin/env python
"""Row-wise (per-channel) cast and encode over row lengths and dtypes; buffers rotate so the working set exceeds L2."""
import json
import os
import statistics
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "fp8-emulation-toolkit_b200")):
    sys.path.insert(0, p)
import fpemu_b200 as F  # noqa: E402


def timed(fn, nbuf, reps=30):
    for i in range(4):
        fn(i % nbuf)
    torch.cuda.synchronize()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    for i, (a, b) in enumerate(evs):
        a.record()
        fn(i % nbuf)
        b.record()
    torch.cuda.synchronize()
    return statistics.median(a.elapsed_time(b) for a, b in evs)


rows_out = []
for dname, dtype in (("fp32", torch.float32), ("bf16", torch.bfloat16)):
    esz = 4 if dtype == torch.float32 else 2
    for cols in (576, 1024, 2048, 4096, 4608, 8192, 11008, 16384):
        rows = max(64, (1 << 27) // cols)                     # ~128 Mi elements per tensor
        n = rows * cols
        nbuf = max(2, min(8, (600 << 20) // (n * esz * 2)))
        xs = [(torch.randn(rows, cols, device="cuda") * 0.02).to(dtype) for _ in range(nbuf)]
        ys = [torch.empty_like(x) for x in xs]
        cs = [torch.empty(rows, cols, dtype=torch.uint8, device="cuda") for _ in range(nbuf)]
        sc = torch.empty(rows, dtype=torch.float32, device="cuda")
        ms = timed(lambda i: F.cast_rowwise(xs[i], "E4M3_RNE", out=ys[i]), nbuf)
        me = timed(lambda i: F.encode_rowwise(xs[i], "E4M3_RNE", codes=cs[i], scales=sc), nbuf)
        r = {"dtype": dname, "rows": rows, "cols": cols, "cast_rowwise_GB/s": round(2 * esz * n / ms / 1e6),
             "encode_rowwise_GB/s": round((esz + 1) * n / me / 1e6)}
        rows_out.append(r)
        print(r, flush=True)
        del xs, ys, cs
        torch.cuda.empty_cache()
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(rows_out, open(os.path.join(ROOT, "gpurun_out", "rowwise.json"), "w"), indent=0)
