#!/usr/bin/env python
"""fpemu_cast_host throughput over staging-chunk sizes (pinned host in/out, 2^28 fp32 elements)."""
import os, sys, time
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "fp8-emulation-toolkit_b200")):
    sys.path.insert(0, p)
import fpemu_b200 as F
n = 1 << 28
h_in = torch.randn(n).pin_memory(); h_out = torch.empty(n).pin_memory()
for lg in (20, 21, 22, 23, 24, 25):
    hc = F.HostCaster(torch.float32, chunk_elems=1 << lg)
    hc.run(h_in, h_out, "E4M3_RNE")
    ts = []
    for _ in range(4):
        t0 = time.perf_counter(); hc.run(h_in, h_out, "E4M3_RNE"); ts.append(time.perf_counter() - t0)
    print("chunk 2^%d elems (%d MiB): %.1f GB/s" % (lg, (4 << lg) >> 20, 8.0 * n / min(ts) / 1e9), flush=True)
