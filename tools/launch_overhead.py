#!/usr/bin/env python
"""Host-side cost of one cast call on a tiny tensor (launch-bound regime), eager vs CUDA-graph replay."""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "fp8-emulation-toolkit_b200")):
    sys.path.insert(0, p)
import fpemu_b200 as F  # noqa: E402
import fpemu_cuda  # noqa: E402

x = torch.randn(4096, device="cuda")
y = torch.empty_like(x)
N = 20000
for name, fn in (("ops.cast(out=)", lambda: F.cast(x, "E4M3_RNE", out=y)),
                 ("fpemu_cuda.forward(out-of-place)", lambda: fpemu_cuda.forward(x, "E4M3_RNE", 4096, False, 1.0, False, 1)),
                 ("torch: y.copy_(x)", lambda: y.copy_(x))):
    for _ in range(200):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(N):
        fn()
    t_host = (time.perf_counter() - t0) / N * 1e6
    torch.cuda.synchronize()
    t_all = (time.perf_counter() - t0) / N * 1e6
    print("%-36s host %.2f us/call, end-to-end %.2f us/call" % (name, t_host, t_all))

# CUDA graph of 100 casts on distinct small tensors (the per-parameter loop of a training step)
xs = [torch.randn(16384, device="cuda") for _ in range(100)]
s = torch.cuda.Stream()
with torch.cuda.stream(s):
    for t in xs:
        F.cast(t, "E5M2_STOCHASTIC", out=t, seed=1)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=s):
        for t in xs:
            F.cast(t, "E5M2_STOCHASTIC", out=t, seed=1)
torch.cuda.synchronize()
for _ in range(5):
    g.replay()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(200):
    g.replay()
torch.cuda.synchronize()
print("graph of 100 small casts: %.2f us per cast" % ((time.perf_counter() - t0) / 200 / 100 * 1e6))
