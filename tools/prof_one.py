#!/usr/bin/env python
"""Run one cast configuration a few times (target for ncu):  python tools/prof_one.py MODE DTYPE LOG2N [SCALE]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "fp8-emulation-toolkit_b200")):
    sys.path.insert(0, p)
import fpemu_b200 as F  # noqa: E402

mode, dt, lg = sys.argv[1], sys.argv[2], int(sys.argv[3])
scale = float(sys.argv[4]) if len(sys.argv) > 4 else 1.0
dtype = {"fp32": torch.float32, "bf16": torch.bfloat16, "fp16": torch.float16}[dt]
x = torch.randn(1 << lg, device="cuda").to(dtype)
y = torch.empty_like(x)
for _ in range(5):
    F.cast(x, mode, out=y, seed=3, scale=scale)
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(10):
    F.cast(x, mode, out=y, seed=3, scale=scale)
b.record()
torch.cuda.synchronize()
ms = a.elapsed_time(b) / 10
print("%s %s 2^%d scale=%g: %.4f ms  %.1f GB/s" % (mode, dt, lg, scale, ms, 2.0 * x.numel() * x.element_size() / ms / 1e6))
