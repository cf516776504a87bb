#!/usr/bin/env python
"""Targets for `ncu --set full` of the kernels added in round 2 (each path runs three times on 2^28 elements):
  python tools/prof_r2.py [fp32|bf16]
    cast_vec_kernel<..., ANYOFF>   stochastic cast of a slab whose element offset is not a multiple of 16
    cast_shift_kernel              input and output off a 32-byte boundary by different amounts
    amax_multi_kernel + cast_multi_kernel (scale derived from the maxima in the cast kernel): 8-slab list, amax-scaled
    abs_sum_partial_kernel + mean_scale_kernel: per-tensor `mean` scaling"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "fp8-emulation-toolkit_b200")):
    sys.path.insert(0, p)
import fpemu_b200 as F  # noqa: E402

dt = {"fp32": torch.float32, "bf16": torch.bfloat16}[sys.argv[1] if len(sys.argv) > 1 else "fp32"]
n = 1 << 28
x = torch.randn(n + 64, device="cuda").to(dt)
y = torch.empty(n + 64, dtype=dt, device="cuda")
parts = [x[k * (n // 8):(k + 1) * (n // 8)] for k in range(8)]
outs = [y[k * (n // 8):(k + 1) * (n // 8)] for k in range(8)]
plan = F.MultiTensorPlan(parts, outs)
for _ in range(3):
    F.cast(x[:n], "E4M3_STOCHASTIC", out=y[:n], seed=7, elem_offset=3)
    F.cast(x[1:n + 1], "E4M3_RNE", out=y[3:n + 3])
    F.cast(x[1:n + 1], "E4M3_STOCHASTIC", out=y[3:n + 3], seed=7, elem_offset=5)
    plan.amax_scaled_cast("E4M3_RNE")
    F.mean_scale(x[:n], 1.5258789e-05)
torch.cuda.synchronize()
print("done")
