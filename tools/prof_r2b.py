#!/usr/bin/env python
"""Targets for `ncu --set full` of the kernels added late in round 2 (each runs twice, 2^28 elements unless noted):
    cast_vec_kernel<FP4,...>   packed FP4 on bf16 and fp32
    rows_tile_kernel           bf16 rows of 576 and 4608 columns
    group_reg_kernel           bf16 rows of 768 columns (for comparison)
    encode_vec_kernel          stochastic E4M3 codes from bf16
    cast_f64_kernel            float64 tensor, scaled E4M3_RNE (2^27 elements)
    amax_multi_kernel, push_kernel, cast_multi_kernel<..., XCHG>   the peer-memory exchange with this GPU as a world of one"""
import ctypes
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "fp8-emulation-toolkit_b200")):
    sys.path.insert(0, p)
import fpemu_b200 as F  # noqa: E402
from fpemu_b200 import _lib as L, ops  # noqa: E402

n = 1 << 28
xb = (torch.randn(n, device="cuda") * 0.3).to(torch.bfloat16)
yb = torch.empty_like(xb)
xf = torch.randn(n, device="cuda")
yf = torch.empty_like(xf)
xd = torch.randn(n // 2, device="cuda", dtype=torch.float64)
yd = torch.empty_like(xd)
codes = torch.empty(n, dtype=torch.uint8, device="cuda")
parts = [xb[k * (n // 8):(k + 1) * (n // 8)] for k in range(8)]
outs = [yb[k * (n // 8):(k + 1) * (n // 8)] for k in range(8)]
plan = F.MultiTensorPlan(parts, outs)
lib = L.lib()
buf = torch.zeros(int(lib.fpemu_exchange_floats(plan.count, 1)), dtype=torch.float32, device="cuda")
ptrs = (ctypes.c_void_p * 1)(buf.data_ptr())
local = torch.zeros(2 * plan.count, dtype=torch.float32, device="cuda")
scales = torch.empty(plan.count, dtype=torch.float32, device="cuda")
stream = ops._stream(plan.d_table)
tab = (plan.d_table.data_ptr(), plan.h_table.data_ptr(), plan.count)
for epoch in (1, 2):
    F.cast(xb, "FP4_NEAREST", out=yb)
    F.cast(xf, "FP4_NEAREST", out=yf)
    for cols in (576, 4608, 768):
        m = (n // cols) * cols
        F.cast_rowwise(xb[:m].view(-1, cols), "E4M3_RNE", out=yb[:m].view(-1, cols))
    F.encode(xb, "E4M3_STOCHASTIC", out=codes, seed=3)
    F.cast(xd, "E4M3_RNE", out=yd, scale=3.7)
    L.check(lib.fpemu_amax_multi_push(*tab, local.data_ptr(), ctypes.addressof(ptrs), 1, 0, epoch, stream), "push")
    L.check(lib.fpemu_cast_multi_xchg(*tab, L.FMT_E4M3, L.RND_RNE, 0, buf.data_ptr(), 1, epoch, 448.0, scales.data_ptr(), 0,
                                      stream), "xchg")
torch.cuda.synchronize()
print("done")
