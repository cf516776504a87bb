#!/usr/bin/env python
"""Host cost per call of the layers above the C ABI (small tensor, so the GPU is never the bottleneck)."""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "fp8-emulation-toolkit_b200")):
    sys.path.insert(0, p)
import fpemu_b200 as F  # noqa: E402
from fpemu_b200 import _lib as L  # noqa: E402
import fpemu_cuda  # noqa: E402
from mpemu.pytquant.cuda.fpemu import FPEmuOp  # noqa: E402
from mpemu import qutils  # noqa: E402


def per_call(fn, n=3000):
    for _ in range(200):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        fn()
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    return (t1 - t0) / n * 1e6


x = torch.randn(4096, device="cuda")
y = torch.empty_like(x)
xg = torch.randn(4096, device="cuda", requires_grad=True)
lib = L.lib()
stream = torch._C._cuda_getCurrentRawStream(0)
work = torch.zeros(4098, device="cuda")
cfg = qutils.TensorQuantConfig("e4m3", "rne", "per-tensor")
cfg_plain = qutils.TensorQuantConfig("e5m2", "stochastic")
rows = {
    "raw ctypes fpemu_cast": lambda: lib.fpemu_cast(x.data_ptr(), y.data_ptr(), 4096, 0, 1, 1, 0, 1.0, None, 0, 0, stream),
    "raw ctypes fpemu_cast_amax_scaled_fused": lambda: lib.fpemu_cast_amax_scaled_fused(
        x.data_ptr(), y.data_ptr(), 4096, 0, 1, 1, 0, 448.0, work.data_ptr(), 4098, 0, 0, stream),
    "ops.cast(out=)": lambda: F.cast(x, "E4M3_RNE", out=y),
    "ops.cast (allocates)": lambda: F.cast(x, "E4M3_RNE"),
    "ops.cast_amax_scaled(out=)": lambda: F.cast_amax_scaled(x, "E4M3_RNE", out=y),
    "fpemu_cuda.forward": lambda: fpemu_cuda.forward(x, "E4M3_RNE", 4096, False, 1.0, False, 1),
    "FPEmuOp.apply (no grad)": lambda: FPEmuOp.apply(x, "E4M3_RNE", False, 1.0),
    "FPEmuOp.apply (grad)": lambda: FPEmuOp.apply(xg, "E4M3_RNE", False, 1.0),
    "quantize_tensor per-tensor (no grad)": lambda: qutils.quantize_tensor(x, cfg, None, inplace=False),
    "quantize_tensor per-tensor (grad)": lambda: qutils.quantize_tensor(xg, cfg, None, inplace=False),
    "quantize_tensor unscaled stochastic": lambda: qutils.quantize_tensor(x, cfg_plain, None, inplace=False),
    "torch x.mul(2.0) for scale": lambda: x.mul(2.0),
    "torch.empty_like": lambda: torch.empty_like(x),
}
for k, fn in rows.items():
    print("%-45s %6.2f us" % (k, per_call(fn)))
