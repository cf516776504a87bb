#!/usr/bin/env python
"""One launch of each register-resident kernel (row-wise cast / encode, blocked, single-launch amax-scaled cast) for ncu."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "fp8-emulation-toolkit_b200")):
    sys.path.insert(0, p)
import fpemu_b200 as F  # noqa: E402

x = torch.randn(32768, 4096, device="cuda") * 0.02            # 512 MiB fp32
y = torch.empty_like(x)
c = torch.empty(x.shape, dtype=torch.uint8, device="cuda")
s = torch.empty(32768, dtype=torch.float32, device="cuda")
for _ in range(2):
    F.cast_rowwise(x, "E4M3_RNE", out=y)
    F.encode_rowwise(x, "E4M3_RNE", codes=c, scales=s)
    F.decode_rowwise(c, s, "E4M3", out=y)
    F.cast_blocked(x.view(-1), "E4M3_RNE", 768, out=y.view(-1), n=(x.numel() // 768) * 768)
    F.cast_rows_scaled(x, "E5M2_RNE", s, out=y)
xs = torch.randn(12 * 384 * 768, device="cuda")               # a BERT activation: 13.5 MiB
ys = torch.empty_like(xs)
for _ in range(2):
    F.cast_amax_scaled(xs, "E4M3_RNE", out=ys)
torch.cuda.synchronize()
print("prof_rows ok")
