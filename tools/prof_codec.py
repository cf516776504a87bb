#!/usr/bin/env python
"""One launch of each codec kernel worth profiling (target for ncu):  python tools/prof_codec.py [LOG2N]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "fp8-emulation-toolkit_b200")):
    sys.path.insert(0, p)
import fpemu_b200 as F  # noqa: E402

lg = int(sys.argv[1]) if len(sys.argv) > 1 else 30
n = 1 << lg
x = torch.randn(n, device="cuda")
c = torch.empty(n, dtype=torch.uint8, device="cuda")
y = torch.empty_like(x)
for mode in ("E4M3_RNE", "E4M3_STOCHASTIC", "E5M2_RNE"):
    F.encode(x, mode, out=c, seed=3)
F.decode(c, "E5M2", out=y)
F.encode(x, "E4M3_RNE", out=c)
F.decode(c, "E4M3", out=y)
del y
xb = x.to(torch.bfloat16)
del x
yb = torch.empty_like(xb)
for mode in ("E4M3_RNE", "E4M3_STOCHASTIC"):
    F.encode(xb, mode, out=c, seed=3)
F.encode(xb, "E4M3_RNE", out=c)
F.decode(c, "E4M3", out=yb)
torch.cuda.synchronize()
print("prof_codec ok")
