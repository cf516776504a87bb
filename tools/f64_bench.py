"""Throughput of the float64 cast (16 B / element), 2^28 elements."""
import os, sys, statistics, json, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "fp8-emulation-toolkit_b200")):
    sys.path.insert(0, p)
import fpemu_b200 as F
n = 1 << 28
x = torch.randn(n, device="cuda", dtype=torch.float64)
y = torch.empty_like(x)
rows = []
for mode, kw in (("E4M3_RNE", {}), ("E4M3_RNE", {"scale": 3.7}), ("E5M2_STOCHASTIC", {"seed": 1}), ("BFLOAT16_RNE", {}), ("FP4_NEAREST", {})):
    for _ in range(3):
        F.cast(x, mode, out=y, **kw)
    torch.cuda.synchronize()
    ts = []
    for _ in range(10):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); F.cast(x, mode, out=y, **kw); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ms = statistics.median(ts)
    rows.append({"mode": mode, "scaled": "scale" in kw, "ms": round(ms, 4), "GB/s": round(16 * n / ms / 1e6, 1)})
    print(rows[-1], flush=True)
