#!/usr/bin/env python
"""BASELINE configs[0] (C1): the reference's own CPU op, unmodified, on this host — 16 Mi fp32 elements, E4M3_RNE (and
E5M2_RNE for the threaded loop), best of 10 after 2 warm-ups, n and n-1, AVX-512 and scalar builds, all threads and 1 —
next to this library on the GPU (device-resident, and host-to-host through fpemu_cast_host)."""
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "fp8-emulation-toolkit_b200")):
    sys.path.insert(0, p)
from oracle import ref_loader  # noqa: E402

N = 16 * 1024 * 1024
x = torch.randn(N, generator=torch.Generator().manual_seed(0))
rows = []
for build in ("fpemu_cpp_ref_avx512", "fpemu_cpp_ref_scalar"):
    if not ref_loader.available(build) or (build.endswith("avx512") and not ref_loader.host_has_avx512()):
        continue
    mod = ref_loader.load(build)
    for threads in (os.cpu_count() or 1, 1):
        torch.set_num_threads(threads)
        os.environ["OMP_NUM_THREADS"] = str(threads)
        for mode in ("E4M3_RNE", "E5M2_RNE"):
            for n in (N, N - 1):
                xx = x[:n].contiguous()
                ts = []
                for _ in range(12):
                    t0 = time.perf_counter()
                    mod.forward(xx, mode, n, False, 1.0, False, 1)
                    ts.append(time.perf_counter() - t0)
                best = min(ts[2:])
                rows.append({"impl": build, "threads": threads, "mode": mode, "n": n, "ms": best * 1e3,
                             "Gelem/s": n / best / 1e9, "GB/s": 8.0 * n / best / 1e9})
                print(rows[-1], flush=True)
if torch.cuda.is_available():
    import fpemu_b200 as F
    xd = x.cuda()
    yd = torch.empty_like(xd)
    for mode in ("E4M3_RNE", "E5M2_RNE"):
        for _ in range(3):
            F.cast(xd, mode, out=yd)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ts = []
        for _ in range(10):
            a.record(); F.cast(xd, mode, out=yd); b.record(); torch.cuda.synchronize()
            ts.append(a.elapsed_time(b))
        rows.append({"impl": "libfpemu_b200 (device-resident, L2-resident at this size)", "mode": mode, "n": N, "ms": min(ts),
                     "Gelem/s": N / min(ts) / 1e6, "GB/s": 8.0 * N / min(ts) / 1e6})
        print(rows[-1], flush=True)
    hx = x.pin_memory()
    hy = torch.empty_like(hx).pin_memory()
    hc = F.HostCaster(dtype=torch.float32, chunk_elems=1 << 22)
    for _ in range(2):
        hc.run(hx, hy, "E4M3_RNE")
    ts = []
    for _ in range(10):
        t0 = time.perf_counter(); hc.run(hx, hy, "E4M3_RNE"); ts.append(time.perf_counter() - t0)
    rows.append({"impl": "libfpemu_b200 fpemu_cast_host (pinned host in, pinned host out)", "mode": "E4M3_RNE", "n": N,
                 "ms": min(ts) * 1e3, "Gelem/s": N / min(ts) / 1e9, "GB/s": 8.0 * N / min(ts) / 1e9})
    print(rows[-1], flush=True)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump({"host_cpus": os.cpu_count(), "rows": rows}, open(os.path.join(ROOT, "gpurun_out", "c1.json"), "w"), indent=0)
