#!/usr/bin/env python
"""Host-to-host path: throughput over staging-chunk sizes, and the box's pinned-copy ceiling.

    python tools/e2e_chunks.py                       # fpemu_cast_host (one cast) over chunk sizes, 2^28 fp32 elements
    python tools/e2e_chunks.py --multi               # fpemu_cast_host_multi (13 casts, one upload) over chunk sizes
    python tools/e2e_chunks.py --copy-only --gpus N  # no kernel at all: concurrent pinned H2D + D2H on N GPUs at once
                                                     # (one thread per GPU), H2D alone, D2H alone, and the bench's 1:13 mix

--copy-only answers "is the end-to-end number the host fabric or the code": every GPU of one box hangs off the same
host memory system, so the aggregate pinned-copy rate at N GPUs is the ceiling of any host-in / host-out path."""
import argparse
import json
import os
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "fp8-emulation-toolkit_b200")):
    sys.path.insert(0, p)
import fpemu_b200 as F  # noqa: E402


def copy_only(ngpus, n, chunk, reps):
    """Aggregate GB/s of chunked pinned copies with no kernel: per GPU one thread, two streams."""
    bufs = []
    for d in range(ngpus):
        torch.cuda.set_device(d)
        bufs.append((torch.randn(n).pin_memory(), torch.empty(n).pin_memory(), torch.empty(n, device="cuda:%d" % d),
                     torch.cuda.Stream(d), torch.cuda.Stream(d)))

    def run(kind):
        def worker(d):
            torch.cuda.set_device(d)
            h_in, h_out, dbuf, s_in, s_out = bufs[d]
            for _ in range(reps):
                for off in range(0, n, chunk):
                    m = min(chunk, n - off)
                    if kind in ("h2d", "duplex"):
                        with torch.cuda.stream(s_in):
                            dbuf[off:off + m].copy_(h_in[off:off + m], non_blocking=True)
                    if kind in ("d2h", "duplex"):
                        with torch.cuda.stream(s_out):
                            h_out[off:off + m].copy_(dbuf[off:off + m], non_blocking=True)
                    if kind == "mix13":                              # the bench's sweep: 1 upload, 13 downloads per chunk
                        with torch.cuda.stream(s_in):
                            dbuf[off:off + m].copy_(h_in[off:off + m], non_blocking=True)
                        with torch.cuda.stream(s_out):
                            for _k in range(13):
                                h_out[off:off + m].copy_(dbuf[off:off + m], non_blocking=True)
            s_in.synchronize()
            s_out.synchronize()
        for d in range(ngpus):
            torch.cuda.synchronize(d)
        ts = [threading.Thread(target=worker, args=(d,)) for d in range(ngpus)]
        t0 = time.perf_counter()
        for t in ts:
            t.start()
        for t in ts:
            t.join()
        dt = time.perf_counter() - t0
        per_rep = {"h2d": 4.0 * n, "d2h": 4.0 * n, "duplex": 8.0 * n, "mix13": 14 * 4.0 * n}[kind]
        return per_rep * reps * ngpus / dt / 1e9
    out = {"gpus": ngpus, "elems": n, "chunk_MiB": (4 * chunk) >> 20}
    for kind in ("h2d", "d2h", "duplex", "mix13"):
        run(kind)
        out[kind + "_GBps_aggregate"] = round(run(kind), 1)
    out["mix13_as_bench_e2e_GBps"] = round(out["mix13_GBps_aggregate"] * 108.0 / 56.0, 1)   # 108 algorithmic B per 56 bus B
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--copy-only", action="store_true")
    ap.add_argument("--multi", action="store_true")
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--elems", type=int, default=1 << 28)
    args = ap.parse_args()
    n = args.elems
    if args.copy_only:
        g = 1
        while g <= min(args.gpus, torch.cuda.device_count()):
            print(json.dumps(copy_only(g, n, 1 << 24, 2)), flush=True)
            g *= 2
        return
    h_in = torch.randn(n).pin_memory()
    modes = ["%s_%s" % (f, r) for f in ("E5M2", "E4M3", "E3M4", "BFLOAT16") for r in ("RNE", "STOCHASTIC", "RTZ")]
    h_outs = [torch.empty(n).pin_memory() for _ in range(13 if args.multi else 1)]
    for lg in (20, 21, 22, 23, 24, 25):
        hc = F.HostCaster(torch.float32, chunk_elems=1 << lg)
        if args.multi:
            jobs = [(h_outs[i], m, 1.0, i) for i, m in enumerate(modes)] + [(h_outs[12], "E4M3_RNE", "amax", 0)]
            fn = lambda: hc.run_multi(h_in, jobs)                       # noqa: E731
            nbytes = 8.0 * n * 12 + 12.0 * n
        else:
            fn = lambda: hc.run(h_in, h_outs[0], "E4M3_RNE")             # noqa: E731
            nbytes = 8.0 * n
        fn()
        ts = []
        for _ in range(3):
            t0 = time.perf_counter()
            fn()
            ts.append(time.perf_counter() - t0)
        print("chunk 2^%d elems (%d MiB): %.1f GB/s (algorithmic)" % (lg, (4 << lg) >> 20, nbytes / min(ts) / 1e9), flush=True)


if __name__ == "__main__":
    main()
