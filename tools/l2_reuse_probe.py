"""Does the cast pass of an amax-scaled sweep find its input in L2 when the tensor list is swept in short runs
(amax of run i, cast of run i, amax of run i+1, ...)?  One GPU, rank 0's slabs of an N-way shard of the Llama-7B weight
set, bf16; runs of about equal size.  Prints ms per sweep for several run counts."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "fp8-emulation-toolkit_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shard", type=int, default=8)
    ap.add_argument("--depths", default="1,8,16,32,64,128")
    ap.add_argument("--reps", type=int, default=10)
    args = ap.parse_args()
    import torch
    import bench
    from fpemu_b200 import dist as fd, ops
    dev = torch.device("cuda", 0)
    slabs = []
    for t, (_, r, c) in enumerate(bench.llama7b_shapes()):
        r0, r1 = fd.slab_bounds(r, args.shard, 0)
        slabs.append((torch.randn((r1 - r0) * c, device=dev) * 0.02).to(torch.bfloat16).view(r1 - r0, c))
    outs = [torch.empty_like(x) for x in slabs]
    plan = ops.MultiTensorPlan(slabs, outs)
    total = sum(x.numel() for x in slabs)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    res = {}
    ref = None
    for rnd in range(2):
        for depth in [int(d) for d in args.depths.split(",")]:
            fn = (lambda: plan.amax_scaled_cast("E4M3_RNE", reduce_fn=(lambda v: None), pipeline=depth))
            for _ in range(3):
                _, sc = fn()
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()                      # replayed from a graph: the host's launch rate is not the subject
            with torch.cuda.graph(g):
                _, sc = fn()
            g.replay()
            torch.cuda.synchronize()
            a.record()
            for _ in range(args.reps):
                g.replay()
            b.record()
            torch.cuda.synchronize()
            ms = a.elapsed_time(b) / args.reps
            res.setdefault(depth, []).append(round(ms, 4))
            if ref is None:
                ref = sc.clone()
            assert torch.equal(ref, sc)
    print(json.dumps({"shard": args.shard, "elements": total, "MB_in": total * 2 / 1e6, "ms_per_sweep_by_runs": res,
                      "GBps_6B_per_elem": {d: round(6 * total / min(v) / 1e6, 1) for d, v in res.items()}}))


if __name__ == "__main__":
    main()
