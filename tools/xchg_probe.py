"""Where the time of a row-sharded sweep goes with either exchange (run under torchrun, N >= 2):
  * whole sweeps, the two exchanges interleaved round after round (so that neither always runs on the warmer GPU);
  * the sweep's stages timed apart with CUDA events on the launching stream: amax (+ push), exchange, cast.
Prints one JSON line on rank 0.  Usage:
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 tools/xchg_probe.py"""
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
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--rounds", type=int, default=3)
    ap.add_argument("--dtype", default="bf16")
    args = ap.parse_args()
    import torch
    import torch.distributed as dist
    import bench
    from fpemu_b200 import dist as fd
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", rank)))
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    dtype = {"bf16": torch.bfloat16, "fp32": torch.float32}[args.dtype]
    slabs, offsets, cols = [], [], []
    for t, (_, r, c) in enumerate(bench.llama7b_shapes()):
        r0, r1 = fd.slab_bounds(r, world, rank)
        g = torch.Generator(device=dev).manual_seed(1000003 * t + rank)
        slabs.append((torch.randn((r1 - r0) * c, device=dev, generator=g) * 0.02).to(dtype).view(r1 - r0, c))
        offsets.append(r0)
        cols.append(c)
    outs = [torch.empty_like(x) for x in slabs]
    kw = dict(row_offsets=offsets, cols=cols, outs=outs)
    plans = {"nccl": fd.ShardedPlan(slabs, exchange="nccl", **kw), "peer": fd.ShardedPlan(slabs, exchange="peer", **kw)}
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def timed(fn):
        for _ in range(3):
            fn()
        dist.barrier()
        torch.cuda.synchronize()
        a.record()
        for _ in range(args.reps):
            fn()
        b.record()
        torch.cuda.synchronize()
        t = torch.tensor([a.elapsed_time(b) / args.reps], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return round(float(t.item()), 4)

    whole = {}
    for rnd in range(args.rounds):
        whole.setdefault("nccl", []).append(timed(lambda: plans["nccl"].amax_scaled_cast("E4M3_RNE")))
        whole.setdefault("peer", []).append(timed(lambda: plans["peer"].amax_scaled_cast("E4M3_RNE")))
    # stages: events between the launches of one sweep, accumulated over reps
    from fpemu_b200 import _lib as L, ops
    import ctypes
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
    stages = {}
    plan = plans["nccl"].plan
    vec = torch.zeros(plan.count, dtype=torch.float32, device=dev)
    scales = torch.empty_like(vec)
    acc = [0.0, 0.0, 0.0]
    for i in range(args.reps + 3):
        vec.zero_()
        evs[0].record()
        plan.amax(out=vec)
        evs[1].record()
        dist.all_reduce(fd._as_bits(vec), op=dist.ReduceOp.MAX)
        evs[2].record()
        plan._from_amax("E4M3_RNE", vec, 448.0, scales, 0)
        evs[3].record()
        torch.cuda.synchronize()
        if i >= 3:
            for k in range(3):
                acc[k] += evs[k].elapsed_time(evs[k + 1]) / args.reps
    stages["nccl"] = {"amax": round(acc[0], 4), "exchange": round(acc[1], 4), "cast": round(acc[2], 4)}
    ex = plans["peer"].peer
    lib = L.lib()
    stream = ops._stream(plan.d_table)
    acc = [0.0, 0.0]
    for i in range(args.reps + 3):
        ex.epoch += 1
        evs[0].record()
        L.check(lib.fpemu_amax_multi_push(plan.d_table.data_ptr(), plan.h_table.data_ptr(), plan.count, ex.local.data_ptr(),
                                          ctypes.addressof(ex.ptrs), ex.world, ex.rank, ex.epoch, stream), "push")
        evs[1].record()
        L.check(lib.fpemu_cast_multi_xchg(plan.d_table.data_ptr(), plan.h_table.data_ptr(), plan.count, L.FMT_E4M3, L.RND_RNE, 0,
                                          ex.buf.data_ptr(), ex.world, ex.epoch, 448.0, scales.data_ptr(), 0, stream), "xchg")
        evs[2].record()
        torch.cuda.synchronize()
        if i >= 3:
            for k in range(2):
                acc[k] += evs[k].elapsed_time(evs[k + 1]) / args.reps
    stages["peer"] = {"amax+push": round(acc[0], 4), "cast (polls the peers' words)": round(acc[1], 4)}
    if rank == 0:
        print(json.dumps({"n_gpus": world, "dtype": args.dtype, "ms_per_sweep_rounds": whole, "stages_ms_rank0": stages}))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
