#!/usr/bin/env python
"""BASELINE.json configs[4]: Llama-style 7B random-init weight set, per-tensor amax-scaled E4M3 QDQ, row-sharded
across the GPUs of one node with ONE NCCL all-reduce(MAX) over the float[291] amax vector.

    python tools/llama7b_sweep.py [--dtype bf16|fp32] [--reps 5]                      # 1 GPU
    python -m torch.distributed.run --nproc-per-node N ... tools/llama7b_sweep.py     # N GPUs (strong scaling)

Shapes: embed 32000x4096; 32 x {q,k,v,o 4096x4096; gate,up 11008x4096; down 4096x11008; 2 norms 4096}; final norm;
lm_head 32000x4096 = 291 tensors, 6.74 G elements.  Weights are generated on the device, per slab, from
(seed, tensor id, rank) as randn * 0.02.  Reported: aggregate algorithmic GB/s (amax read + cast read + write =
3 x element size per element), device time (max over ranks), time of the all-reduce alone."""
import argparse
import json
import os
import sys
import time

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "fp8-emulation-toolkit_b200")):
    sys.path.insert(0, p)
import fpemu_b200 as F  # noqa: E402
from fpemu_b200 import dist as fd  # noqa: E402


def llama7b_shapes():
    shapes = [("embed_tokens", 32000, 4096)]
    for i in range(32):
        for nm in ("q_proj", "k_proj", "v_proj", "o_proj"):
            shapes.append(("layers.%d.%s" % (i, nm), 4096, 4096))
        shapes.append(("layers.%d.gate_proj" % i, 11008, 4096))
        shapes.append(("layers.%d.up_proj" % i, 11008, 4096))
        shapes.append(("layers.%d.down_proj" % i, 4096, 11008))
        shapes.append(("layers.%d.input_layernorm" % i, 4096, 1))
        shapes.append(("layers.%d.post_attention_layernorm" % i, 4096, 1))
    shapes.append(("norm", 4096, 1))
    shapes.append(("lm_head", 32000, 4096))
    return shapes


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--dtype", default="bf16", choices=["bf16", "fp32", "fp16"])
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--mode", default="E4M3_RNE")
    ap.add_argument("--multi", action="store_true", help="multi-tensor launches (3 launches per sweep instead of 583)")
    ap.add_argument("--packed", action="store_true", help="write 8-bit codes instead of de-quantised values (implies --multi)")
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    dtype = {"bf16": torch.bfloat16, "fp32": torch.float32, "fp16": torch.float16}[args.dtype]
    esz = torch.empty(0, dtype=dtype).element_size()

    shapes = llama7b_shapes()
    assert len(shapes) == 291
    slabs, offsets, cols = [], [], []
    total = 0
    for t, (_, r, c) in enumerate(shapes):
        r0, r1 = fd.slab_bounds(r, world, rank)
        g = torch.Generator(device=dev).manual_seed(1000003 * t + rank)
        slabs.append((torch.randn((r1 - r0) * c, device=dev, dtype=torch.float32, generator=g) * 0.02).to(dtype).view(r1 - r0, c))
        offsets.append(r0)
        cols.append(c)
        total += r * c
    if args.packed:
        args.multi = True
        outs = [torch.empty(s.shape, dtype=torch.uint8, device=dev) for s in slabs]
    else:
        outs = [torch.empty_like(s) for s in slabs]
    local_elems = sum(s.numel() for s in slabs)

    plan = fd.ShardedPlan(slabs, row_offsets=offsets, cols=cols, outs=outs) if args.multi else None

    def run():
        if args.packed:
            return plan.amax_scaled_encode(args.mode, seed=7)
        if plan is not None:
            return plan.amax_scaled_cast(args.mode, seed=7)
        return fd.sharded_amax_scaled_cast(slabs, args.mode, row_offsets=offsets, cols=cols, outs=outs, seed=7)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(3):
        run()
    barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    a.record()
    for _ in range(args.reps):
        run()
    b.record()
    barrier()
    wall = (time.perf_counter() - t0) / args.reps
    ms = torch.tensor([a.elapsed_time(b) / args.reps], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms.item())

    # the exchange step alone
    vec = torch.zeros(len(slabs), dtype=torch.float32, device=dev)
    ar_us = 0.0
    if world > 1:
        for _ in range(5):
            dist.all_reduce(vec, op=dist.ReduceOp.MAX)
        torch.cuda.synchronize()
        a.record()
        for _ in range(50):
            dist.all_reduce(vec, op=dist.ReduceOp.MAX)
        b.record()
        torch.cuda.synchronize()
        ar_us = a.elapsed_time(b) / 50 * 1e3

    if rank == 0:
        gbs = ((2.0 * esz + 1.0) if args.packed else 3.0 * esz) * total / (ms * 1e-3) / 1e9
        print(json.dumps({"workload": "llama7b per-tensor amax-scaled %s%s" % (args.mode, " -> 8-bit codes" if args.packed else ""), "dtype": args.dtype, "n_gpus": world,
                          "tensors": len(shapes), "elements": total, "elements_per_rank": local_elems,
                          "ms_per_sweep_device": ms, "ms_per_sweep_wall": wall * 1e3,
                          "aggregate_GBps_algorithmic": gbs, "per_gpu_GBps": gbs / world,
                          "allreduce_us": ar_us, "launches_per_rank": 3 if args.multi else 2 * len(shapes) + 1,
                          "scaling": "strong"}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
