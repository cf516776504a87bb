#!/usr/bin/env python
"""BASELINE.json configs[4]: Llama-style 7B random-init weight set, per-tensor amax-scaled E4M3 QDQ, row-sharded
across the GPUs of one node with one NCCL all-reduce(MAX) per run of tensors over its slice of the float[291] amax vector.

    python tools/llama7b_sweep.py [--dtype bf16|fp32] [--reps 5] [--multi] [--pipeline K]       # 1 GPU
    python -m torch.distributed.run --nproc-per-node N ... tools/llama7b_sweep.py ...          # N GPUs (strong scaling)

The measurement itself is bench.py's `c5_sweep` (the driver-run bench reports the same numbers under detail.c5); this
script sweeps its options: per-tensor launches vs multi-tensor launches, pipeline depth, values vs 8-bit codes."""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "fp8-emulation-toolkit_b200")):
    sys.path.insert(0, p)
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--dtype", default="bf16", choices=["bf16", "fp32", "fp16"])
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--mode", default="E4M3_RNE")
    ap.add_argument("--multi", action="store_true", help="multi-tensor launches instead of 2 x 291 + 1 per-tensor launches")
    ap.add_argument("--packed", action="store_true", help="write 8-bit codes instead of de-quantised values (implies --multi)")
    ap.add_argument("--exchange", default="auto", choices=["auto", "nccl", "peer"])
    ap.add_argument("--pipeline", default="1", help="runs of tensors per sweep, comma-separated list to sweep (e.g. 1,2,4)")
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    for depth in [int(v) for v in args.pipeline.split(",")]:
        res = bench.c5_sweep(dev, world, rank, args.dtype, args.reps, depth, mode=args.mode, packed=args.packed,
                             multi=args.multi or args.packed, exchange=args.exchange)
        if rank == 0:
            print(json.dumps(res), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
