#!/usr/bin/env python
"""BASELINE configs[2] and [3] on one GPU through this repo's `mpemu` drop-in: time per forward / training step with
and without emulation, number of cast launches, time spent inside the cast kernels.

    python tools/module_bench.py [--resnet-batch 64] [--bert-batch 12 --bert-seq 384]"""
import argparse
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "fp8-emulation-toolkit_b200")):
    sys.path.insert(0, p)
from mpemu import mpt_emu  # noqa: E402


def timed(fn, reps=5, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps * 1e3


def kernel_time_ms(fn):
    """Sum of device time of this library's kernels (names contain 'fpemu') and of all kernels, one call."""
    from torch.profiler import ProfilerActivity, profile
    fn()
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        fn()
        torch.cuda.synchronize()
    ours = total = 0.0
    n_ours = 0
    for e in prof.events():
        if e.device_type == torch.autograd.DeviceType.CUDA:
            total += e.device_time
            if "fpemu" in e.name:
                ours += e.device_time
                n_ours += 1
    return ours / 1e3, total / 1e3, n_ours


def kernel_breakdown(fn, top=14):
    """Device time by kernel name for one call: [(name, ms, launches)], longest first."""
    from torch.profiler import ProfilerActivity, profile
    fn()
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        fn()
        torch.cuda.synchronize()
    acc = {}
    for e in prof.events():
        if e.device_type == torch.autograd.DeviceType.CUDA:
            t = acc.setdefault(e.name[:90], [0.0, 0])
            t[0] += e.device_time / 1e3
            t[1] += 1
    rows = sorted(((k, round(v[0], 3), v[1]) for k, v in acc.items()), key=lambda r: -r[1])
    return rows[:top], round(sum(v[0] for v in acc.values()), 3), sum(v[1] for v in acc.values())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--breakdown", action="store_true", help="also list device time by kernel name")
    ap.add_argument("--resnet-batch", type=int, default=64)
    ap.add_argument("--bert-batch", type=int, default=12)
    ap.add_argument("--bert-seq", type=int, default=384)
    args = ap.parse_args()
    out = {}
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False

    import torchvision
    torch.manual_seed(0)
    rn = torchvision.models.resnet50(weights=None).eval().cuda()
    x = torch.randn(args.resnet_batch, 3, 224, 224, device="cuda")
    with torch.no_grad():
        base = timed(lambda: rn(x))
        t0 = time.perf_counter()
        rn, emu = mpt_emu.quantize_model(rn, dtype="e4m3", list_exempt_layers=["conv1", "bn1", "fc"], device="cuda")
        torch.cuda.synchronize()
        prep = (time.perf_counter() - t0) * 1e3
        # the first call in a process also pays CUDA's lazy loading of the kernels it touches: time a second model too
        torch.manual_seed(0)
        rn2 = torchvision.models.resnet50(weights=None).eval().cuda()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        mpt_emu.quantize_model(rn2, dtype="e4m3", list_exempt_layers=["conv1", "bn1", "fc"], device="cuda")
        torch.cuda.synchronize()
        prep_warm = (time.perf_counter() - t0) * 1e3
        del rn2
        emul = timed(lambda: rn(x))
        ours, total, n = kernel_time_ms(lambda: rn(x))
    out["resnet50_ptq_e4m3"] = {"batch": args.resnet_batch, "fp32_forward_ms": base, "emulated_forward_ms": emul,
                                "quantize_model_ms (53 per-channel weight casts + hooks), first call in the process": prep,
                                "quantize_model_ms, second model": prep_warm,
                                "cast_kernel_ms_per_forward": ours, "all_kernel_ms_per_forward": total, "cast_launches_per_forward": n}

    from transformers import BertConfig, BertForQuestionAnswering
    torch.manual_seed(0)
    cfg = BertConfig()
    bert = BertForQuestionAnswering(cfg).train().cuda()
    opt = torch.optim.SGD(bert.parameters(), lr=1e-3)
    ids = torch.randint(0, cfg.vocab_size, (args.bert_batch, args.bert_seq), device="cuda")
    sp = torch.randint(0, args.bert_seq, (args.bert_batch,), device="cuda")
    ep = torch.randint(0, args.bert_seq, (args.bert_batch,), device="cuda")

    def step():
        opt.zero_grad(set_to_none=True)
        loss = bert(input_ids=ids, start_positions=sp, end_positions=ep).loss
        loss.backward()
        opt.step()
    base = timed(step)
    if args.breakdown:
        out["bert_plain_breakdown"] = kernel_breakdown(step)
    bert, emu = mpt_emu.initialize(bert, None, training_algo="hybrid", list_exempt_layers=["qa_outputs"], device="cuda")

    def estep():
        opt.zero_grad(set_to_none=True)
        loss = bert(input_ids=ids, start_positions=sp, end_positions=ep).loss
        loss.backward()
        emu.optimizer_step(opt)
    emul = timed(estep)
    ours, total, n = kernel_time_ms(estep)
    if args.breakdown:
        out["bert_emulated_breakdown"] = kernel_breakdown(estep)
    from mpemu import qutils
    qutils.set_output_cast_out_of_place(True)
    emul_oop = timed(estep)
    ours_oop, total_oop, n_oop = kernel_time_ms(estep)
    qutils.set_output_cast_out_of_place(False)
    out["bert_base_hybrid_step_out_of_place_outputs"] = {"emulated_step_ms": emul_oop, "cast_kernel_ms_per_step": ours_oop,
                                                         "all_kernel_ms_per_step": total_oop, "cast_launches_per_step": n_oop}
    # the whole emulated training step as ONE CUDA graph: the library neither synchronises nor allocates, the hooks are
    # sync-free, and the device seed word gives the stochastic gradient casts fresh bits on every replay
    try:
        import fpemu_b200 as F
        qutils.set_output_cast_out_of_place(True)
        F.enable_device_seed(value=1)
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(3):
                estep()
                F.advance_device_seed()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        opt.zero_grad(set_to_none=True)
        with torch.cuda.graph(graph):
            loss = bert(input_ids=ids, start_positions=sp, end_positions=ep).loss
            loss.backward()
            emu.optimizer_step(opt)
            F.advance_device_seed()
        losses = []
        for _ in range(3):
            graph.replay()
            losses.append(float(loss))
        graph_ms = timed(graph.replay)
        out["bert_base_hybrid_step_one_cuda_graph"] = {"replay_ms": graph_ms, "losses_of_3_replays": losses}
    except Exception as e:  # noqa: BLE001
        out["bert_base_hybrid_step_one_cuda_graph"] = {"failed": "%s: %s" % (type(e).__name__, str(e)[:300])}
    finally:
        qutils.set_output_cast_out_of_place(False)
        try:
            F.disable_device_seed()
        except Exception:  # noqa: BLE001
            pass
    out["bert_base_hybrid_step"] = {"batch": args.bert_batch, "seq": args.bert_seq, "fp32_step_ms": base, "emulated_step_ms": emul,
                                    "cast_kernel_ms_per_step": ours, "all_kernel_ms_per_step": total, "cast_launches_per_step": n}
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
