"""In-kernel cost of the peer exchange without any peer: ONE GPU plays a world of one (its own buffer is the only
"peer"), so fpemu_amax_multi_push + fpemu_cast_multi_xchg do everything they do at N > 1 except cross NVLink, and can be
compared, alternating, with fpemu_amax_multi + fpemu_cast_multi_amax on the same slabs (rank 0's slabs of a 2-way shard
of the Llama-7B weight set).  Prints one JSON line."""
import argparse
import ctypes
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "fp8-emulation-toolkit_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=10)
    ap.add_argument("--rounds", type=int, default=8)
    ap.add_argument("--once", action="store_true", help="one launch of each kernel and exit (for a run under ncu)")
    args = ap.parse_args()
    import torch
    import bench
    from fpemu_b200 import dist as fd, ops, _lib as L
    dev = torch.device("cuda", 0)
    slabs = []
    for t, (_, r, c) in enumerate(bench.llama7b_shapes()):
        r0, r1 = fd.slab_bounds(r, 2, 0)
        slabs.append((torch.randn((r1 - r0) * c, device=dev) * 0.02).to(torch.bfloat16).view(r1 - r0, c))
    outs = [torch.empty_like(x) for x in slabs]
    plan = ops.MultiTensorPlan(slabs, outs)
    lib = L.lib()
    buf = torch.zeros(int(lib.fpemu_exchange_floats(plan.count, 1)), dtype=torch.float32, device=dev)
    ptrs = (ctypes.c_void_p * 1)(buf.data_ptr())
    local = torch.zeros(2 * plan.count, dtype=torch.float32, device=dev)
    vec = torch.zeros(plan.count, dtype=torch.float32, device=dev)
    scales = torch.empty(plan.count, dtype=torch.float32, device=dev)
    stream = ops._stream(plan.d_table)
    tab = (plan.d_table.data_ptr(), plan.h_table.data_ptr(), plan.count)
    epoch = [0]

    def plain():
        vec.zero_()
        plan.amax(out=vec)
        plan._from_amax("E4M3_RNE", vec, 448.0, scales, 0)

    def xchg():
        epoch[0] += 1
        L.check(lib.fpemu_amax_multi_push(*tab, local.data_ptr(), ctypes.addressof(ptrs), 1, 0, epoch[0], stream), "push")
        L.check(lib.fpemu_cast_multi_xchg(*tab, L.FMT_E4M3, L.RND_RNE, 0, buf.data_ptr(), 1, epoch[0], 448.0,
                                          scales.data_ptr(), 0, stream), "xchg")

    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def timed(fn):
        torch.cuda.synchronize()
        a.record()
        for _ in range(args.reps):
            fn()
        b.record()
        torch.cuda.synchronize()
        return round(a.elapsed_time(b) / args.reps, 4)

    def amax_only():
        vec.zero_()
        plan.amax(out=vec)

    def push_only():
        epoch[0] += 1
        L.check(lib.fpemu_amax_multi_push(*tab, local.data_ptr(), ctypes.addressof(ptrs), 1, 0, epoch[0], stream), "push")

    def cast_plain():
        plan._from_amax("E4M3_RNE", vec, 448.0, scales, 0)

    def cast_xchg():                                        # the buffer already holds the words of epoch[0]
        L.check(lib.fpemu_cast_multi_xchg(*tab, L.FMT_E4M3, L.RND_RNE, 0, buf.data_ptr(), 1, epoch[0], 448.0,
                                          scales.data_ptr(), 0, stream), "xchg")

    if args.once:
        amax_only()
        push_only()
        cast_plain()
        cast_xchg()
        torch.cuda.synchronize()
        return
    for _ in range(20):                                    # reach the steady power state before comparing
        plain()
    res = {}
    for _ in range(args.rounds):
        res.setdefault("plain", []).append(timed(plain))
        res.setdefault("xchg", []).append(timed(xchg))
    vec.zero_()
    plan.amax(out=vec)
    for _ in range(args.rounds):
        res.setdefault("amax_only", []).append(timed(amax_only))
        res.setdefault("amax_push_only", []).append(timed(push_only))
        res.setdefault("cast_plain_only", []).append(timed(cast_plain))
        res.setdefault("cast_xchg_only", []).append(timed(cast_xchg))
    want = scales.clone()
    plain()
    torch.cuda.synchronize()
    assert torch.equal(want, scales)
    med = {k: sorted(v)[len(v) // 2] for k, v in res.items()}
    print(json.dumps({"ms_per_sweep_median": med, "rounds": res}))


if __name__ == "__main__":
    main()
