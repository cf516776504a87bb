"""The exchange step of a row-sharded sweep done by the kernels around it over peer memory (fpemu_amax_multi_push /
fpemu_cast_multi_xchg, fpemu_b200.dist.PeerExchange) against the NCCL all-reduce path: two ranks on two GPUs (skipped on
a one-GPU box; bench.py repeats the check at N > 1 where the driver's scaling run sees it).  Values, 8-bit codes and the
scale vector must be identical bit for bit, sweep after sweep (the buffer halves alternate with the sweep number), with
a rank that holds an empty slab, an odd row length, stochastic rounding and a NaN in one tensor."""
import os
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, q):
    try:
        os.environ["MASTER_ADDR"] = "127.0.0.1"
        os.environ["MASTER_PORT"] = str(port)
        for p in (ROOT, os.path.join(ROOT, "fp8-emulation-toolkit_b200")):
            if p not in sys.path:
                sys.path.insert(0, p)
        torch.cuda.set_device(rank)
        dev = torch.device("cuda", rank)
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
        from fpemu_b200 import dist as fd
        g = torch.Generator().manual_seed(17)
        shapes = [(37, 48), (64, 4096), (1, 64), (300, 1001), (8, 16)]       # 1 row: one rank's slab is empty; 1001: odd
        full = [torch.randn(s, generator=g) * (0.02 * (i + 1)) for i, s in enumerate(shapes)]
        full[4][3, 5] = float("nan")
        for dtype in (torch.float32, torch.bfloat16):
            bounds = [fd.slab_bounds(t.shape[0], world, rank) for t in full]
            slabs = [t[r0:r1].to(dtype).contiguous().to(dev) for t, (r0, r1) in zip(full, bounds)]
            kw = dict(row_offsets=[b[0] for b in bounds], cols=[t.shape[1] for t in full])
            outs_a = [torch.empty_like(x) for x in slabs]
            outs_b = [torch.empty_like(x) for x in slabs]
            a = fd.ShardedPlan(slabs, outs=outs_a, exchange="nccl", **kw)
            b = fd.ShardedPlan(slabs, outs=outs_b, exchange="peer", **kw)
            assert b.peer is not None
            for sweep, mode in enumerate(["E4M3_RNE", "E4M3_STOCHASTIC", "E5M2_RNE", "E3M4_STOCHASTIC", "E4M3_RNE"]):
                _, sa = a.amax_scaled_cast(mode, seed=11 + sweep)
                _, sb = b.amax_scaled_cast(mode, seed=11 + sweep)
                torch.cuda.synchronize()
                assert torch.equal(sa.view(torch.int32), sb.view(torch.int32)), (dtype, mode, sa, sb)
                assert torch.isnan(sb[4]).item() and not torch.isnan(sb[:4]).any().item()
                for x, y in zip(outs_a, outs_b):
                    assert torch.equal(x.view(torch.uint8), y.view(torch.uint8)), (dtype, mode)
            codes_a = [torch.empty(x.shape, dtype=torch.uint8, device=dev) for x in slabs]
            codes_b = [torch.empty(x.shape, dtype=torch.uint8, device=dev) for x in slabs]
            ea = fd.ShardedPlan(slabs[:4], outs=codes_a[:4], exchange="nccl", row_offsets=kw["row_offsets"][:4], cols=kw["cols"][:4])
            eb = fd.ShardedPlan(slabs[:4], outs=codes_b[:4], exchange="peer", row_offsets=kw["row_offsets"][:4], cols=kw["cols"][:4])
            for sweep in range(3):
                _, sa = ea.amax_scaled_encode("E4M3_RNE", seed=5)
                _, sb = eb.amax_scaled_encode("E4M3_RNE", seed=5)
                torch.cuda.synchronize()
                assert torch.equal(sa, sb)
                for x, y in zip(codes_a[:4], codes_b[:4]):
                    assert torch.equal(x, y)
        # many short sweeps back to back (the ranks drift against each other, the buffer halves alternate every sweep,
        # nothing synchronises the host in between): every sweep must reproduce the first one's scales and codes
        want_s, want_c = sb.clone(), [c.clone() for c in codes_b[:4]]
        for sweep in range(400):
            _, sk = eb.amax_scaled_encode("E4M3_RNE", seed=5)
            if sweep % 37 == 0 or sweep == 399:
                torch.cuda.synchronize()
                assert torch.equal(sk, want_s), sweep
                for x, y in zip(codes_b[:4], want_c):
                    assert torch.equal(x, y), sweep
                if rank == sweep % world:
                    torch.cuda._sleep(2000000)             # ~1 ms: push this rank behind its peer
        torch.cuda.synchronize()
        dist.barrier()
        q.put((rank, "ok"))
    except Exception as e:  # noqa: BLE001
        import traceback
        q.put((rank, "FAILED: %s\n%s" % (e, traceback.format_exc())))
    finally:
        try:
            dist.destroy_process_group()
        except Exception:  # noqa: BLE001
            pass


def test_peer_exchange_equals_the_all_reduce_path_on_two_gpus():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import queue
    import time
    world = 2
    port = 29600 + (os.getpid() % 300)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q), daemon=True) for r in range(world)]
    for p in procs:
        p.start()
    results, deadline = [], time.time() + 200
    try:
        # a rank that fails leaves its peer waiting in a collective: report the first failure at once and stop the
        # other rank, instead of waiting for a timeout the peer can never meet
        while len(results) < world and time.time() < deadline:
            try:
                results.append(q.get(timeout=2))
            except queue.Empty:
                if any(p.exitcode not in (None, 0) for p in procs):
                    break
                continue
            if results[-1][1] != "ok":
                break
    finally:
        for p in procs:
            p.join(timeout=20 if len(results) == world and all(m == "ok" for _, m in results) else 0.1)
            if p.is_alive():
                p.kill()                                   # this exact child, by handle
    for rank, msg in results:
        assert msg == "ok", "rank %d: %s" % (rank, msg)
    assert len(results) == world, "ranks finished: %r, exit codes %r" % (results, [p.exitcode for p in procs])
