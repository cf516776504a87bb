"""Host-side logic of the multi-GPU path on CPU: world_size-2 `gloo` process group, the collective plumbing
of fpemu_b200.dist (slab bounds, one all-reduce(MAX) for the global per-tensor amax, element offsets for
partition-independent stochastic rounding) driven with an oracle-backed kernel provider in place of the CUDA
library.  The sharded result must equal the unsharded oracle result bit for bit."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class OracleKernels(object):
    """Kernel provider with the same call shapes as fpemu_b200.dist.CudaKernels, on CPU tensors."""

    @staticmethod
    def amax(x, out=None):
        from oracle import fpemu_oracle as O
        a = O.amax(x.numpy())
        out[0] = max(float(out[0]), a)
        return out

    @staticmethod
    def scale_from_amax(vec, flt_max, dtype=torch.float32):
        from oracle import fpemu_oracle as O
        return torch.tensor([O.scale_from_amax(float(a), flt_max) for a in vec], dtype=torch.float32)

    @staticmethod
    def cast(x, mode, out=None, scale=1.0, d_scale=None, seed=0, elem_offset=0, unscale_fp32=False):
        from oracle import fpemu_oracle as O
        s = float(d_scale[0]) if d_scale is not None else scale
        y = torch.from_numpy(O.cast(x.numpy(), mode, scale=s, seed=seed, elem_offset=elem_offset, unscale_fp32=unscale_fp32))
        y = y.reshape(x.shape)
        if out is not None:
            out.copy_(y)
            return out
        return y


    @staticmethod
    def encode(x, mode, out=None, scale=1.0, d_scale=None, seed=0, elem_offset=0, expect_normals=False):
        from oracle import fpemu_oracle as O
        s = float(d_scale[0]) if d_scale is not None else scale
        c = torch.from_numpy(O.encode(x.numpy(), mode, scale=s, seed=seed, elem_offset=elem_offset)).reshape(x.shape)
        if out is not None:
            out.copy_(c)
            return out
        return c


def _tensors():
    g = torch.Generator().manual_seed(17)
    shapes = [(37, 16), (64, 48), (5, 32), (1, 16)]          # 1 row: one rank gets an empty slab
    return [torch.randn(s, generator=g) * (0.02 * (i + 1)) for i, s in enumerate(shapes)]


def _worker(rank, world, port, mode, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    for p in (ROOT, os.path.join(ROOT, "fp8-emulation-toolkit_b200")):
        if p not in sys.path:
            sys.path.insert(0, p)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from fpemu_b200 import dist as fd
        full = _tensors()
        bounds = [fd.slab_bounds(t.shape[0], world, rank) for t in full]
        slabs = [t[r0:r1].contiguous() for t, (r0, r1) in zip(full, bounds)]
        outs, scales = fd.sharded_amax_scaled_cast(slabs, mode, row_offsets=[b[0] for b in bounds],
                                                   cols=[t.shape[1] for t in full], seed=11, kernels=OracleKernels)
        # the pipelined schedule (runs of tensors, asynchronous all-reduce per run) must not change a bit
        for depth in (2, 3, 8):
            outs_p, scales_p = fd.sharded_amax_scaled_cast(slabs, mode, row_offsets=[b[0] for b in bounds],
                                                           cols=[t.shape[1] for t in full], seed=11,
                                                           kernels=OracleKernels, pipeline=depth)
            assert torch.equal(scales_p, scales)
            for a, b in zip(outs, outs_p):
                assert torch.equal(a.view(torch.int32), b.view(torch.int32))
        plain = fd.sharded_cast(slabs, "E5M2_STOCHASTIC", row_offsets=[b[0] for b in bounds],
                                cols=[t.shape[1] for t in full], seed=5, kernels=OracleKernels)
        codes, cscales = fd.sharded_amax_scaled_encode(slabs, mode, row_offsets=[b[0] for b in bounds],
                                                       cols=[t.shape[1] for t in full], seed=11, kernels=OracleKernels,
                                                       pipeline=2)
        assert torch.equal(cscales, scales)
        q.put((rank, [o.numpy() for o in outs], scales.numpy(), [p.numpy() for p in plain], bounds,
               [c.numpy() for c in codes]))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("mode", ["E4M3_RNE", "E4M3_STOCHASTIC"])
def test_two_rank_sharded_cast_equals_unsharded(mode):
    from oracle import fpemu_oracle as O
    world = 2
    port = 29500 + (os.getpid() % 2000)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, mode, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = sorted([q.get(timeout=180) for _ in range(world)], key=lambda r: r[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    full = _tensors()
    for t, x in enumerate(full):
        s = O.scale_from_amax(O.amax(x.numpy()), 448.0)                   # GLOBAL amax, not the slab's
        want = O.cast(x.numpy().reshape(-1), mode, scale=s, seed=11 + t).reshape(x.shape)
        got = np.concatenate([r[1][t] for r in results], axis=0)
        assert O.same_bits_or_both_nan(got, want).all(), "tensor %d" % t
        for r in results:
            assert r[2][t] == np.float32(s)                               # every rank derived the same scale
        want_codes = O.encode(x.numpy().reshape(-1), mode, scale=s, seed=11 + t).reshape(x.shape)
        assert (np.concatenate([r[5][t] for r in results], axis=0) == want_codes).all(), "codes of tensor %d" % t
        want_plain = O.cast(x.numpy().reshape(-1), "E5M2_STOCHASTIC", seed=5 + t).reshape(x.shape)
        got_plain = np.concatenate([r[3][t] for r in results], axis=0)
        assert O.same_bits_or_both_nan(got_plain, want_plain).all()


def test_slab_bounds_cover_all_rows_exactly_once():
    sys.path.insert(0, os.path.join(ROOT, "fp8-emulation-toolkit_b200"))
    from fpemu_b200.dist import slab_bounds
    for rows in (0, 1, 7, 8, 4096, 11008, 32000):
        for world in (1, 2, 3, 4, 8):
            cuts = [slab_bounds(rows, world, r) for r in range(world)]
            assert cuts[0][0] == 0 and cuts[-1][1] == rows
            for a, b in zip(cuts, cuts[1:]):
                assert a[1] == b[0]
            sizes = [b - a for a, b in cuts]
            assert max(sizes) - min(sizes) <= 1


def test_run_splitting_and_sweep_order():
    sys.path.insert(0, os.path.join(ROOT, "fp8-emulation-toolkit_b200"))
    from fpemu_b200.ops import pipelined_sweep, split_runs
    assert split_runs([10, 10, 10, 10], 2) == [(0, 2), (2, 4)]
    assert split_runs([1, 1, 1, 100], 3) == [(0, 2), (2, 3), (3, 4)]
    assert split_runs([5], 4) == [(0, 1)]
    rng = np.random.default_rng(0)
    for _ in range(500):
        sizes = [int(v) for v in rng.choice([0, 1, 5, 100], size=rng.integers(1, 12))]
        depth = int(rng.integers(1, 7))
        runs = split_runs(sizes, depth)
        assert len(runs) == min(depth, len(sizes)) and runs[0][0] == 0 and runs[-1][1] == len(sizes)
        assert all(a[1] == b[0] for a, b in zip(runs, runs[1:])) and all(b > a for a, b in runs)
    log = []

    class Work(object):
        def __init__(self, tag):
            self.tag = tag

        def wait(self):
            log.append(("wait",) + self.tag)
    pipelined_sweep([(0, 2), (2, 3)], lambda a, b: log.append(("amax", a, b)),
                    lambda a, b: log.append(("reduce", a, b)) or Work((a, b)), lambda a, b: log.append(("apply", a, b)))
    assert log == [("amax", 0, 2), ("reduce", 0, 2), ("amax", 2, 3), ("reduce", 2, 3),
                   ("wait", 0, 2), ("apply", 0, 2), ("wait", 2, 3), ("apply", 2, 3)]
