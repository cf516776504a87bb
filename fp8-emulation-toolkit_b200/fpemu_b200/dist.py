"""Row-slab sharding of the cast across the GPUs of one node.

Every element is independent, so a tensor of R rows is cut into contiguous row slabs, one per rank; the
only exchange step is the per-tensor absolute maximum: each rank reduces its slabs into one float32 vector
(one entry per tensor), ONE all-reduce(MAX) over that vector (NCCL over NVLink: 4*T bytes), then every rank
derives the scales on its own device and casts its slabs.  `elem_offset` = the slab's first global element
index keeps stochastic rounding identical for any world size.

The kernel calls go through a small provider object so that the collective plumbing can be exercised with
any tensor backend; the default provider is the CUDA library (there is no other implementation here)."""
import torch
import torch.distributed as dist

from . import ops as _ops
from . import _lib as L


def slab_bounds(rows, world_size, rank):
    """Rows [r0, r1) owned by `rank`: r0 = floor(rank * rows / world_size)."""
    return (rank * rows) // world_size, ((rank + 1) * rows) // world_size


class CudaKernels(object):
    """Default kernel provider: the C ABI through fpemu_b200.ops."""
    amax = staticmethod(_ops.amax)
    scale_from_amax = staticmethod(_ops.scale_from_amax)
    cast = staticmethod(_ops.cast)
    encode = staticmethod(_ops.encode)


def _as_bits(vec):
    """The vector of maxima as int32 bit patterns: for non-negative floats the integer order is the numeric order and every
    NaN pattern sorts above +Inf, so an integer MAX all-reduce propagates a NaN from any rank (what the kernels' own
    reductions do), which a floating-point MAX does not promise."""
    return vec.view(torch.int32)


def _world(group):
    if dist.is_available() and dist.is_initialized():
        return dist.get_world_size(group), dist.get_rank(group)
    return 1, 0


def global_amax(slabs, group=None, kernels=CudaKernels, out=None):
    """max|x| of each tensor over ALL ranks: local reductions into one vector, one all-reduce(MAX)."""
    if not slabs:
        return out
    dev = slabs[0].device
    vec = torch.zeros(len(slabs), dtype=torch.float32, device=dev) if out is None else out.zero_()
    for t, x in enumerate(slabs):
        if x.numel():
            kernels.amax(x, out=vec[t:t + 1])
    world, _ = _world(group)
    if world > 1:
        dist.all_reduce(_as_bits(vec), op=dist.ReduceOp.MAX, group=group)
    return vec


def _sharded_amax_scaled(slabs, mode, row_offsets, cols, outs, group, seed, flt_max, kernels, pipeline, apply_one):
    fmt, _, _ = L.parse_mode(mode)
    if flt_max is None:
        flt_max = _ops.FLT_MAX[fmt]
    n = len(slabs)
    row_offsets = row_offsets or [0] * n
    cols = cols or [(x.shape[-1] if x.dim() > 1 else 1) for x in slabs]
    world, _ = _world(group)
    vec = torch.zeros(n, dtype=torch.float32, device=slabs[0].device)
    scales = torch.empty(n, dtype=torch.float32, device=slabs[0].device)
    results = [None] * n

    def amax_run(t0, t1):
        for t in range(t0, t1):
            if slabs[t].numel():
                kernels.amax(slabs[t], out=vec[t:t + 1])

    def reduce_run(t0, t1):
        if world > 1:
            return dist.all_reduce(_as_bits(vec[t0:t1]), op=dist.ReduceOp.MAX, group=group, async_op=True)
        return None

    def apply_run(t0, t1):
        scales[t0:t1] = kernels.scale_from_amax(vec[t0:t1], flt_max, dtype=slabs[0].dtype)
        for t in range(t0, t1):
            results[t] = apply_one(t, slabs[t], None if outs is None else outs[t], scales[t:t + 1], seed + t,
                                   row_offsets[t] * cols[t])

    _ops.pipelined_sweep(_ops.split_runs([x.numel() for x in slabs], pipeline if world > 1 else 1),
                         amax_run, reduce_run, apply_run)
    return results, scales


def sharded_amax_scaled_cast(slabs, mode, row_offsets=None, cols=None, outs=None, group=None, seed=0,
                             flt_max=None, kernels=CudaKernels, unscale_fp32=False, pipeline=1):
    """Per-tensor amax-scaled cast of row-sharded tensors.

    slabs        list of this rank's slabs (contiguous; slab t holds rows [row_offsets[t], ...) of tensor t)
    row_offsets  first global row of each slab (default 0: unsharded); cols: row length of each tensor
    outs         optional preallocated outputs (pass the slabs themselves for in-place)
    pipeline     k > 1: the list is cut into k runs whose all-reduces overlap the neighbouring runs' kernels
                 (fpemu_b200.ops.pipelined_sweep); results do not depend on k
    Returns (outs, scales) where scales is the float32 device vector of per-tensor scales."""
    return _sharded_amax_scaled(
        slabs, mode, row_offsets, cols, outs, group, seed, flt_max, kernels, pipeline,
        lambda t, x, out, d_scale, sd, off: kernels.cast(x, mode, out=out, d_scale=d_scale, seed=sd, elem_offset=off,
                                                         unscale_fp32=unscale_fp32))


def sharded_amax_scaled_encode(slabs, mode, row_offsets=None, cols=None, outs=None, group=None, seed=0,
                               flt_max=None, kernels=CudaKernels, pipeline=1):
    """As sharded_amax_scaled_cast, but every rank ends up with the 8-bit codes of its slabs (uint8) plus the common
    float32[T] scale vector — the wire form of a row-sharded weight set.  Same all-reduce(MAX)."""
    return _sharded_amax_scaled(
        slabs, mode, row_offsets, cols, outs, group, seed, flt_max, kernels, pipeline,
        lambda t, x, out, d_scale, sd, off: kernels.encode(x, mode, out=out, d_scale=d_scale, seed=sd, elem_offset=off,
                                                           expect_normals=True))


class PeerExchange(object):
    """The all-reduce(MAX) of a sharded sweep done by the kernels on either side of it (fpemu_amax_multi_push /
    fpemu_cast_multi_xchg): every rank owns an exchange buffer in symmetric memory (torch.distributed._symmetric_memory:
    peer-mapped over NVLink), a one-CTA launch behind the amax launch stores the local maxima, each paired with the
    sweep number in one 8-byte word, into every rank's buffer, and the cast launch's warps poll those words and take
    the maximum over ranks themselves.
    One sweep = amax, push, cast: no collective, no host synchronisation, no memset.  Raises RuntimeError when symmetric memory is not available."""

    def __init__(self, plan, group=None):
        import ctypes
        try:
            import torch.distributed._symmetric_memory as symm_mem
        except Exception as e:  # noqa: BLE001
            raise RuntimeError("fpemu: torch symmetric memory is not available (%s)" % e)
        self.plan = plan
        self.world, self.rank = _world(group)
        if self.world < 2 or self.world > 16:
            raise RuntimeError("fpemu: the peer exchange needs 2..16 ranks")
        grp = group if group is not None else dist.group.WORLD
        nfl = int(L.lib().fpemu_exchange_floats(plan.count, self.world))
        self.buf = symm_mem.empty(nfl, dtype=torch.float32, device=plan.device)
        self.hdl = symm_mem.rendezvous(self.buf, grp.group_name)
        self.buf.zero_()
        torch.cuda.synchronize(plan.device)
        self.hdl.barrier()                                   # every rank's buffer is zero before anybody pushes
        self.ptrs = (ctypes.c_void_p * self.world)(*[int(p) for p in self.hdl.buffer_ptrs])
        self.local = torch.zeros(2 * plan.count, dtype=torch.float32, device=plan.device)
        self.epoch = 0

    def sweep(self, mode, flt_max, seed, flags_extra=0):
        """amax -> exchange -> cast / encode of the plan's tensors; returns (outs, scales).  Every rank must call the
        sweeps of one PeerExchange in the same order."""
        import ctypes
        plan = self.plan
        fmt, rnd, flags = L.parse_mode(mode)
        if flt_max is None:
            flt_max = _ops.FLT_MAX[fmt]
        if rnd == L.RND_STOCHASTIC and seed is None:
            seed = _ops.next_seed()
        self.epoch += 1
        scales = torch.empty(plan.count, dtype=torch.float32, device=plan.device)
        stream = _ops._stream(plan.d_table)
        with _ops._on_device(plan.d_table):
            st = L.lib().fpemu_amax_multi_push(plan.d_table.data_ptr(), plan.h_table.data_ptr(), plan.count,
                                               self.local.data_ptr(),
                                               ctypes.addressof(self.ptrs), self.world, self.rank, self.epoch, stream)
            L.check(st, "fpemu_amax_multi_push")
            fn = L.lib().fpemu_encode_multi_xchg if plan.packed else L.lib().fpemu_cast_multi_xchg
            st = fn(plan.d_table.data_ptr(), plan.h_table.data_ptr(), plan.count, fmt, rnd, flags | flags_extra,
                    self.buf.data_ptr(), self.world, self.epoch, float(flt_max), scales.data_ptr(),
                    int(seed or 0) & 0xFFFFFFFFFFFFFFFF, stream)
            L.check(st, "fpemu_%s_multi_xchg(%s)" % ("encode" if plan.packed else "cast", mode))
        return plan.outs, scales


class ShardedPlan(object):
    """The multi-tensor version of sharded_amax_scaled_cast: the slab list is described to the library once and
    every sweep is then two launches per run of tensors (amax, cast — the cast kernel derives the scales) plus ONE
    all-reduce(MAX) per run over its slice of the float32[T] vector, instead of 2T + 1 launches.  With `pipeline` = k
    the list is cut into k consecutive runs and each run's all-reduce is issued asynchronously right after its amax
    launch, so it executes behind the next run's amax / the previous run's cast instead of idling the GPU between
    two kernels (tools/llama7b_sweep.py measures k = 1, 2, 4)."""

    def __init__(self, slabs, row_offsets=None, cols=None, outs=None, group=None, pipeline=1, exchange="auto"):
        """exchange: "nccl" (one all-reduce(MAX) between the two launches), "peer" (the maxima pushed over NVLink peer
        memory behind the amax launch and awaited inside the cast launch, PeerExchange) or "auto": peer when there is
        more than one rank on NCCL and symmetric memory works here, else nccl.  Measured on the Llama-7B sweep (bf16,
        both exchanges timed alternately in one run, profiles/bench_r2_n*.json): 2.835 against 2.859 ms at 2 GPUs,
        1.439 against 1.460 ms at 4, 0.739 against 0.770 ms at 8."""
        n = len(slabs)
        row_offsets = row_offsets or [0] * n
        cols = cols or [(x.shape[-1] if x.dim() > 1 else 1) for x in slabs]
        self.group = group
        self.pipeline = pipeline
        self.plan = _ops.MultiTensorPlan(slabs, outs, [r * c for r, c in zip(row_offsets, cols)])
        self.peer = None
        world, _ = _world(group)
        if exchange not in ("nccl", "peer", "auto"):
            raise ValueError("exchange must be 'nccl', 'peer' or 'auto'")
        # measured exception: 8-bit codes from fp32 tensors (a purely HBM-bound encode kernel) lose 3-4 % with the
        # in-kernel wait at 2 GPUs (4.27 against 4.12 ms), bf16 codes gain 0.8 %
        slow_case = self.plan.packed and self.plan.dtype == torch.float32
        want_peer = exchange == "peer" or (exchange == "auto" and world >= 2 and self.plan.device.type == "cuda"
                                           and dist.get_backend(group) == "nccl" and not slow_case)
        if want_peer and world > 1:
            try:
                self.peer = PeerExchange(self.plan, group)
            except Exception:  # noqa: BLE001
                if exchange == "peer":
                    raise

    def _allreduce(self, vec):
        world, _ = _world(self.group)
        if world > 1:
            return dist.all_reduce(_as_bits(vec), op=dist.ReduceOp.MAX, group=self.group, async_op=True)
        return None

    def _depth(self):
        world, _ = _world(self.group)
        return self.pipeline if world > 1 else 1

    def amax_scaled_cast(self, mode, seed=0, flt_max=None, unscale_fp32=False):
        if self.peer is not None:
            return self.peer.sweep(mode, flt_max, seed, L.FLAG_UNSCALE_FP32 if unscale_fp32 else 0)
        return self.plan.amax_scaled_cast(mode, flt_max=flt_max, seed=seed, reduce_fn=self._allreduce,
                                          unscale_fp32=unscale_fp32, pipeline=self._depth())

    def cast(self, mode, seed=0):
        return self.plan.cast(mode, seed=seed)

    def amax_scaled_encode(self, mode, seed=0, flt_max=None):
        """As amax_scaled_cast, for a plan whose outputs are uint8 code buffers: every rank ends up with the 8-bit
        codes of its slabs and the same float32[T] scale vector."""
        if self.peer is not None:
            return self.peer.sweep(mode, flt_max, seed, L.FLAG_EXPECT_NORMALS)
        return self.plan.amax_scaled_encode(mode, flt_max=flt_max, seed=seed, reduce_fn=self._allreduce,
                                            pipeline=self._depth())


def sharded_cast(slabs, mode, row_offsets=None, cols=None, outs=None, scale=1.0, seed=0, kernels=CudaKernels):
    """Constant-scale cast of row-sharded tensors: no communication at all."""
    n = len(slabs)
    row_offsets = row_offsets or [0] * n
    cols = cols or [(x.shape[-1] if x.dim() > 1 else 1) for x in slabs]
    return [kernels.cast(x, mode, out=None if outs is None else outs[t], scale=scale, seed=seed + t,
                         elem_offset=row_offsets[t] * cols[t]) for t, x in enumerate(slabs)]
