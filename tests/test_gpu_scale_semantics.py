"""The amax-derived scale, pinned against what the reference EXECUTES on a CUDA device.

`quantize_tensor` (mpemu/qutils.py:264-345) computes `scale = flt_max / vmax` with `vmax` a 0-dim CUDA tensor, i.e.
Tensor.__rtruediv__ = `vmax.reciprocal() * flt_max` in the tensor's dtype, then calls the op once per tensor /
channel / group.  The functions below restate those three branches line by line with torch-on-CUDA doing the scale
arithmetic and the reference's own CUDA kernels (oracle/_ref/fpemu_cuda_ref.so, built unmodified) doing the casts;
this library's single-launch kernels must reproduce the loops bit for bit (default un-scale semantics = the
reference CUDA op's).  A second group of tests pins NaN handling: torch.max propagates NaN, so a NaN anywhere in a
tensor (row) makes its scale, and with it every output of that tensor (row), NaN."""
import numpy as np
import pytest
import torch

from oracle import fpemu_oracle as O
from oracle import ref_loader

pytestmark = pytest.mark.gpu

FLT_MAX = {"E5M2": 57344.0, "E4M3": 448.0, "E3M4": 30.0}


@pytest.fixture(scope="module")
def F():
    import fpemu_b200
    fpemu_b200._lib.lib()
    return fpemu_b200


@pytest.fixture(scope="module")
def ref_cuda():
    if not ref_loader.available("fpemu_cuda_ref"):
        pytest.skip("oracle/_ref/fpemu_cuda_ref.so not built")
    return ref_loader.load("fpemu_cuda_ref")


def same(a, b):
    bits = torch.int32 if a.dtype == torch.float32 else torch.int16
    return bool(((a.view(bits) == b.view(bits)) | (a.isnan() & b.isnan())).all().item())


# ---- the reference's three scaled branches, restated over its own CUDA op (qutils.py:274-287, :289-311, :313-337)
def ref_per_tensor(op, tensor, mode, flt_max):
    vmax = torch.max(abs(torch.flatten(tensor.detach())))
    scale = flt_max / vmax
    scale = 6.55e+04 if scale > 3.275e+04 else scale
    return op.forward(tensor.contiguous(), mode, tensor.nelement(), False, scale, False, 1)[0], float(scale)


def ref_per_channel(op, tensor, mode, flt_max):
    out = torch.zeros_like(tensor)
    scales = []
    for c in range(tensor.shape[0]):
        sub = tensor.select(0, c).detach()
        vmax = torch.max(abs(torch.flatten(sub)))
        scale = flt_max / vmax
        scale = 6.55e+04 if scale > 3.275e+04 else scale
        scales.append(float(scale))
        out.select(0, c).data.copy_(op.forward(sub.contiguous(), mode, sub.nelement(), False, scale, False, 1)[0])
    return out, scales


def ref_fine_grained(op, tensor, mode, flt_max, group_size):
    out = torch.zeros_like(tensor)
    for k in range(tensor.shape[0]):
        crs = tensor.select(0, k).detach().clone()
        chunks = max(1, int(crs.shape[0] / group_size))
        for sub in crs.chunk(chunks, 0):
            vmax = torch.max(abs(torch.flatten(sub)))
            scale = flt_max / vmax
            sub.copy_(op.forward(sub.contiguous(), mode, sub.nelement(), False, scale, False, 1)[0])
        out.select(0, k).data.copy_(crs)
    return out


def _data(shape, seed, dtype):
    g = torch.Generator(device="cuda").manual_seed(seed)
    x = torch.randn(shape, device="cuda", generator=g)
    x = x * torch.exp2(torch.randint(-5, 6, (shape[0],) + (1,) * (len(shape) - 1), device="cuda", generator=g).float())
    return x.to(dtype)


@pytest.mark.parametrize("mode", ["E4M3_RNE", "E5M2_RNE", "E3M4_RNE"])
@pytest.mark.parametrize("dtype", [torch.float32, torch.float16])
def test_per_tensor_branch_equals_the_reference_on_cuda(F, ref_cuda, mode, dtype):
    flt_max = FLT_MAX[mode.split("_")[0]]
    for i, n in enumerate((1000, 4096, 70001, 1 << 20, 6000000)):          # single-launch and three-launch paths
        x = (_data((n,), 100 + i, torch.float32) * (3.0 ** (i - 2))).to(dtype)
        want, ref_scale = ref_per_tensor(ref_cuda, x, mode, flt_max)
        for fused in (True, False):
            got, d_scale = F.cast_amax_scaled(x, mode, return_scale=True, fused=fused)
            assert d_scale.item() == np.float32(ref_scale), (mode, n, fused, d_scale.item(), ref_scale)
            assert same(got, want), (mode, dtype, n, fused)
        assert same(F.cast_amax_scaled_noalloc(x, mode, out=torch.empty_like(x)), want), (mode, dtype, n, "noalloc")
        assert same(F.cast_amax_scaled(x, mode, d_amax=F.amax(x)), want), (mode, dtype, n, "external amax")
    # many maxima: the scale alone, against torch's own `flt_max / vmax` + clamp on this device (for half tensors the
    # comparison with 3.275e4 runs in half, so the values around the clamp threshold are swept densely)
    v = (torch.rand(1500, device="cuda", generator=torch.Generator(device="cuda").manual_seed(7)) * 900 + 1e-3).to(dtype)
    edge = flt_max / 3.275e4
    near = torch.linspace(edge * 0.98, edge * 1.02, 600, device="cuda").to(dtype)
    v = torch.cat([v, near])
    want = []
    for t in v:
        scale = flt_max / t
        scale = 6.55e+04 if scale > 3.275e+04 else scale
        want.append(float(scale))
    got = F.scale_from_amax(v.float().contiguous(), flt_max, dtype=dtype)
    assert got.tolist() == want


@pytest.mark.parametrize("mode", ["E4M3_RNE", "E3M4_RNE", "E5M2_RNE"])
@pytest.mark.parametrize("dtype", [torch.float32, torch.float16])
def test_per_channel_branch_equals_the_reference_loop_on_cuda(F, ref_cuda, mode, dtype):
    flt_max = FLT_MAX[mode.split("_")[0]]
    # register-resident rows, the one-warp-per-row path (rows > 16 384 elements) and the element-wise path (odd length)
    for i, shape in enumerate([(24, 16, 3, 3), (64, 576), (7, 4608), (3, 5), (2, 20000), (5, 1001), (16, 64, 1, 1)]):
        x = _data(shape, 200 + i, dtype)
        want, ref_scales = ref_per_channel(ref_cuda, x, mode, flt_max)
        got, scales = F.cast_rowwise(x, mode, return_scales=True)
        assert scales.tolist() == [float(np.float32(s)) for s in ref_scales], (mode, shape)
        assert same(got, want), (mode, dtype, shape)


@pytest.mark.parametrize("dtype", [torch.float32, torch.float16])
def test_fine_grained_branch_equals_the_reference_loop_on_cuda(F, ref_cuda, dtype):
    for i, (shape, gs) in enumerate([((24, 16, 3, 3), 4), ((8, 64, 1, 1), 16), ((6, 32, 5), 8)]):
        x = _data(shape, 300 + i, dtype)
        want = ref_fine_grained(ref_cuda, x, "E4M3_RNE", 448.0, gs)
        chunks = max(1, int(shape[1] / gs))
        rows = shape[0] * chunks
        got = F.cast_rowwise(x, "E4M3_RNE", clamp=False, rows=rows, cols=x.numel() // rows)
        assert same(got, want), (shape, gs, dtype)


def test_quantize_tensor_scaled_branches_equal_the_reference_loops_on_cuda(F, ref_cuda):
    """The same three pins through the drop-in `mpemu.qutils.quantize_tensor`."""
    from mpemu.qutils import TensorQuantConfig, quantize_tensor
    w = _data((24, 16, 3, 3), 5, torch.float32)
    a = _data((4, 16, 8, 8), 6, torch.float32) * 3.0
    assert same(quantize_tensor(a, TensorQuantConfig("e4m3", "rne", "per-tensor"), None), ref_per_tensor(ref_cuda, a, "E4M3_RNE", 448.0)[0])
    assert same(quantize_tensor(w, TensorQuantConfig("e4m3", "rne", "per-channel"), None), ref_per_channel(ref_cuda, w, "E4M3_RNE", 448.0)[0])
    assert same(quantize_tensor(w, TensorQuantConfig("e3m4", "rne", "per-channel"), None), ref_per_channel(ref_cuda, w, "E3M4_RNE", 30.0)[0])
    assert same(quantize_tensor(w, TensorQuantConfig("e4m3", "rne", "fine-grained", 4), None), ref_fine_grained(ref_cuda, w, "E4M3_RNE", 448.0, 4))


# ---- NaN in the tensor: torch.max(abs(x)) is NaN, the scale is NaN, every output is NaN (qutils.py:283-285)
@pytest.mark.parametrize("dtype", [torch.float32, torch.float16, torch.bfloat16])
@pytest.mark.parametrize("mode", ["E4M3_RNE", "E5M2_RNE"])
def test_nan_in_the_tensor_reaches_the_scale_on_every_amax_path(F, mode, dtype):
    for n, where in ((70001, 12345), (1 << 20, (1 << 20) - 1), (6000000, 0), (1000, 999)):
        x = torch.randn(n, device="cuda").to(dtype)
        x[where] = float("nan")
        assert torch.isnan(F.amax(x)).item()
        assert torch.isnan(F.amax(x[1:]) if where else F.amax(x)).item()
        for fused in (True, False):
            y, s = F.cast_amax_scaled(x, mode, return_scale=True, fused=fused)
            assert torch.isnan(s).item() and torch.isnan(y).all().item(), (mode, dtype, n, fused)
        assert torch.isnan(F.cast_amax_scaled_noalloc(x, mode, out=torch.empty_like(x))).all().item()
    # oracle agrees (float32): amax NaN -> scale NaN -> all NaN
    xn = np.random.default_rng(0).standard_normal(4096).astype(np.float32)
    xn[77] = np.nan
    s = O.scale_from_amax(O.amax(xn), 448.0)
    assert np.isnan(s) and np.isnan(O.cast(xn, "E4M3_RNE", scale=s)).all()
    # a NaN host scale passes like in the reference (`if (scale > 65504)` is false for NaN) and gives all NaN
    assert torch.isnan(F.cast(torch.randn(4096, device="cuda"), mode, scale=float("nan"))).all().item()


def test_nan_rows_and_nan_tensors_of_a_list(F, ref_cuda):
    x = torch.randn(9, 1024, device="cuda")
    x[4, 100] = float("nan")
    for cols in (1024, 1023):                                            # register-resident and element-wise kernels
        xc = x[:, :cols].contiguous()
        y, s = F.cast_rowwise(xc, "E4M3_RNE", return_scales=True)
        assert torch.isnan(s[4]).item() and torch.isnan(y[4]).all().item()
        keep = [r for r in range(9) if r != 4]
        assert not torch.isnan(y[keep]).any().item() and not torch.isnan(s[keep]).any().item()
        assert same(y, ref_per_channel(ref_cuda, xc, "E4M3_RNE", 448.0)[0])
    big = torch.randn(2, 20000, device="cuda")                           # one warp per row
    big[1, 19999] = float("nan")
    y = F.cast_rowwise(big, "E4M3_RNE")
    assert torch.isnan(y[1]).all().item() and not torch.isnan(y[0]).any().item()
    xs = [torch.randn(5000, device="cuda"), torch.randn(33, device="cuda"), torch.randn(70000, device="cuda")]
    xs[1][32] = float("nan")
    plan = F.MultiTensorPlan(xs, [torch.empty_like(t) for t in xs])
    vec = plan.amax()
    assert torch.isnan(vec).tolist() == [False, True, False]
    outs, scales = plan.amax_scaled_cast("E4M3_RNE")
    assert torch.isnan(outs[1]).all().item() and not torch.isnan(outs[0]).any().item() and not torch.isnan(outs[2]).any().item()
    # the blocked mode keeps the reference kernel's fmaxf (NaN ignored, fpemu_kernels.cu:150-170)
    b = torch.randn(64 * 32, device="cuda")
    b[5] = float("nan")
    got = F.cast_blocked(b, "E4M3_RNE", 64)
    want = ref_cuda.forward(b.clone(), "E4M3_RNE", b.numel(), False, 1.0, True, 64)[0]
    assert same(got, want)


@pytest.mark.parametrize("mode", ["E4M3_RNE", "E5M2_STOCHASTIC", "E3M4_RTZ"])
def test_rows_tile_kernel_on_bfloat16_equals_row_by_row_casts(F, mode):
    """bf16 rows that fit no power-of-two group shape (36, 37, 72 or 288 vectors per row) take the whole-rows-per-CTA tile
    kernel: same bits and scales as one amax-scaled cast per row (the per-tensor path, pinned above), row counts that
    leave the last CTA short, a NaN in one row, stochastic rounding keyed by the global element index."""
    for i, (rows, cols) in enumerate([(45, 592), (130, 576), (7, 4608), (29, 1152), (1, 576)]):
        x = _data((rows, cols), 400 + i, torch.float32).to(torch.bfloat16)
        if rows > 3:
            x[3, 5] = float("nan")
        got, scales = F.cast_rowwise(x, mode, seed=9, return_scales=True)
        for r in range(rows):
            want, s = F.cast_amax_scaled(x[r].contiguous(), mode, seed=9, elem_offset=r * cols, return_scale=True, fused=False)
            assert same(got[r], want), (mode, rows, cols, r)
            assert (scales[r].item() == s.item()) or (np.isnan(scales[r].item()) and np.isnan(s.item())), (mode, rows, cols, r)
