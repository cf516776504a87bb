"""float64 tensors (fpemu_cast_f64): the reference dispatches its kernels for double as well
(cuda/fpemu_kernels.cu:1059 ...) — the element is narrowed to float, for the scaled formats AFTER a double-precision
multiplication by the scale (:405, :420), rounded like a float element and widened back.  Bit-exact against
  (a) the oracle applied to the narrowed values (numpy does the double multiply and the narrowing), and
  (b) the reference's own CUDA kernels run on the same double tensor (oracle/_ref/fpemu_cuda_ref.so), when built."""
import numpy as np
import pytest
import torch

from oracle import fpemu_oracle as O
from oracle import ref_loader
from f64_cases import adversarial_f64

pytestmark = pytest.mark.gpu

MODES = ["E5M2_RNE", "E5M2_RTZ", "E4M3_RNE", "E4M3_RTZ", "E3M4_RNE", "BFLOAT16_RNE", "FLOAT16_RNE", "E4M3_IEEE_RNE",
         "FP4_NEAREST", "E5M2_RNAZ"]


@pytest.fixture(scope="module")
def F():
    import fpemu_b200
    fpemu_b200._lib.lib()
    return fpemu_b200


def wide64(n, seed):
    rng = np.random.default_rng(seed)
    x = rng.standard_normal(n) * np.exp2(rng.integers(-20, 18, n))          # full double mantissas
    x[:8] = [0.0, -0.0, np.inf, -np.inf, 448.0000000001, 1e-300, -65504.0000001, 3.5e38]
    return x


def expected(x64, mode, scale):
    O.set_nan_mode(False)
    return O.cast_f64(x64, mode, scale=scale)


@pytest.mark.parametrize("mode", MODES)
def test_double_tensors_vs_oracle(F, mode):
    for n, seed in ((1 << 16, 1), (1000003, 2), (7, 3), (0, 4)):
        x = wide64(max(n, 8), seed)[:n]
        xt = torch.from_numpy(x).cuda()
        for scale in (1.0, 0.125, 3.7):
            got = F.cast(xt, mode, scale=scale)
            assert got.dtype == torch.float64
            want = expected(x, mode, scale)
            ok = O.same_bits_or_both_nan(got.cpu().numpy().astype(np.float32), want.astype(np.float32))
            assert ok.all(), "%s n=%d s=%g: %d mismatches" % (mode, n, scale, (~ok).sum())
            assert np.array_equal(got.cpu().numpy()[ok & ~np.isnan(want)], want[ok & ~np.isnan(want)])


@pytest.mark.parametrize("mode", ["E5M2_RNE", "E4M3_RNE", "E3M4_RNE", "BFLOAT16_RNE", "E5M2_RTZ", "E4M3_IEEE_RNE"])
def test_double_tensors_vs_reference_cuda_kernels(F, mode):
    ref = ref_loader.load("fpemu_cuda_ref")
    if ref is None:
        pytest.skip("oracle/_ref/fpemu_cuda_ref.so not built")
    x = wide64(1 << 16, 9)
    x[2:4] = [1e5, -1e5]                              # keep Inf out: the reference's `scale * Inf` paths agree, NaN payloads need not
    xt = torch.from_numpy(x).cuda()
    for scale in (1.0, 3.7, 94.09):
        if mode.startswith("BFLOAT16") and scale != 1.0:
            continue
        want = ref.forward(xt.clone(), mode, xt.numel(), False, scale, False, 1)[0]
        got = F.cast(xt, mode, scale=scale)
        assert torch.equal(got.view(torch.int64), want.view(torch.int64)), "%s s=%g" % (mode, scale)
        import fpemu_cuda
        again = fpemu_cuda.forward(xt.clone(), mode, xt.numel(), False, scale, False, 1)[0]
        assert torch.equal(again.view(torch.int64), want.view(torch.int64))


def test_double_in_place_views_and_stochastic(F):
    x = torch.from_numpy(wide64(300007, 5)).cuda()
    want = F.cast(x, "E4M3_RNE", scale=3.7)
    y = x.clone()
    F.cast(y, "E4M3_RNE", out=y, scale=3.7)
    assert torch.equal(y.view(torch.int64), want.view(torch.int64))
    v = x[3:]                                          # 8-byte aligned only: element-wise kernel
    assert torch.equal(F.cast(v, "E4M3_RNE", scale=3.7).view(torch.int64), want[3:].view(torch.int64))
    # stochastic: values that are floats already give the float tensor's result (same random field per element index),
    # in one piece or in slabs with any element offset
    xf = (torch.randn(200003, device="cuda") * 3).float()
    xd = xf.double()
    for mode in ("E5M2_STOCHASTIC", "E4M3_STOCHASTIC", "BFLOAT16_STOCHASTIC", "FLOAT16_STOCHASTIC"):
        a = F.cast(xf, mode, seed=77).double()
        b = F.cast(xd, mode, seed=77)
        assert torch.equal(a.view(torch.int64), b.view(torch.int64)), mode
        cut = 1003
        c = torch.cat([F.cast(xd[:cut].clone(), mode, seed=77), F.cast(xd[cut:].clone(), mode, seed=77, elem_offset=cut)])
        assert torch.equal(c.view(torch.int64), b.view(torch.int64)), mode
    with pytest.raises(ValueError):
        F.cast(x, "E4M3_RNE", scale=70000.0)


@pytest.mark.parametrize("mode,dropped", [("E5M2_RNE", 8), ("E4M3_RNE", 7), ("E3M4_RNE", 6)])
def test_double_multiply_semantics_on_adversarial_inputs(F, mode, dropped):
    """Inputs built so that `(float)(x * (double)scale)` and `(float)x * scale` round to different grid points for a large
    share of the elements (tests/f64_cases.py): the kernel must follow the double multiply, like the reference's."""
    scale = 3.7
    x = adversarial_f64(1 << 16, 7, scale, dropped)
    xt = torch.from_numpy(x).cuda()
    got = F.cast(xt, mode, scale=scale)
    want = expected(x, mode, scale)
    assert np.array_equal(got.cpu().numpy(), want)
    narrowed_first = F.cast(xt.float(), mode, scale=scale).double().cpu().numpy()
    if mode != "E3M4_RNE":
        assert (narrowed_first != want).mean() > 0.1          # the distinction is real on this input
    ref = ref_loader.load("fpemu_cuda_ref")
    if ref is not None:
        r = ref.forward(xt.clone(), mode, xt.numel(), False, scale, False, 1)[0]
        assert torch.equal(got.view(torch.int64), r.view(torch.int64))


def test_quantize_tensor_on_double_tensors_equals_the_reference_branch(F):
    """`mpemu.qutils.quantize_tensor` on float64 tensors: the per-tensor `max` branch (qutils.py:282-287) restated over the
    reference's own CUDA op — torch derives `flt_max / vmax` in double on the device, the op's `float scale` argument
    narrows it — and the unscaled cast; bit for bit.  Per-channel modes go through a float copy and stay finite."""
    ref = ref_loader.load("fpemu_cuda_ref")
    if ref is None:
        pytest.skip("oracle/_ref/fpemu_cuda_ref.so not built")
    from mpemu.qutils import TensorQuantConfig, quantize_tensor
    g = torch.Generator(device="cuda").manual_seed(3)
    for i, shape in enumerate([(4, 16, 8, 8), (70001,), (24, 16, 3, 3)]):
        x = torch.randn(shape, device="cuda", dtype=torch.float64, generator=g) * (3.0 ** (i - 1))
        for dt, mode, flt_max in (("e4m3", "E4M3_RNE", 448.0), ("e5m2", "E5M2_RNE", 57344.0), ("e3m4", "E3M4_RNE", 30.0)):
            vmax = torch.max(abs(torch.flatten(x.detach())))
            scale = flt_max / vmax
            scale = 6.55e+04 if scale > 3.275e+04 else scale
            want = ref.forward(x.contiguous(), mode, x.nelement(), False, scale, False, 1)[0]
            got = quantize_tensor(x, TensorQuantConfig(dt, "rne", "per-tensor"), None)
            assert got.dtype == torch.float64 and torch.equal(got.view(torch.int64), want.view(torch.int64)), (shape, mode)
            plain = quantize_tensor(x, TensorQuantConfig(dt, "rne"), None)
            want_plain = ref.forward(x.contiguous(), mode, x.nelement(), False, 1.0, False, 1)[0]
            assert torch.equal(plain.view(torch.int64), want_plain.view(torch.int64)), (shape, mode)
        y = x.clone()
        out = quantize_tensor(y, TensorQuantConfig("e4m3", "rne", "per-tensor"), None, inplace=True)
        assert out.data_ptr() == y.data_ptr() and not torch.equal(y, x)
    w = torch.randn(24, 16, 3, 3, device="cuda", dtype=torch.float64, generator=g)
    q = quantize_tensor(w, TensorQuantConfig("e4m3", "rne", "per-channel"), None)
    assert q.dtype == torch.float64 and torch.isfinite(q).all()
