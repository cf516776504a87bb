"""GPU parity tests proper: the CUDA path, called through the C ABI (fpemu_b200.ops -> ctypes ->
libfpemu_b200.so), against
  (a) the C oracle (oracle/fpemu_oracle.c) on the same seeded inputs,
  (b) the committed golden vectors the reference CPU op produced (tests/golden/),
  (c) the reference's own CUDA kernels compiled unmodified for sm_100a (oracle/_ref/fpemu_cuda_ref.so),
bit-exact for every deterministic mode (NaN compared as NaN: payloads are ISA-specific)."""
import os

import numpy as np
import pytest
import torch

from oracle import fpemu_oracle as O
from oracle import ref_loader

pytestmark = pytest.mark.gpu

DET_MODES = ["E5M2_RTZ", "E5M2_RNE", "E5M2_RNAZ", "E5M2_RNTZ", "E5M2_RPINF", "E5M2_RNINF",
             "E5M2_DAZ_RNE", "E5M2_DAZ_RNAZ", "E5M2_DAZ_RNTZ", "E4M3_RNE", "E3M4_RNE",
             "FLOAT16_RNE", "FLOAT16_DAZ_RNE", "BFLOAT16_RNE", "E4M3_RTZ", "E3M4_RTZ", "BFLOAT16_RTZ",
             "E4M3_IEEE_RNE", "FP4_NEAREST"]
GOLDEN_MODES = ["E5M2_RTZ", "E5M2_RNE", "E5M2_RNAZ", "E5M2_RNTZ", "E5M2_RPINF", "E5M2_RNINF",
                "E5M2_DAZ_RNE", "E5M2_DAZ_RNAZ", "E5M2_DAZ_RNTZ", "E4M3_RNE", "E3M4_RNE",
                "FLOAT16_RNE", "FLOAT16_DAZ_RNE"]
SR_MODES = ["E5M2_STOCHASTIC", "E5M2_DAZ_STOCHASTIC", "E4M3_STOCHASTIC", "E3M4_STOCHASTIC", "E4M3_IEEE_STOCHASTIC",
            "BFLOAT16_STOCHASTIC", "FLOAT16_STOCHASTIC"]
SCALES = [1.0, 0.125, 3.7, 94.09]


@pytest.fixture(scope="module")
def F():
    import fpemu_b200
    fpemu_b200._lib.lib()          # raises if the CUDA library is missing: no fallback
    return fpemu_b200


@pytest.fixture(autouse=True)
def _cuda_nan_mode():
    O.set_nan_mode(False)
    yield


def dev(x):
    return torch.from_numpy(np.ascontiguousarray(x)).cuda()


def host(t):
    if t.dtype == torch.bfloat16:
        return t.view(torch.int16).cpu().numpy().view(np.uint16)
    return t.cpu().numpy()


def assert_same(got, want, what):
    ok = O.same_bits_or_both_nan(got, want)
    if not ok.all():
        bad = np.nonzero(~ok)[0]
        i = bad[0]
        raise AssertionError("%s: %d / %d mismatches, first at %d: got %r want %r" %
                             (what, len(bad), ok.size, i, got.flat[i], want.flat[i]))


def fp16_patterns():
    return np.arange(65536, dtype=np.uint16).view(np.float16).astype(np.float32)


def wide(n=1 << 16, seed=11):
    rng = np.random.default_rng(seed)
    x = rng.standard_normal(n).astype(np.float32) * np.exp2(rng.integers(-30, 20, n)).astype(np.float32)
    spec = np.array([0.0, -0.0, np.inf, -np.inf, np.nan, -np.nan, 1e-40, -1e-40, 65504, 65520, -65520, 3.4e38, 1e-45,
                     448, 464, 480, -448, 30, 31, -31, 57344, 61440, -61440], dtype=np.float32)
    k = min(n, len(spec))
    x[:k] = spec[:k]
    return x


# ------------------------------------------------------------------------------------------ vs oracle
@pytest.mark.parametrize("mode", DET_MODES)
def test_exhaustive_fp16_patterns_vs_oracle(F, mode):
    x = fp16_patterns()
    got = host(F.cast(dev(x), mode))
    assert_same(got, O.cast(x, mode), mode)


@pytest.mark.parametrize("mode", DET_MODES)
@pytest.mark.parametrize("scale", SCALES)
def test_wide_vs_oracle_both_unscale_semantics(F, mode, scale):
    if mode.startswith(("FLOAT16", "BFLOAT16")) and scale != 1.0:
        pytest.skip("format takes no scale")
    x = wide()
    for fp32_unscale in (False, True):
        got = host(F.cast(dev(x), mode, scale=scale, unscale_fp32=fp32_unscale))
        assert_same(got, O.cast(x, mode, scale=scale, unscale_fp32=fp32_unscale), "%s s=%g fp32=%s" % (mode, scale, fp32_unscale))


@pytest.mark.parametrize("mode", DET_MODES)
@pytest.mark.parametrize("dtype", ["f16", "bf16"])
def test_half_and_bfloat_tensors_vs_oracle(F, mode, dtype):
    rng = np.random.default_rng(5)
    if dtype == "f16":
        xb = rng.integers(0, 65536, 1 << 16).astype(np.uint16)
        xb[:65536] = np.arange(65536, dtype=np.uint16)           # every fp16 pattern
        x_np = xb.view(np.float16)
        xt = dev(x_np)
    else:
        x_np = np.arange(65536, dtype=np.uint16)                 # every bf16 pattern (uint16 = bf16 bits)
        xt = dev(x_np.view(np.int16)).view(torch.bfloat16)
    for scale in (1.0, 3.7):
        if mode.startswith(("FLOAT16", "BFLOAT16")) and scale != 1.0:
            continue
        got = host(F.cast(xt, mode, scale=scale))
        assert_same(got, O.cast(x_np, mode, scale=scale), "%s %s s=%g" % (mode, dtype, scale))


# ------------------------------------------------------------------------------------------ vs golden
@pytest.mark.parametrize("mode", GOLDEN_MODES)
def test_golden_fp16_patterns(F, golden_dir, mode):
    g = np.load(os.path.join(golden_dir, "fp16_patterns.npz"))
    x = fp16_patterns()
    want = g[mode].view(np.float16).astype(np.float32)
    got = host(F.cast(dev(x), mode, unscale_fp32=True))
    ok = O.same_bits_or_both_nan(got, want)
    ok |= np.isnan(x) & np.signbit(x)        # -NaN: x86 keeps the sign through float->half, CUDA does not
    assert ok.all(), "%s: %d mismatches" % (mode, (~ok).sum())


@pytest.mark.parametrize("mode", GOLDEN_MODES)
def test_golden_wide_all_scales_cpu_unscale(F, golden_dir, mode):
    g = np.load(os.path.join(golden_dir, "wide.npz"))
    x = g["input"].view(np.float32)
    xt = dev(x)
    for key in g.files:
        if not key.startswith(mode + "@"):
            continue
        scale = float(key.split("@")[1])
        got = host(F.cast(xt, mode, scale=scale, unscale_fp32=True))
        ok = O.same_bits_or_both_nan(got, g[key].view(np.float32))
        ok |= np.isnan(x) & np.signbit(x)
        assert ok.all(), "%s s=%g: %d mismatches" % (mode, scale, (~ok).sum())


def test_golden_bf16(F, golden_dir):
    g = np.load(os.path.join(golden_dir, "wide.npz"))
    if "BFLOAT16_RNE@1" not in g.files:
        pytest.skip("no bf16 fixture")
    x = g["bf16_input"].view(np.float32)
    got = host(F.cast(dev(x), "BFLOAT16_RNE"))
    ok = O.same_bits_or_both_nan(got, g["BFLOAT16_RNE@1"].view(np.float32)) | (x == 0)
    assert ok.all()


@pytest.mark.parametrize("mode", ["E5M2_RNE", "E4M3_RNE", "E3M4_RNE", "E5M2_RTZ"])
@pytest.mark.parametrize("bs", [32, 64, 1024])
def test_golden_blocked(F, golden_dir, mode, bs):
    g = np.load(os.path.join(golden_dir, "blocked.npz"))
    x = g["input"].view(np.float32)
    got = host(F.cast_blocked(dev(x), mode, bs, unscale_fp32=True))
    assert_same(got, g["%s@%d" % (mode, bs)].view(np.float32), "%s bs=%d" % (mode, bs))


def test_golden_reference_64_value_fixture(F, golden_dir):
    """The reference's own 64-value test vector (mpemu/pytquant/test.py:120-135): CUDA path vs the outputs its CPU op
    produced (fp32 un-scale), every deterministic mode, scales 1 and 3.7, BFLOAT16, 16-element blocks."""
    g = np.load(os.path.join(golden_dir, "fixture64.npz"))
    x = g["input"].view(np.float32)
    xt = dev(x)
    for key in g.files:
        if key == "input":
            continue
        mode, arg = key.split("@")
        if arg.startswith("block"):
            got = F.cast_blocked(xt, mode, int(arg[5:]), unscale_fp32=True)
        else:
            got = F.cast(xt, mode, scale=float(arg), unscale_fp32=True)
        assert_same(host(got), g[key].view(np.float32), key)


# ------------------------------------------------------------------------------------------ vs reference CUDA kernels
REF_CUDA_MODES = ["E5M2_RTZ", "E5M2_RNE", "E5M2_RNAZ", "E5M2_RNTZ", "E5M2_RPINF", "E5M2_RNINF",
                  "E5M2_DAZ_RNE", "E5M2_DAZ_RNAZ", "E5M2_DAZ_RNTZ", "E4M3_RNE", "E3M4_RNE", "E4M3_IEEE_RNE",
                  "FLOAT16_RNE", "FLOAT16_DAZ_RNE", "BFLOAT16_RNE", "FP4_NEAREST"]


@pytest.fixture(scope="module")
def ref_cuda():
    if not ref_loader.available("fpemu_cuda_ref"):
        pytest.skip("oracle/_ref/fpemu_cuda_ref.so not built")
    return ref_loader.load("fpemu_cuda_ref")


@pytest.mark.parametrize("mode", REF_CUDA_MODES)
def test_bit_exact_vs_reference_cuda_kernels(F, ref_cuda, mode):
    for name, x in (("fp16pat", fp16_patterns()), ("wide", wide())):
        xt = dev(x)
        for scale in SCALES:
            if mode.startswith(("FLOAT16", "BFLOAT16")) and scale != 1.0:
                continue
            want = ref_cuda.forward(xt.clone(), mode, xt.numel(), False, scale, False, 1)[0]
            got = F.cast(xt, mode, scale=scale)
            assert_same(host(got), host(want), "%s %s s=%g" % (mode, name, scale))


@pytest.mark.parametrize("mode", ["E5M2_RNE", "E4M3_RNE", "E3M4_RNE", "E5M2_DAZ_RNE", "FLOAT16_RNE", "BFLOAT16_RNE"])
def test_bit_exact_vs_reference_cuda_kernels_half_tensors(F, ref_cuda, mode):
    xt = dev(np.arange(65536, dtype=np.uint16).view(np.float16))
    for scale in (1.0, 0.125, 3.7):
        if mode.startswith(("FLOAT16", "BFLOAT16")) and scale != 1.0:
            continue
        want = ref_cuda.forward(xt.clone(), mode, xt.numel(), False, scale, False, 1)[0]
        got = F.cast(xt, mode, scale=scale)
        assert_same(host(got), host(want), "%s half s=%g" % (mode, scale))


@pytest.mark.parametrize("mode", ["E5M2_RNE", "E4M3_RNE", "E3M4_RNE"])
@pytest.mark.parametrize("bs", [1, 4, 16, 32, 64, 256, 1024])
def test_blocked_vs_reference_cuda_kernels(F, ref_cuda, mode, bs):
    # power-of-two block sizes only: the reference's smem tree over-reads otherwise
    rng = np.random.default_rng(2)
    x = (rng.standard_normal(1 << 16).astype(np.float32) * np.repeat(np.exp2(rng.integers(-6, 6, 64)), 1024).astype(np.float32))
    xt = dev(x)
    want = ref_cuda.forward(xt.clone(), mode, xt.numel(), False, 1.0, True, bs)[0]
    got = F.cast_blocked(xt, mode, bs)
    assert_same(host(got), host(want), "%s bs=%d" % (mode, bs))


@pytest.mark.parametrize("mode", ["E5M2_RNE", "E4M3_RNE", "E3M4_RNE"])
@pytest.mark.parametrize("bs", [32, 256, 1024])
def test_blocked_half_tensors_vs_reference_cuda_kernels(F, ref_cuda, mode, bs):
    """fp16 tensors: the E5M2 kernel keeps its block scale in half precision (`scalar_t scale`, fpemu_kernels.cu:405-413),
    E4M3 / E3M4 keep it in float."""
    rng = np.random.default_rng(9)
    x = (rng.standard_normal(1 << 16) * np.repeat(np.exp2(rng.integers(-12, 6, 64)), 1024)).astype(np.float16)
    xt = dev(x)
    want = ref_cuda.forward(xt.clone(), mode, xt.numel(), False, 1.0, True, bs)[0]
    got = F.cast_blocked(xt, mode, bs)
    assert_same(host(got), host(want), "%s half bs=%d" % (mode, bs))
    assert_same(host(got), O.cast_blocked(x, mode, bs), "%s half bs=%d vs oracle" % (mode, bs))


# ------------------------------------------------------------------------------------------ shapes, alignment, in-place
@pytest.mark.parametrize("n", [0, 1, 7, 8, 9, 15, 16, 17, 31, 255, 1000003])
@pytest.mark.parametrize("mode", ["E4M3_RNE", "E5M2_STOCHASTIC", "BFLOAT16_RNE"])
def test_ragged_sizes(F, mode, n):
    x = wide(max(n, 1), seed=n)[:n]
    got = host(F.cast(dev(x), mode, seed=7, scale=1.0)) if n else np.zeros(0, np.float32)
    if n:
        assert_same(got, O.cast(x, mode, seed=7), "%s n=%d" % (mode, n))
    for dt in (torch.float16, torch.bfloat16):
        xt = dev(x).to(dt)
        y = F.cast(xt, mode, seed=7)
        x_np = host(xt)
        assert_same(host(y), O.cast(x_np, mode, seed=7), "%s n=%d %s" % (mode, n, dt))


@pytest.mark.parametrize("mode", ["E4M3_RNE", "E3M4_STOCHASTIC", "E5M2_RNE"])
def test_unaligned_views_take_the_generic_kernel(F, mode):
    x = wide(4099)
    xt = dev(x)
    for off in (1, 3, 5):
        sub = xt[off:]
        out = torch.empty_like(xt)[off:]
        F.cast(sub, mode, out=out, seed=3, scale=3.7)
        assert_same(host(out), O.cast(x[off:], mode, seed=3, scale=3.7), "%s off=%d" % (mode, off))


@pytest.mark.parametrize("mode", ["E4M3_RNE", "E5M2_RNE", "E3M4_RNE", "BFLOAT16_RNE", "E4M3_STOCHASTIC"])
@pytest.mark.parametrize("dt", [torch.float32, torch.float16, torch.bfloat16])
def test_inplace_equals_out_of_place(F, mode, dt):
    xt = dev(wide(100003)).to(dt)
    a = F.cast(xt, mode, seed=9, scale=2.5)
    b = xt.clone()
    r = F.cast(b, mode, out=b, seed=9, scale=2.5)
    assert r.data_ptr() == b.data_ptr()
    assert_same(host(a), host(b), mode)


# ------------------------------------------------------------------------------------------ stochastic rounding
@pytest.mark.parametrize("mode", SR_MODES)
@pytest.mark.parametrize("dt", ["f32", "f16", "bf16"])
def test_stochastic_bits_match_oracle_given_the_documented_philox_stream(F, mode, dt):
    x = wide(1 << 15, seed=21)
    xt = dev(x)
    if dt == "f16":
        xt = xt.half()
    elif dt == "bf16":
        xt = xt.bfloat16()
    x_np = host(xt)
    for scale in (1.0, 3.7):
        if mode.startswith(("FLOAT16", "BFLOAT16")) and scale != 1.0:
            continue
        got = host(F.cast(xt, mode, seed=0xC0FFEE1234, scale=scale))
        assert_same(got, O.cast(x_np, mode, seed=0xC0FFEE1234, scale=scale), "%s %s s=%g" % (mode, dt, scale))


@pytest.mark.parametrize("mode", ["E5M2_STOCHASTIC", "E4M3_STOCHASTIC", "BFLOAT16_STOCHASTIC"])
def test_stochastic_is_partition_independent(F, mode):
    x = dev(wide(1 << 16, seed=4))
    whole = F.cast(x, mode, seed=77)
    for cut in (16 * 1000, 16 * 1000 + 8, 12345):       # aligned, half-aligned, arbitrary slab boundary
        a = F.cast(x[:cut].clone(), mode, seed=77, elem_offset=0)
        b = F.cast(x[cut:].clone(), mode, seed=77, elem_offset=cut)
        assert_same(host(torch.cat([a, b])), host(whole), "%s cut=%d" % (mode, cut))


def test_stochastic_distribution_matches_reference_counts(F, golden_dir):
    """Tolerance: |p_hat - p_ref| <= 5 * sqrt(2 p(1-p)/N) with N = 2^18 draws per probe value (the fixture holds
    the reference CPU op's observed counts); the two possible outputs must be exactly the reference's."""
    g = np.load(os.path.join(golden_dir, "stochastic.npz"))
    N = int(g["N"][0])
    for key in g.files:
        if key == "N":
            continue
        mode, v = key.split("@")
        ref_vals, ref_cnt = g[key][0].astype(np.uint32), g[key][1].astype(np.float64)
        x = torch.full((N,), float(v), dtype=torch.float32, device="cuda")
        got = host(F.cast(x, mode, seed=4242, unscale_fp32=True))
        u, c = np.unique(got.view(np.uint32), return_counts=True)
        assert set(u.tolist()) == set(ref_vals.tolist()), key
        for val, cnt in zip(u, c):
            p_ref = ref_cnt[list(ref_vals).index(val)] / N
            sigma = np.sqrt(max(p_ref * (1 - p_ref), 1e-9) / N)
            assert abs(cnt / N - p_ref) <= 5 * np.sqrt(2) * sigma, (key, cnt / N, p_ref)


def test_stochastic_is_unbiased_on_a_large_tensor(F):
    """E[q - x] = 0 for in-range values: |mean(q - x)| <= 5 sigma / sqrt(n), n = 16 Mi N(0,1) samples."""
    torch.manual_seed(0)
    x = torch.randn(1 << 24, device="cuda")
    for mode in ("E5M2_STOCHASTIC", "E4M3_STOCHASTIC", "E3M4_STOCHASTIC"):
        q = F.cast(x, mode, seed=1)
        keep = (x.abs() > 0.26) & (x.abs() < 16)                 # away from pseudo-subnormals / saturation
        d = (q - x.half().float())[keep].double()                # rounding error w.r.t. the fp16 value being rounded
        assert abs(d.mean().item()) <= 5 * d.std().item() / np.sqrt(d.numel()), mode


# ------------------------------------------------------------------------------------------ amax / scale / rowwise
@pytest.mark.parametrize("dt", [torch.float32, torch.float16, torch.bfloat16])
@pytest.mark.parametrize("n", [1, 5, 4096, 1000003])
def test_amax(F, dt, n):
    torch.manual_seed(n)
    x = (torch.randn(n, device="cuda") * 37).to(dt)
    got = F.amax(x).item()
    assert got == x.float().abs().max().item()
    off = F.amax(x[1:] if n > 1 else x).item()                  # unaligned view
    assert off == (x[1:] if n > 1 else x).float().abs().max().item()
    if n > 3:
        x[n - 2] = float("nan")                                 # propagates, like torch.max(abs(x)) (qutils.py:283)
        assert np.isnan(F.amax(x).item()) and np.isnan(F.amax(x[1:]).item())
        assert not np.isnan(F.amax(x[: n - 2]).item())


def test_per_tensor_amax_scaled_cast_without_host_sync(F):
    x = wide(1 << 18, seed=8)
    x = x[np.isfinite(x)]
    xt = dev(x)
    y, d_scale = F.cast_amax_scaled(xt, "E4M3_RNE", return_scale=True)
    s = O.scale_from_amax(O.amax(x), 448.0)
    assert d_scale.item() == np.float32(s)
    assert_same(host(y), O.cast(x, "E4M3_RNE", scale=s), "amax-scaled")
    z = torch.zeros(1000, device="cuda")                         # amax 0 -> inf -> clamp to 6.55e4
    _, d_scale = F.cast_amax_scaled(z, "E4M3_RNE", return_scale=True)
    assert d_scale.item() == np.float32(6.55e4)


@pytest.mark.parametrize("mode", ["E4M3_RNE", "E3M4_RNE", "E5M2_RNE", "E4M3_STOCHASTIC"])
@pytest.mark.parametrize("shape", [(64, 147), (128, 576), (7, 4608), (1000, 16), (3, 5), (16, 1280), (4, 12288),
                                   (3, 12296), (5, 24576), (2, 24592), (9, 1024), (9, 1032),
                                   (45, 592), (130, 272), (67, 1152), (1, 576)])   # whole-rows-per-CTA tile kernel
@pytest.mark.parametrize("dt", ["f32", "f16"])
def test_rowwise_per_channel_cast(F, mode, shape, dt):
    rng = np.random.default_rng(shape[0])
    x = (rng.standard_normal(shape) * np.exp2(rng.integers(-4, 4, (shape[0], 1)))).astype(np.float32)
    if dt == "f16":
        x = x.astype(np.float16)
    fmt = O.parse_mode(mode)[0]
    flt_max = {0: 57344.0, 1: 448.0, 2: 30.0}[fmt]
    got, scales = F.cast_rowwise(dev(x), mode, seed=5, return_scales=True)
    want = np.empty_like(x)
    for r in range(shape[0]):
        s = O.scale_from_amax(O.amax(x[r]), flt_max, like=x)
        assert scales[r].item() == np.float32(s)
        want[r] = O.cast(x[r], mode, scale=s, seed=5, elem_offset=r * shape[1])
    assert_same(host(got), want, "%s %s" % (mode, shape))


@pytest.mark.parametrize("mode", ["E5M2_RNE", "E4M3_RNE", "E3M4_RNE", "E4M3_STOCHASTIC", "E4M3_IEEE_RNE", "FP4_NEAREST"])
@pytest.mark.parametrize("bs", [1, 2, 5, 8, 16, 32, 128, 256, 768, 1000, 4096])
@pytest.mark.parametrize("dt", ["f32", "f16", "bf16"])
def test_blocked_vs_oracle(F, mode, bs, dt):
    rng = np.random.default_rng(bs)
    n = bs * 24
    x = (rng.standard_normal(n) * np.repeat(np.exp2(rng.integers(-6, 6, 24)), bs)).astype(np.float32)
    x[:bs] = 0.0                                                 # an all-zero block gives NaN, as in the reference
    xt = dev(x)
    if dt == "f16":
        xt = xt.half()
    elif dt == "bf16":
        xt = xt.bfloat16()
    got = host(F.cast_blocked(xt, mode, bs, seed=6))
    assert_same(got, O.cast_blocked(host(xt), mode, bs, seed=6), "%s bs=%d %s" % (mode, bs, dt))


# ------------------------------------------------------------------------------------------ error behaviour
def test_errors_are_exceptions_not_exit(F):
    x = torch.ones(16, device="cuda")
    with pytest.raises(ValueError):
        F.cast(x, "E9M9_RNE")
    with pytest.raises(ValueError):
        F.cast(x, "E4M3_RNE", scale=70000.0)
    with pytest.raises(RuntimeError):
        F.cast(torch.ones(16), "E4M3_RNE")
    with pytest.raises(RuntimeError):
        F.cast(torch.ones(4, 4, device="cuda").t(), "E4M3_RNE")
    assert F.cast(x.double(), "E4M3_RNE").dtype == torch.float64       # doubles have their own entry point (test_gpu_f64.py)
    with pytest.raises(TypeError):
        F.cast(x.to(torch.int32), "E4M3_RNE")
    with pytest.raises(TypeError):
        F.cast_amax_scaled(x.double(), "E4M3_RNE")                     # ... and only that one
    with pytest.raises(ValueError):
        F.cast_blocked(x, "E4M3_RNE", 5)


# ------------------------------------------------------------------------------------------ multi-tensor launches
@pytest.mark.parametrize("mode", ["E4M3_RNE", "E5M2_STOCHASTIC", "E3M4_RTZ", "BFLOAT16_RNE", "FLOAT16_STOCHASTIC"])
@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16, torch.float16])
def test_multi_tensor_launch_equals_per_tensor_calls(F, mode, dt):
    """One launch over a ragged list of tensors (sizes 0 .. 300 001, several CTAs per tensor and several tensors per
    CTA-size) == the per-tensor entry point, bit for bit, including per-tensor seeds and element offsets."""
    torch.manual_seed(1)
    sizes = [1, 7, 16, 17, 4096, 0, 300001, 65536, 8, 123457, 32, 2049] + [1000 + 16 * i for i in range(300)]
    xs = [(torch.randn(n, device="cuda") * (1 + (i % 5))).to(dt) for i, n in enumerate(sizes)]
    offs = [16 * (i * 3) for i in range(len(xs))]
    plan = F.MultiTensorPlan(xs, [torch.empty_like(x) for x in xs], offs)
    scales = (torch.rand(len(xs), device="cuda") * 4 + 0.25)
    outs = plan.cast(mode, d_scales=scales, seed=1234)
    for t, (x, y) in enumerate(zip(xs, outs)):
        want = F.cast(x, mode, d_scale=scales[t:t + 1], seed=1234 + t, elem_offset=offs[t])
        assert_same(host(y), host(want), "%s tensor %d (n=%d)" % (mode, t, x.numel()))
    outs = plan.cast(mode, seed=99)                                   # unscaled
    for t in (0, 3, 6, 11, 200):
        assert_same(host(outs[t]), host(F.cast(xs[t], mode, seed=99 + t, elem_offset=offs[t])), "unscaled %d" % t)
    vec = plan.amax()
    for t, x in enumerate(xs):
        assert vec[t].item() == (x.float().abs().max().item() if x.numel() else 0.0)


def test_multi_tensor_amax_scaled_cast_and_rejects_unaligned(F):
    xs = [torch.randn(n, device="cuda") * 0.02 for n in (4096 * 11, 777, 1 << 20)]
    plan = F.MultiTensorPlan(xs, [torch.empty_like(x) for x in xs])
    outs, scales = plan.amax_scaled_cast("E4M3_RNE")
    for x, y, s in zip(xs, outs, scales):
        want, ws = F.cast_amax_scaled(x, "E4M3_RNE", return_scale=True)
        assert ws.item() == s.item()
        assert_same(host(y), host(want), "amax-scaled multi")
    # two launches per sweep (the cast kernel derives the scales): same bits as amax + scale kernel + cast with scales
    vec = plan.amax()
    ref_scales = F.scale_from_amax(vec, 448.0)
    ref = [t.clone() for t in plan.cast("E4M3_RNE", d_scales=ref_scales)]
    assert torch.equal(ref_scales, scales)
    for y, r in zip(outs, ref):
        assert_same(host(y), host(r), "scales derived in the cast kernel")
    # pipelined schedule (runs of tensors, an exchange step per run) changes nothing; stochastic seeds stay per tensor
    seen = []
    for depth in (1, 2, 3, 7):
        o, sc = plan.amax_scaled_cast("E4M3_STOCHASTIC", seed=77, reduce_fn=lambda v: seen.append(v.numel()), pipeline=depth)
        got = [host(t.clone()) for t in o]
        if depth == 1:
            first, first_sc = got, sc.clone()
        assert torch.equal(sc, first_sc)
        for a, b in zip(got, first):
            assert_same(a, b, "pipeline depth %d" % depth)
    assert seen == [3, 2, 1, 1, 1, 1, 1, 1, 1] or seen == [3, 1, 2, 1, 1, 1, 1, 1, 1]
    for t, x in enumerate(xs):
        assert_same(first[t], host(F.cast(x, "E4M3_STOCHASTIC", d_scale=scales[t:t + 1], seed=77 + t)), "per-tensor seed")
    # half tensors: the kernel-derived scale follows the half arithmetic of torch's `flt_max / vmax`
    xh = [x.half() for x in xs]
    ph = F.MultiTensorPlan(xh, [torch.empty_like(x) for x in xh])
    oh, sh = ph.amax_scaled_cast("E5M2_RNE")
    for x, y, s in zip(xh, oh, sh):
        want, ws = F.cast_amax_scaled(x, "E5M2_RNE", return_scale=True)
        assert ws.item() == s.item()
        assert_same(host(y), host(want), "amax-scaled multi, half")
    # a tensor with no elements here (an empty row slab of a sharded list) has no CTA in the cast launch, but its
    # scale — from the maximum the other ranks contributed — is still part of the result
    xe = [xs[0], torch.empty(0, device="cuda"), xs[1]]
    pe = F.MultiTensorPlan(xe, [torch.empty_like(x) for x in xe])

    def other_ranks(v):
        v[1:2].fill_(0.75)
    for _ in range(2):
        _, se = pe.amax_scaled_cast("E4M3_RNE", reduce_fn=other_ranks)
        assert se[1].item() == F.scale_from_amax(torch.tensor([0.75], device="cuda"), 448.0).item()
        assert se[0].item() == scales[0].item() and se[2].item() == scales[1].item()
    with pytest.raises(ValueError):
        F.MultiTensorPlan([xs[0][1:]])                                # not 32-byte aligned
    # element offsets that are not on a Philox-block boundary (slabs of a sharded tensor with an odd row length):
    # the list launch draws the straddled blocks (ANYOFF kernel instance), same bits as the per-tensor casts
    offs = [8, 3, 1 << 33 | 5]
    for dt in (torch.float32, torch.bfloat16):
        xo = [x.to(dt) for x in xs]
        po = F.MultiTensorPlan(xo, [torch.empty_like(x) for x in xo], elem_offsets=offs)
        for mode in ("E4M3_STOCHASTIC", "BFLOAT16_STOCHASTIC", "E5M2_RNE"):
            outs_o = po.cast(mode, seed=31)
            for t, (x, y) in enumerate(zip(xo, outs_o)):
                assert_same(host(y), host(F.cast(x, mode, seed=31 + t, elem_offset=offs[t])), "%s odd offset %d" % (mode, t))
    with pytest.raises(ValueError):
        F.MultiTensorPlan(xs, [torch.empty(x.shape, dtype=torch.uint8, device="cuda") for x in xs],
                          elem_offsets=offs).encode("E4M3_STOCHASTIC")      # code kernels: aligned offsets only


# ------------------------------------------------------------------------------------------ out-of-bounds guard
# (compute-sanitizer is closed on this GPU pool, so writes past the end are caught with sentinel-filled guard zones)
@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16])
def test_no_writes_outside_the_output_buffer(F, dt):
    guard = 64
    sent = 123.0
    for n in (1, 5, 8, 9, 15, 16, 17, 255, 256, 257, 4095, 4097, 65537):
        for off in (0, 1, 3, 8):
            big = torch.full((guard + off + n + guard,), sent, dtype=dt, device="cuda")
            out = big[guard + off: guard + off + n]
            x = (torch.randn(n, device="cuda") * 3).to(dt)
            for mode in ("E4M3_RNE", "E5M2_STOCHASTIC", "BFLOAT16_RNE"):
                F.cast(x, mode, out=out, seed=1)
                assert bool((big[:guard + off] == sent).all()) and bool((big[guard + off + n:] == sent).all()), (n, off, mode)
            if n % 8 == 0:
                F.cast_blocked(x, "E4M3_RNE", 8, out=out)
                assert bool((big[:guard + off] == sent).all()) and bool((big[guard + off + n:] == sent).all()), (n, off, "blocked")
            F.cast_rowwise(x.view(1, n), "E4M3_RNE", out=out.view(1, n))
            F.cast_amax_scaled(x, "E3M4_RNE", out=out)
            F.cast_amax_scaled_noalloc(x, "E4M3_RNE", out=out)
            assert bool((big[:guard + off] == sent).all()) and bool((big[guard + off + n:] == sent).all()), (n, off, "scaled")
    # multi-tensor: outputs packed back to back inside one guarded arena (32-byte aligned pieces)
    sizes = [16, 48, 1000 * 16, 7 * 16, 33 * 16]
    arena = torch.full((guard + sum(sizes) + guard,), sent, dtype=dt, device="cuda")
    pos, outs = guard, []
    for n in sizes:
        outs.append(arena[pos: pos + n])
        pos += n
    xs = [(torch.randn(n, device="cuda")).to(dt) for n in sizes]
    if all(o.data_ptr() % 32 == 0 for o in outs):
        F.MultiTensorPlan(xs, outs).cast("E4M3_RNE")
        assert bool((arena[:guard] == sent).all()) and bool((arena[pos:] == sent).all())
        for x, o in zip(xs, outs):
            assert_same(host(o), host(F.cast(x, "E4M3_RNE")), "arena piece")


# ------------------------------------------------------------------------------------------ streams and graphs
def test_runs_on_the_current_torch_stream_without_device_sync(F):
    """The reference launches on the legacy default stream and synchronises the device (fpemu_kernels.cu:1060-1071);
    this library must enqueue on torch's CURRENT stream.  A non-blocking side stream first sleeps, then produces the
    input: a cast issued on any other stream would read the stale zeros."""
    s = torch.cuda.Stream()
    x = torch.zeros(1 << 20, device="cuda")
    with torch.cuda.stream(s):                             # first use loads the kernels (lazy module loading may sync)
        F.cast(x, "E4M3_RNE")
        F.cast_amax_scaled(x, "E5M2_RNE")
    torch.cuda.synchronize()
    with torch.cuda.stream(s):
        torch.cuda._sleep(400_000_000)                     # ~0.2 s of GPU time on the side stream
        x.fill_(1.3)
        y = F.cast(x, "E4M3_RNE")
        z = F.cast_amax_scaled(x, "E5M2_RNE")
    assert not s.query()                                   # the call returned while the stream is still busy: no sync
    s.synchronize()
    want = O.cast(np.full(4, 1.3, dtype=np.float32), "E4M3_RNE")[0]
    assert bool((y == float(want)).all())
    assert bool((z == 1.3125).all()) or bool((z == z[0]).all() and z[0] != 0)


def test_cuda_graph_capture_and_replay(F):
    """Plain, amax-scaled, row-wise and multi-tensor casts are capture-safe (no allocation, no sync inside the
    library): capture once, change the input, replay, compare with eager results."""
    x = torch.randn(1 << 18, device="cuda")
    y1, y2 = torch.empty_like(x), torch.empty_like(x)
    m = torch.randn(64, 1024, device="cuda")
    ym = torch.empty_like(m)
    parts = [torch.randn(4096, device="cuda") for _ in range(5)]
    plan = F.MultiTensorPlan(parts, [torch.empty_like(p) for p in parts])
    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        F.cast(x, "E4M3_STOCHASTIC", out=y1, seed=11)       # warm-up (scratch for the stream gets created here)
        F.cast_amax_scaled(x, "E4M3_RNE", out=y2)
        F.cast_rowwise(m, "E3M4_RNE", out=ym)
        plan.cast("E5M2_RNE")
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=side):
            F.cast(x, "E4M3_STOCHASTIC", out=y1, seed=11)
            F.cast_amax_scaled(x, "E4M3_RNE", out=y2)
            F.cast_rowwise(m, "E3M4_RNE", out=ym)
            plan.cast("E5M2_RNE")
    torch.cuda.synchronize()
    x.mul_(3.0).add_(0.25)
    m.mul_(0.5)
    for p in parts:
        p.add_(1.0)
    g.replay()
    torch.cuda.synchronize()
    assert_same(host(y1), host(F.cast(x, "E4M3_STOCHASTIC", seed=11)), "graph plain")
    assert_same(host(y2), host(F.cast_amax_scaled(x, "E4M3_RNE")), "graph amax-scaled")
    assert_same(host(ym), host(F.cast_rowwise(m, "E3M4_RNE")), "graph row-wise")
    for p, o in zip(parts, plan.outs):
        assert_same(host(o), host(F.cast(p, "E5M2_RNE")), "graph multi")


@pytest.mark.parametrize("mode", ["E5M2_RNE", "E4M3_RNE", "E3M4_RNE"])
@pytest.mark.parametrize("dt", ["f32", "f16"])
def test_blocked_with_one_block_spanning_a_large_tensor(F, mode, dt):
    """block_size == size: one block = the whole tensor (the reference CPU op's reading, cpp/fpemu_impl.cpp:778-779)."""
    rng = np.random.default_rng(4)
    x = (rng.standard_normal(200000) * 0.37).astype(np.float32)
    if dt == "f16":
        x = x.astype(np.float16)
    got = host(F.cast_blocked(dev(x), mode, x.size))
    assert_same(got, O.cast_blocked(x, mode, x.size), "%s %s whole-tensor block" % (mode, dt))
    import fpemu_cuda
    y = fpemu_cuda.forward(dev(x), mode, x.size, False, 1.0, True, x.size)[0]
    assert_same(host(y), got, "through fpemu_cuda.forward")


@pytest.mark.parametrize("mode", ["E4M3_RNE", "E5M2_STOCHASTIC", "E3M4_RTZ", "E4M3_STOCHASTIC"])
@pytest.mark.parametrize("dt", ["f32", "f16", "bf16"])
def test_single_launch_amax_scaled_cast_equals_the_three_launch_path(F, mode, dt):
    """fpemu_cast_amax_scaled_fused: one cooperative launch for tensors that fit in registers, the three-launch path
    above that; the result must not depend on which one ran (sizes straddle the capacity of about 4.8 M fp32
    elements), must match the oracle, and amax / scale must come back in the work buffer."""
    tdt = {"f32": torch.float32, "f16": torch.float16, "bf16": torch.bfloat16}[dt]
    g = torch.Generator(device="cuda").manual_seed(12)
    for n in (1, 7, 8, 4097, 65536 + 3, 768 * 768, 12 * 384 * 768, 4849664, 4849664 + 8, 9699328, 9699328 + 16, 11000000):
        x = (torch.randn(n, device="cuda", generator=g) * 0.37).to(tdt)
        a, sa = F.cast_amax_scaled(x, mode, seed=5, return_scale=True, fused=True)
        b, sb = F.cast_amax_scaled(x, mode, seed=5, return_scale=True, fused=False)
        bits = torch.int32 if dt == "f32" else torch.int16
        assert torch.equal(a.view(bits), b.view(bits)), "%s %s n=%d" % (mode, dt, n)
        assert torch.equal(sa, sb)
        y = x.clone()
        F.cast_amax_scaled(y, mode, out=y, seed=5)                                   # in place
        assert torch.equal(y.view(bits), b.view(bits))
    x = dev(wide(300001, 77) * np.float32(1e-3))
    got = host(F.cast_amax_scaled(x, "E4M3_RNE"))
    xn = host(x)
    s = O.scale_from_amax(O.amax(xn), 448.0)
    assert_same(got, O.cast(xn, "E4M3_RNE", scale=s), "fused vs oracle")


def test_single_launch_amax_scaled_cast_in_a_cuda_graph(F):
    x = torch.randn(1 << 20, device="cuda")
    y = torch.empty_like(x)
    want = F.cast_amax_scaled(x, "E4M3_RNE", fused=False)
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        F.cast_amax_scaled(x, "E4M3_RNE", out=y)                                     # warm-up on the capture stream
        s.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=s):
            F.cast_amax_scaled(x, "E4M3_RNE", out=y)
    y.zero_()
    x.mul_(3.0)
    want = F.cast_amax_scaled(x, "E4M3_RNE", fused=False)
    graph.replay()
    torch.cuda.synchronize()
    assert torch.equal(y.view(torch.int32), want.view(torch.int32))


def test_concurrent_host_threads_on_their_own_streams(F):
    """Four host threads, each with its own stream and tensors, hammer the entry points at once (ctypes drops the GIL
    during a call): per-device caches inside the library are lock-protected, per-call state lives in the arguments."""
    import threading
    xs = [dev(wide(200003, 60 + i)) for i in range(4)]
    want = [(F.cast(x, "E4M3_STOCHASTIC", seed=5 + i).clone(), F.cast_amax_scaled(x, "E4M3_RNE", fused=True).clone(),
             F.encode(x, "E5M2_RNE").clone()) for i, x in enumerate(xs)]
    torch.cuda.synchronize()
    errors = []

    def worker(i):
        try:
            s = torch.cuda.Stream()
            with torch.cuda.stream(s):
                for _ in range(100):
                    a = F.cast(xs[i], "E4M3_STOCHASTIC", seed=5 + i)
                    b = F.cast_amax_scaled(xs[i], "E4M3_RNE")
                    c = F.encode(xs[i], "E5M2_RNE")
                s.synchronize()
                bits = torch.int32
                if not (torch.equal(a.view(bits), want[i][0].view(bits)) and torch.equal(b.view(bits), want[i][1].view(bits))
                        and torch.equal(c, want[i][2])):
                    errors.append("thread %d: results differ" % i)
        except Exception as e:  # noqa: BLE001
            errors.append("thread %d: %r" % (i, e))
    threads = [threading.Thread(target=worker, args=(i,)) for i in range(4)]
    for t in threads:
        t.start()
    for t in threads:
        t.join(timeout=120)
    assert not errors, errors


def test_out_of_place_cast_of_a_view_gets_an_equally_aligned_output(F):
    """x[1:] starts 4 bytes off a 32-byte boundary; the output the library allocates starts off by the same amount so
    that the vector kernels run (input and output misaligned alike), and holds the right values."""
    import fpemu_cuda
    for dt in (torch.float32, torch.bfloat16, torch.float16):
        base = torch.randn(100003, device="cuda").to(dt)
        for off in (1, 3, 8, 13):
            x = base[off:]
            y = F.cast(x, "E4M3_RNE")
            assert y.data_ptr() % 32 == x.data_ptr() % 32 and y.shape == x.shape and y.is_contiguous()
            want = F.cast(x.clone(), "E4M3_RNE")                      # aligned copy of the same values
            bits = torch.int32 if dt == torch.float32 else torch.int16
            assert torch.equal(y.view(bits), want.view(bits))
            z = fpemu_cuda.forward(x, "E5M2_RNE", x.numel(), False, 1.0, False, 1)[0]
            assert z.data_ptr() % 32 == x.data_ptr() % 32
            assert torch.equal(z.view(bits), F.cast(x.clone(), "E5M2_RNE").view(bits))
