"""GPU tests of the 8-bit wire form (fpemu_encode / fpemu_decode through the C ABI):
  * codes bit-exact against the C oracle (itself pinned to the reference's values in tests/test_oracle_codec.py);
  * decode bit-exact against the oracle on all 256 codes x dtype x scale x un-scale semantics;
  * decode(encode(x)) == cast(x): always for E5M2, for every value that is a member of the format otherwise."""
import numpy as np
import pytest
import torch

from oracle import fpemu_oracle as O
from test_gpu_cast_parity import assert_same, dev, fp16_patterns, host, wide

pytestmark = pytest.mark.gpu

FORMATS = ["E5M2", "E4M3", "E3M4"]
MODES = ["%s_%s" % (f, r) for f in FORMATS for r in ("RNE", "RTZ", "STOCHASTIC")] + \
        ["E5M2_DAZ_RNE", "E5M2_DAZ_STOCHASTIC", "E5M2_RNAZ", "E5M2_RPINF"]
MIN_NORMAL = {"E5M2": 2.0 ** -14, "E4M3": 2.0 ** -6, "E3M4": 2.0 ** -2}
TORCH_DT = {"f32": torch.float32, "f16": torch.float16, "bf16": torch.bfloat16}


@pytest.fixture(scope="module")
def F():
    import fpemu_b200
    fpemu_b200._lib.lib()          # raises if the CUDA library is missing: no fallback
    return fpemu_b200


@pytest.fixture(autouse=True)
def _cuda_nan_mode():
    O.set_nan_mode(False)
    yield


def as_dtype(x, dt):
    """fp32 numpy -> (numpy array the oracle reads, torch CUDA tensor the library reads)."""
    if dt == "f32":
        return x, dev(x)
    if dt == "f16":
        h = x.astype(np.float16)
        return h, dev(h)
    t = torch.from_numpy(x).to(torch.bfloat16)
    return t.view(torch.int16).numpy().view(np.uint16), t.cuda()


def codes_equal(got, want, what):
    got, want = got.cpu().numpy().ravel(), want.ravel()
    bad = np.nonzero(got != want)[0]
    assert bad.size == 0, "%s: %d / %d codes differ, first at %d: got 0x%02x want 0x%02x" % (
        what, bad.size, got.size, bad[0], got[bad[0]], want[bad[0]])


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("dt", ["f32", "f16", "bf16"])
def test_codes_vs_oracle(F, mode, dt):
    for x, scale in ((fp16_patterns(), 1.0), (wide(70001), 1.0), (wide(4096 + 13, 5), 3.7), (wide(999, 6), 0.125)):
        xn, xt = as_dtype(x, dt)
        got = F.encode(xt, mode, scale=scale, seed=77, elem_offset=48)
        codes_equal(got, O.encode(xn, mode, scale=scale, seed=77, elem_offset=48), "%s %s scale %g n %d" % (mode, dt, scale, x.size))


@pytest.mark.parametrize("mode", ["E5M2_RNE", "E4M3_STOCHASTIC", "E3M4_RNE"])
def test_codes_ragged_sizes_unaligned_pointers_and_device_scale(F, mode):
    base = wide(5000, 8)
    d_scale = torch.tensor([3.7], device="cuda")
    for n in (0, 1, 7, 8, 9, 15, 16, 17, 31, 33, 2047, 2048, 2049, 4099):
        x = base[:n].copy()
        got = F.encode(dev(x), mode, seed=5) if n else torch.empty(0, dtype=torch.uint8, device="cuda")
        if n:
            codes_equal(got, O.encode(x, mode, seed=5), "%s n=%d" % (mode, n))
    xt = dev(base)
    for off in (1, 3, 8):                                     # input and codes start off the vector alignment
        out = torch.zeros(base.size + 64, dtype=torch.uint8, device="cuda")
        view = out[off:off + base.size - off]
        F.encode(xt[off:], mode, out=view, d_scale=d_scale, seed=9, elem_offset=off)
        codes_equal(view, O.encode(base[off:], mode, scale=3.7, seed=9, elem_offset=off), "%s offset %d" % (mode, off))
        assert int(out[:off].sum()) == 0 and int(out[base.size:].sum()) == 0          # nothing written outside


@pytest.mark.parametrize("fmt", FORMATS)
@pytest.mark.parametrize("dt", ["f32", "f16", "bf16"])
def test_decode_all_codes_vs_oracle(F, fmt, dt):
    np_dt = {"f32": np.float32, "f16": np.float16, "bf16": np.uint16}[dt]
    codes = np.tile(np.arange(256, dtype=np.uint8), 9)[:2303]               # vector body + ragged tail
    for scale in (1.0, 0.125, 3.7, 94.09):
        for unscale_fp32 in (False, True):
            got = host(F.decode(dev(codes), fmt, dtype=TORCH_DT[dt], scale=scale, unscale_fp32=unscale_fp32))
            want = O.decode(codes, fmt, dtype=np_dt, scale=scale, unscale_fp32=unscale_fp32)
            if dt == "bf16":
                got, want = (got.astype(np.uint32) << 16).view(np.float32), (want.astype(np.uint32) << 16).view(np.float32)
            assert_same(got, want, "%s %s scale %g fp32-unscale %s" % (fmt, dt, scale, unscale_fp32))
    # unaligned code pointer: element-wise kernel
    got = host(F.decode(dev(codes)[3:], fmt, dtype=TORCH_DT[dt]))
    want = O.decode(codes[3:], fmt, dtype=np_dt)
    if dt == "bf16":
        got, want = (got.astype(np.uint32) << 16).view(np.float32), (want.astype(np.uint32) << 16).view(np.float32)
    assert_same(got, want, "%s %s unaligned" % (fmt, dt))


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("dt", ["f32", "f16", "bf16"])
def test_round_trip_reproduces_the_cast(F, mode, dt):
    fmt = mode.split("_")[0]
    for x, scale in ((fp16_patterns(), 1.0), (wide(1 << 17, 21), 1.0), (wide(1 << 16, 22), 3.7)):
        _, xt = as_dtype(x, dt)
        for unscale_fp32 in (False, True):
            q = F.cast(xt, mode, scale=scale, seed=31, unscale_fp32=unscale_fp32)
            back = F.decode(F.encode(xt, mode, scale=scale, seed=31), fmt, dtype=xt.dtype, scale=scale,
                            unscale_fp32=unscale_fp32)
            same = (q.view(torch.int16 if xt.element_size() == 2 else torch.int32) ==
                    back.view(torch.int16 if xt.element_size() == 2 else torch.int32)) | (q.isnan() & back.isnan())
            if fmt == "E5M2":
                assert bool(same.all()), "%s %s scale %g: %d differ" % (mode, dt, scale, int((~same).sum()))
            else:
                # at or above the smallest normal (after scaling, with slack for the scale's own rounding) every
                # rounded value is a member of the format
                big = (xt.float().abs() * scale) >= 2.0 * MIN_NORMAL[fmt]
                assert bool(same[big].all()), "%s %s scale %g: %d differ" % (mode, dt, scale, int((~same[big]).sum()))
                zero = xt.float() == 0
                assert bool(same[zero].all())


def test_codes_do_not_depend_on_how_the_tensor_is_split(F):
    x = dev(wide(1 << 16, 30))
    whole = F.encode(x, "E4M3_STOCHASTIC", seed=1234)
    parts = [F.encode(x[a:b], "E4M3_STOCHASTIC", seed=1234, elem_offset=a) for a, b in ((0, 4096), (4096, 50000 - 50000 % 16), (50000 - 50000 % 16, 1 << 16))]
    assert torch.equal(whole, torch.cat(parts))


def test_codec_rejects_other_formats_and_bad_arguments(F):
    x = torch.zeros(64, device="cuda")
    for mode in ("BFLOAT16_RNE", "FLOAT16_RNE", "E4M3_IEEE_RNE", "FP4_NEAREST"):
        with pytest.raises(ValueError):
            F.encode(x, mode)
    with pytest.raises(ValueError):
        F.encode(x, "E4M3_RNE", scale=1e6)
    with pytest.raises(ValueError):
        F.encode(x, "E4M3_RNE", out=torch.zeros(8, dtype=torch.uint8, device="cuda"))
    with pytest.raises(ValueError):
        F.decode(torch.zeros(8, dtype=torch.float32, device="cuda"), "E4M3")


def test_full_size_round_trip(F):
    """2^30 elements (BASELINE configs[1] size), size-independent properties: the E5M2 round trip equals the cast
    everywhere; an E4M3 code is a fixed point of decode -> encode; the E4M3 round trip equals the cast at and above
    twice the smallest normal."""
    n = 1 << 30
    g = torch.Generator(device="cuda").manual_seed(3)
    x = torch.randn(n, device="cuda", generator=g)
    codes = F.encode(x, "E5M2_RNE")
    back = F.decode(codes, "E5M2")
    q = F.cast(x, "E5M2_RNE")
    assert torch.equal(back.view(torch.int32), q.view(torch.int32))
    del back, q
    codes = F.encode(x, "E4M3_RNE", out=codes)
    back = F.decode(codes, "E4M3")
    assert torch.equal(F.encode(back, "E4M3_RNE"), codes)                  # a code is a fixed point
    q = F.cast(x, "E4M3_RNE")
    big = x.abs() >= 2.0 ** -5
    assert torch.equal(back[big].view(torch.int32), q[big].view(torch.int32))


@pytest.mark.parametrize("mode", ["E4M3_RNE", "E5M2_STOCHASTIC", "E3M4_RTZ"])
@pytest.mark.parametrize("dt", ["f32", "bf16"])
def test_multi_tensor_encode_equals_per_tensor_encode(F, mode, dt):
    sizes = [4096, 16, 70001, 8 * 256 * 2 * 3, 5, 33333, 1 << 18]
    xs = [as_dtype(wide(n, 40 + i), dt)[1] for i, n in enumerate(sizes)]
    codes = [torch.zeros(n + 32, dtype=torch.uint8, device="cuda")[:n] for n in sizes]
    plan = F.MultiTensorPlan(xs, codes)
    scales = torch.tensor([1.0, 3.7, 0.125, 94.09, 2.0, 1.0, 0.5], device="cuda")
    plan.encode(mode, d_scales=scales, seed=900)
    for t, x in enumerate(xs):
        want = F.encode(x, mode, d_scale=scales[t:t + 1].clone(), seed=900 + t)
        assert torch.equal(codes[t], want), "tensor %d of %s %s" % (t, mode, dt)
    plan.encode(mode, seed=17)                                       # unscaled
    for t, x in enumerate(xs):
        assert torch.equal(codes[t], F.encode(x, mode, seed=17 + t))
    with pytest.raises(ValueError):
        plan.cast(mode)
    with pytest.raises(ValueError):
        F.MultiTensorPlan(xs).encode(mode)


def test_multi_tensor_amax_scaled_codes_decode_to_the_amax_scaled_cast(F):
    g = torch.Generator(device="cuda").manual_seed(8)
    xs = [torch.randn(n, device="cuda", generator=g) * s for n, s in ((1 << 16, 0.02), (4096 * 3, 5.0), (100000, 1e-3))]
    codes = [torch.empty(x.numel(), dtype=torch.uint8, device="cuda") for x in xs]
    got, scales = F.MultiTensorPlan(xs, codes).amax_scaled_encode("E4M3_RNE")
    outs = [torch.empty_like(x) for x in xs]
    want, scales2 = F.MultiTensorPlan(xs, outs).amax_scaled_cast("E4M3_RNE")
    assert torch.equal(scales, scales2)
    for t, x in enumerate(xs):
        back = F.decode(got[t], "E4M3", scale=float(scales[t]))
        big = (x.abs() * scales[t]) >= 2.0 ** -5
        assert torch.equal(back[big].view(torch.int32), want[t][big].view(torch.int32))
        assert float(big.float().mean()) > 0.99


@pytest.mark.parametrize("dt", ["f32", "f16", "bf16"])
def test_expect_normals_hint_never_changes_a_code(F, dt):
    """FPEMU_FLAG_EXPECT_NORMALS switches E4M3_RNE to the hardware fp16 -> e4m3 conversion for vectors without
    sub-normal-range lanes or NaN: every fp16 pattern, wide data (hint badly wrong) and amax-scaled data (hint right)."""
    for x, scale in ((fp16_patterns(), 1.0), (wide(1 << 17, 51), 1.0), (wide(1 << 16, 52), 3.7)):
        xn, xt = as_dtype(x, dt)
        got = F.encode(xt, "E4M3_RNE", scale=scale, expect_normals=True)
        codes_equal(got, O.encode(xn, "E4M3_RNE", scale=scale), "hint on, %s scale %g" % (dt, scale))
    g = torch.Generator(device="cuda").manual_seed(2)
    x = (torch.randn((1 << 20) + 5, device="cuda", generator=g) * 0.02).to(TORCH_DT[dt])
    s = F.scale_from_amax(F.amax(x), 448.0)
    assert torch.equal(F.encode(x, "E4M3_RNE", d_scale=s, expect_normals=True), F.encode(x, "E4M3_RNE", d_scale=s))
    for mode in ("E4M3_STOCHASTIC", "E5M2_RNE", "E3M4_RNE"):                          # ignored for other modes
        assert torch.equal(F.encode(x, mode, seed=3, expect_normals=True), F.encode(x, mode, seed=3))


@pytest.mark.parametrize("mode", ["E4M3_RNE", "E5M2_STOCHASTIC", "E3M4_RTZ", "E4M3_STOCHASTIC", "E5M2_RNAZ"])
@pytest.mark.parametrize("dt", ["f32", "f16", "bf16"])
@pytest.mark.parametrize("shape", [(7, 64), (64, 576), (33, 4096), (5, 12288), (3, 12288 + 32), (9, 100)])
def test_rowwise_codes_and_scales_vs_oracle(F, mode, dt, shape):
    """Per-channel wire form: codes and per-row scales against the oracle's per-row loop; rows in registers (warp- and
    CTA-sized), rows too long for registers, and a column count that is not a vector multiple (one warp per row)."""
    fmt = mode.split("_")[0]
    flt_max = {"E5M2": 57344.0, "E4M3": 448.0, "E3M4": 30.0}[fmt]
    rng = np.random.default_rng(shape[0] * 1000 + shape[1])
    x = (rng.standard_normal(shape) * np.exp2(rng.integers(-6, 3, (shape[0], 1)))).astype(np.float32)
    xn, xt = as_dtype(x.reshape(-1), dt)
    xn, xt = xn.reshape(shape), xt.reshape(shape)
    codes, scales = F.encode_rowwise(xt, mode, seed=21)
    want_codes, want_scales = O.encode_rowwise(xn, mode, flt_max, seed=21)
    assert (scales.cpu().numpy() == want_scales).all()
    codes_equal(codes, want_codes, "%s %s %s" % (mode, dt, shape))
    np_dt = {"f32": np.float32, "f16": np.float16, "bf16": np.uint16}[dt]
    for unscale_fp32 in (False, True):
        back = F.decode_rowwise(codes, scales, fmt, dtype=TORCH_DT[dt], unscale_fp32=unscale_fp32)
        got = host(back)
        want = O.decode_rowwise(want_codes, want_scales, fmt, dtype=np_dt, unscale_fp32=unscale_fp32)
        if dt == "bf16":
            got, want = (got.astype(np.uint32) << 16).view(np.float32), (want.astype(np.uint32) << 16).view(np.float32)
        assert_same(got.reshape(-1), want.reshape(-1), "decode %s %s %s" % (mode, dt, shape))


def test_rowwise_round_trip_reproduces_cast_rowwise(F):
    g = torch.Generator(device="cuda").manual_seed(4)
    for shape in ((4096, 4096), (512, 11008), (1000, 576)):
        x = torch.randn(shape, device="cuda", generator=g) * 0.02
        q, s0 = F.cast_rowwise(x, "E4M3_RNE", return_scales=True)
        codes, s1 = F.encode_rowwise(x, "E4M3_RNE")
        assert torch.equal(s0, s1)
        back = F.decode_rowwise(codes, s1, "E4M3")
        big = (x.abs() * s1[:, None]) >= 2.0 ** -5
        assert torch.equal(back[big].view(torch.int32), q[big].view(torch.int32))
        assert float(big.float().mean()) > 0.99
        q5, s5 = F.cast_rowwise(x, "E5M2_RNE", return_scales=True)
        c5, t5 = F.encode_rowwise(x, "E5M2_RNE")
        assert torch.equal(F.decode_rowwise(c5, t5, "E5M2").view(torch.int32), q5.view(torch.int32))


def test_codec_calls_inside_a_cuda_graph(F):
    x = torch.randn(1 << 18, device="cuda")
    x2 = torch.randn(256, 1024, device="cuda")
    codes = torch.empty(x.numel(), dtype=torch.uint8, device="cuda")
    back = torch.empty_like(x)
    rc = torch.empty(x2.shape, dtype=torch.uint8, device="cuda")
    rs = torch.empty(256, dtype=torch.float32, device="cuda")
    rb = torch.empty_like(x2)
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        def work():
            F.encode(x, "E4M3_STOCHASTIC", out=codes, seed=3)
            F.decode(codes, "E4M3", out=back)
            F.encode_rowwise(x2, "E5M2_RNE", codes=rc, scales=rs)
            F.decode_rowwise(rc, rs, "E5M2", out=rb)
        work()
        s.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=s):
            work()
    x.mul_(1.7)
    x2.add_(0.3)
    graph.replay()
    torch.cuda.synchronize()
    assert torch.equal(codes, F.encode(x, "E4M3_STOCHASTIC", seed=3))
    assert torch.equal(back, F.decode(codes, "E4M3"))
    assert torch.equal(rb.view(torch.int32), F.cast_rowwise(x2, "E5M2_RNE").view(torch.int32))
