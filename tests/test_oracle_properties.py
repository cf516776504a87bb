"""Size-independent properties of the restated algorithm (CPU, oracle only): every deterministic mode is
idempotent and monotone, nearest modes are sign-symmetric and within half an ulp of the target grid, truncation never
increases the magnitude, stochastic rounding lands on one of the two neighbours, blocked / row-wise variants reduce
to the plain cast with the documented scale."""
import numpy as np
import pytest

from oracle import fpemu_oracle as O

MODES = ["E5M2_RTZ", "E5M2_RNE", "E5M2_RNAZ", "E5M2_RNTZ", "E5M2_DAZ_RNE", "E4M3_RNE", "E4M3_RTZ", "E3M4_RNE", "E3M4_RTZ",
         "BFLOAT16_RNE", "BFLOAT16_RTZ", "FLOAT16_RNE", "FLOAT16_DAZ_RNE", "E4M3_IEEE_RNE"]
MANT = {"E5M2": 2, "E4M3": 3, "E3M4": 4, "BFLOAT16": 7, "FLOAT16": 10}


def data(n=1 << 15, seed=0):
    rng = np.random.default_rng(seed)
    return (rng.standard_normal(n) * np.exp2(rng.integers(-12, 6, n))).astype(np.float32)


@pytest.fixture(autouse=True)
def _cuda_nan_mode():
    O.set_nan_mode(False)
    yield


@pytest.mark.parametrize("mode", MODES)
def test_idempotent(mode):
    x = data()
    q = O.cast(x, mode)
    assert np.array_equal(O.cast(q, mode).view(np.uint32), q.view(np.uint32))


@pytest.mark.parametrize("mode", MODES)
def test_monotone(mode):
    x = np.sort(data(seed=1))
    q = O.cast(x, mode)
    assert (np.diff(q.astype(np.float64)) >= 0).all()


@pytest.mark.parametrize("mode", ["E5M2_RNE", "E4M3_RNE", "E3M4_RNE", "E5M2_RTZ", "E4M3_RTZ", "FLOAT16_RNE"])
def test_sign_symmetric(mode):
    x = data(seed=2)
    assert np.array_equal(O.cast(-x, mode).view(np.uint32), (-O.cast(x, mode)).view(np.uint32))


@pytest.mark.parametrize("mode", ["E5M2_RNE", "E4M3_RNE", "E3M4_RNE", "BFLOAT16_RNE"])
def test_nearest_is_within_half_an_ulp_and_truncation_shrinks(mode):
    fmt = mode.rsplit("_", 1)[0]
    x = data(seed=3)
    h = x.astype(np.float16).astype(np.float32) if fmt != "BFLOAT16" else x
    inside = (np.abs(h) > 0.3) & (np.abs(h) < 16)
    q = O.cast(x, mode)
    bound = np.exp2(np.floor(np.log2(np.abs(h[inside]))) - (MANT[fmt] + 1))
    assert (np.abs(q[inside] - h[inside]) <= bound).all()
    t = O.cast(x, fmt + "_RTZ")
    assert (np.abs(t[inside]) <= np.abs(h[inside])).all()
    assert (np.abs(t[inside] - h[inside]) < 2 * bound).all()


@pytest.mark.parametrize("mode", ["E5M2_STOCHASTIC", "E4M3_STOCHASTIC", "E3M4_STOCHASTIC", "BFLOAT16_STOCHASTIC"])
def test_stochastic_lands_on_a_neighbour_and_is_unbiased(mode):
    fmt = mode.rsplit("_", 1)[0]
    x = data(n=1 << 17, seed=4)
    h = x.astype(np.float16).astype(np.float32) if fmt != "BFLOAT16" else x
    inside = (np.abs(h) > 0.3) & (np.abs(h) < 16)
    lo = O.cast(x, fmt + "_RTZ")
    q = O.cast(x, mode, seed=11)
    ulp = np.exp2(np.floor(np.log2(np.abs(h[inside]))) - MANT[fmt])
    d = np.abs(q[inside]) - np.abs(lo[inside])
    assert ((d == 0) | (np.abs(d - ulp) <= 1e-12)).all()                   # the truncated value or one step further
    err = (q[inside] - h[inside]).astype(np.float64)
    assert abs(err.mean()) <= 5 * err.std() / np.sqrt(err.size)           # unbiased
    assert not np.array_equal(q, O.cast(x, mode, seed=12))                # and the seed matters


def test_blocked_equals_plain_cast_with_the_block_scale():
    rng = np.random.default_rng(5)
    x = (rng.standard_normal(4096) * np.repeat(np.exp2(rng.integers(-5, 5, 64)), 64)).astype(np.float32)
    q = O.cast_blocked(x, "E4M3_RNE", 64)
    for b in range(64):
        blk = x[b * 64:(b + 1) * 64]
        p2 = np.float32(2.0) ** np.floor(np.log2(np.abs(blk).max()))
        s = np.float32(2.0) * p2 / np.float32(8.0)                         # fpemu_kernels.cu:577-581
        assert np.array_equal(q[b * 64:(b + 1) * 64].view(np.uint32), O.cast(blk, "E4M3_RNE", scale=float(s)).view(np.uint32))


def test_rowwise_equals_per_row_amax_scaled_cast():
    rng = np.random.default_rng(6)
    x = (rng.standard_normal((16, 96)) * np.exp2(rng.integers(-4, 4, (16, 1)))).astype(np.float32)
    q, scales = O.cast_rowwise(x, "E4M3_RNE", 448.0)
    for r in range(16):
        s = (np.float32(1.0) / np.float32(np.abs(x[r]).max())) * np.float32(448.0)   # torch: vmax.reciprocal() * flt_max
        assert scales[r] == s
        assert np.array_equal(q[r].view(np.uint32), O.cast(x[r], "E4M3_RNE", scale=float(s)).view(np.uint32))
    assert O.scale_from_amax(0.0, 448.0) == np.float32(6.55e4)             # amax 0 -> inf -> clamp, qutils.py:285
    assert O.scale_from_amax(1e-3, 448.0) == np.float32(6.55e4)            # 4.48e5 > 3.275e4 -> 6.55e4


def test_scale_from_amax_is_what_torch_computes_for_flt_max_over_a_0dim_tensor():
    """qutils.py:283-285 on the reference's own arithmetic: `flt_max / vmax` with vmax a 0-dim tensor (torch evaluates
    Tensor.__rtruediv__ as reciprocal * flt_max in the tensor's dtype), then the clamp — float32 and float16, with a
    dense sweep around the clamp threshold (the comparison runs in half for half tensors)."""
    import torch
    rng = np.random.default_rng(3)
    for dtype, like in ((torch.float32, np.zeros(1, np.float32)), (torch.float16, np.zeros(1, np.float16))):
        for flt_max in (448.0, 57344.0, 30.0):
            edge = flt_max / 3.275e4
            v = np.concatenate([rng.uniform(1e-3, 900.0, 1500), np.linspace(edge * 0.97, edge * 1.03, 800), [0.0]])
            v = torch.from_numpy(v).to(dtype)
            for t in v:
                scale = flt_max / t
                scale = 6.55e+04 if scale > 3.275e+04 else scale
                got = O.scale_from_amax(float(t), flt_max, like=like)
                assert got == float(scale), (dtype, flt_max, float(t), got, float(scale))
                unclamped = O.scale_from_amax(float(t), flt_max, like=like, clamp=False)
                assert unclamped == float(flt_max / t)                      # fine-grained branch, qutils.py:327-328


def test_amax_propagates_nan_like_torch_max():
    import torch
    x = np.random.default_rng(1).standard_normal(1000).astype(np.float32)
    assert O.amax(x) == torch.max(abs(torch.from_numpy(x))).item()
    x[500] = np.nan
    assert np.isnan(O.amax(x)) and np.isnan(torch.max(abs(torch.from_numpy(x))).item())
    s = O.scale_from_amax(O.amax(x), 448.0)
    assert np.isnan(s)                                                      # NaN > 3.275e4 is False: no clamp
    assert np.isnan(O.cast(x, "E4M3_RNE", scale=s)).all() and np.isnan(O.cast(x, "E5M2_RNE", scale=s)).all()


def test_scaled_branches_of_quantize_tensor_equal_the_reference_stack_golden(golden_dir):
    """tests/golden/modules.npz holds outputs of the reference's own quantize_tensor (CPU op, fp32 un-scale); the
    oracle's row loop must reproduce the per-tensor, per-channel and fine-grained branches bit for bit."""
    import os
    G = np.load(os.path.join(golden_dir, "modules.npz"))
    O.set_nan_mode(True)
    try:
        w, a = G["qt_weight"], G["qt_act"]
        for key, x, rows, fmax, clamp in (("qt|e4m3|rne|per-channel|1|qt_weight", w, 24, 448.0, True),
                                          ("qt|e3m4|rne|per-channel|1|qt_weight", w, 24, 30.0, True),
                                          ("qt|e4m3|rne|fine-grained|4|qt_weight", w, 24 * 4, 448.0, False),
                                          ("qt|e4m3|rne|per-tensor|1|qt_act", a, 1, 448.0, True),
                                          ("qt|e5m2|rne|per-tensor|1|qt_act", a, 1, 57344.0, True)):
            mode = key.split("|")[1].upper() + "_RNE"
            q, _ = O.cast_rowwise(x.reshape(rows, -1), mode, fmax, unscale_fp32=True, clamp=clamp)
            assert np.array_equal(q.reshape(x.shape).view(np.uint32), G[key].view(np.uint32)), key
    finally:
        O.set_nan_mode(False)
