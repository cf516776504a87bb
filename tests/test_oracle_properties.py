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
        s = np.float32(448.0) / np.float32(np.abs(x[r]).max())
        assert scales[r] == s
        assert np.array_equal(q[r].view(np.uint32), O.cast(x[r], "E4M3_RNE", scale=float(s)).view(np.uint32))
    assert O.scale_from_amax(0.0, 448.0) == np.float32(6.55e4)             # amax 0 -> inf -> clamp, qutils.py:285
    assert O.scale_from_amax(1e-3, 448.0) == np.float32(6.55e4)            # 4.48e5 > 3.275e4 -> 6.55e4
