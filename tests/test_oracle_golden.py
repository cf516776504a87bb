"""The C restatement (oracle/fpemu_oracle.c) against golden vectors produced by the reference itself
(tests/golden/make_golden.py ran the unmodified reference CPU op).  CPU-only, bit-exact."""
import os

import numpy as np
import pytest

from oracle import fpemu_oracle as O

DET_MODES = ["E5M2_RTZ", "E5M2_RNE", "E5M2_RNAZ", "E5M2_RNTZ", "E5M2_RPINF", "E5M2_RNINF",
             "E5M2_DAZ_RNE", "E5M2_DAZ_RNAZ", "E5M2_DAZ_RNTZ", "E4M3_RNE", "E3M4_RNE",
             "FLOAT16_RNE", "FLOAT16_DAZ_RNE"]


@pytest.fixture(autouse=True)
def _x86_nan_mode():
    # the golden vectors came from the x86 build: float->half keeps the NaN sign there
    O.set_nan_mode(True)
    yield
    O.set_nan_mode(False)


@pytest.mark.parametrize("mode", DET_MODES)
def test_exhaustive_fp16_patterns(golden_dir, mode):
    g = np.load(os.path.join(golden_dir, "fp16_patterns.npz"))
    x = np.arange(65536, dtype=np.uint16).view(np.float16).astype(np.float32)
    want = g[mode].view(np.float16).astype(np.float32)
    got = O.cast(x, mode, scale=1.0, unscale_fp32=True)
    ok = O.same_bits_or_both_nan(got, want)
    assert ok.all(), "%s: %d mismatches, first at pattern 0x%04x" % (mode, (~ok).sum(), np.nonzero(~ok)[0][0])


@pytest.mark.parametrize("mode", DET_MODES)
def test_wide_inputs_all_scales(golden_dir, mode):
    g = np.load(os.path.join(golden_dir, "wide.npz"))
    x = g["input"].view(np.float32)
    seen = 0
    for key in g.files:
        if not key.startswith(mode + "@"):
            continue
        scale = float(key.split("@")[1])
        got = O.cast(x, mode, scale=scale, unscale_fp32=True)
        ok = O.same_bits_or_both_nan(got, g[key].view(np.float32))
        assert ok.all(), "%s scale %g: %d mismatches" % (mode, scale, (~ok).sum())
        seen += 1
    assert seen >= 1


def test_bf16_rne_matches_reference_avx512(golden_dir):
    g = np.load(os.path.join(golden_dir, "wide.npz"))
    if "BFLOAT16_RNE@1" not in g.files:
        pytest.skip("fixture generated on a host without AVX-512")
    x = g["bf16_input"].view(np.float32)
    got = O.cast(x, "BFLOAT16_RNE")
    want = g["BFLOAT16_RNE@1"].view(np.float32)
    # the CPU AVX-512 op leaves NaN payloads alone and then masks them like the CUDA text does
    ok = O.same_bits_or_both_nan(got, want)
    # zeros take the CUDA text's "denormal" branch (exponent field 0 -> `uf.u = 0`, fpemu_kernels.cu:273,298),
    # which drops the sign of -0.0; the AVX-512 CPU op keeps it.  Target = CUDA text, so skip x == 0 here.
    ok |= (x == 0)
    assert ok.all(), "%d mismatches" % (~ok).sum()
    assert O.cast(np.array([-0.0], dtype=np.float32), "BFLOAT16_RNE").view(np.uint32)[0] == 0


def test_bf16_flushes_fp32_denormals_like_cuda_text():
    x = np.array([1e-40, -1e-40, 1.17549435e-38, 1.1754942e-38], dtype=np.float32)
    got = O.cast(x, "BFLOAT16_RNE")
    assert got[0] == 0 and got[1] == 0 and not np.signbit(got[1])  # `uf.u = 0`, fpemu_kernels.cu:298
    assert got[2] == np.float32(1.17549435e-38)


@pytest.mark.parametrize("mode", ["E5M2_RNE", "E4M3_RNE", "E3M4_RNE", "E5M2_RTZ"])
@pytest.mark.parametrize("bs", [32, 64, 1024])
def test_blocked(golden_dir, mode, bs):
    g = np.load(os.path.join(golden_dir, "blocked.npz"))
    x = g["input"].view(np.float32)
    got = O.cast_blocked(x, mode, bs, unscale_fp32=True)
    ok = O.same_bits_or_both_nan(got, g["%s@%d" % (mode, bs)].view(np.float32))
    assert ok.all(), "%d mismatches" % (~ok).sum()


def test_stochastic_distribution_matches_reference(golden_dir):
    """P(round up) of the restated formula == the reference's observed frequency, and the two possible
    outputs are exactly the reference's two outputs (tolerance: 5 sigma of the binomial)."""
    g = np.load(os.path.join(golden_dir, "stochastic.npz"))
    N = int(g["N"][0])
    for key in g.files:
        if key == "N":
            continue
        mode, v = key.split("@")
        v = np.float32(float(v))
        ref_vals, ref_cnt = g[key][0].astype(np.uint32), g[key][1].astype(np.float64)
        x = np.full(N, v, dtype=np.float32)
        got = O.cast(x, mode, seed=99, unscale_fp32=True)
        u, c = np.unique(got.view(np.uint32), return_counts=True)
        assert set(u.tolist()) == set(ref_vals.tolist()), (key, u, ref_vals)
        for val, cnt in zip(u, c):
            p_ref = ref_cnt[list(ref_vals).index(val)] / N
            p = cnt / N
            sigma = np.sqrt(max(p_ref * (1 - p_ref), 1e-9) / N)
            assert abs(p - p_ref) <= 5 * sigma * np.sqrt(2), (key, p, p_ref)


def test_known_answers_survey_appendix_a():
    """SURVEY Appendix A rows (reference scalar CPU build, scale 1)."""
    O.set_nan_mode(True)
    def q(v, mode):
        return O.cast(np.array([v], dtype=np.float32), mode, unscale_fp32=True)[0]
    assert q(-61440.0, "E5M2_RNE") == -np.inf and q(-61440.0, "E5M2_RTZ") == -57344.0
    assert q(65504.0, "E5M2_RNE") == np.inf and q(65504.0, "E4M3_RNE") == 448.0 and q(65504.0, "E3M4_RNE") == 30.0
    assert q(-31.0, "E4M3_RNE") == -32.0 and q(-31.0, "E3M4_RNE") == -30.0
    assert q(0.1, "E4M3_RNE") == np.float32(0.1015625) and q(0.1, "E5M2_RNE") == np.float32(0.09375)
    assert q(0.0029296875, "E4M3_RNE") == np.float32(0.0029296875)       # not a true E4M3 value
    assert q(0.0009765625, "E4M3_RNE") == 0.0 and q(0.00146484375, "E4M3_RNE") == 0.0
    assert q(1.52587890625e-05, "E5M2_DAZ_RNE") == 0.0 and q(1.52587890625e-05, "E5M2_RNE") == np.float32(2.0 ** -16)
    z = q(-2.3020705e-06, "E5M2_DAZ_RNE")
    assert z == 0.0 and not np.signbit(z)                                # DAZ loses the sign
    assert np.signbit(q(-2.3020705e-06, "E5M2_RNE"))
    assert q(np.inf, "E4M3_RNE") == 448.0 and q(np.nan, "E4M3_RNE") == 448.0 and q(np.nan, "E3M4_RNE") == 30.0
    assert np.isnan(q(np.nan, "E5M2_RNE"))
    # scale 3.7, CPU un-scale: 448/3.7 in fp32
    got = O.cast(np.array([57344.0], dtype=np.float32), "E4M3_RNE", scale=3.7, unscale_fp32=True)
    assert got.view(np.uint32)[0] == 0x42F22983


def test_fp16_unscale_is_cpu_result_rounded_through_half():
    """CUDA semantics (fpemu_kernels.cu:652) == CPU semantics followed by an fp16 round trip."""
    rng = np.random.default_rng(3)
    x = rng.standard_normal(1 << 14).astype(np.float32)
    for mode in ("E4M3_RNE", "E5M2_RNE", "E3M4_RNE"):
        cpu = O.cast(x, mode, scale=94.09, unscale_fp32=True)
        cuda = O.cast(x, mode, scale=94.09, unscale_fp32=False)
        assert np.array_equal(cpu.astype(np.float16).astype(np.float32), cuda)


def test_philox_known_answer():
    # Random123 kat_vectors: philox4x32-10, counter = key = 0
    assert O.philox(0, 0) == [0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8]


def test_rand_bits_partition_independent():
    a = O.rand_bits(5, 0, 1000, O.FMT["E4M3"])
    b = np.concatenate([O.rand_bits(5, 0, 333, O.FMT["E4M3"]), O.rand_bits(5, 333, 667, O.FMT["E4M3"])])
    assert np.array_equal(a, b)
    a = O.rand_bits(5, 0, 1000, O.FMT["BF16"])
    b = np.concatenate([O.rand_bits(5, 0, 501, O.FMT["BF16"]), O.rand_bits(5, 501, 499, O.FMT["BF16"])])
    assert np.array_equal(a, b)


def test_reference_64_value_fixture(golden_dir):
    """The input vector of the reference's own test script (mpemu/pytquant/test.py:120-135) through every
    deterministic mode at scale 1 and 3.7, BFLOAT16 (AVX-512 build) and blocked mode with 16-element blocks."""
    g = np.load(os.path.join(golden_dir, "fixture64.npz"))
    x = g["input"].view(np.float32)
    assert x.size == 64 and x[0] == 57344.0 and x[-1] == np.float32(2.1336775e-06)
    seen = 0
    for key in g.files:
        if key == "input":
            continue
        mode, arg = key.split("@")
        want = g[key].view(np.float32)
        if arg.startswith("block"):
            got = O.cast_blocked(x, mode, int(arg[5:]), unscale_fp32=True)
        else:
            got = O.cast(x, mode, scale=float(arg), unscale_fp32=True)
        assert O.same_bits_or_both_nan(got, want).all(), key
        seen += 1
    assert seen >= 2 * 11 + 2 + 3
