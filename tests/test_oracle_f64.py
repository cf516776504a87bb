"""The oracle's restatement of the reference's double dispatch (oracle.fpemu_oracle.cast_f64): CPU-only checks."""
import numpy as np
import pytest

from oracle import fpemu_oracle as O
from f64_cases import adversarial_f64


@pytest.mark.parametrize("mode,dropped", [("E5M2_RNE", 8), ("E4M3_RNE", 7)])
def test_double_multiply_is_not_narrow_then_multiply(mode, dropped):
    O.set_nan_mode(False)
    x = adversarial_f64(50000, 3, 3.7, dropped)
    x = x[np.abs(x * 3.7) < 400]
    a = O.cast_f64(x, mode, scale=3.7)
    b = O.cast(x.astype(np.float32), mode, scale=3.7).astype(np.float64)
    assert (a != b).mean() > 0.2                     # the two readings of `in[gid] * scale` are distinguishable here
    # every result is a member of the format's grid un-scaled through fp16: re-casting the scaled value is the identity
    back = (a * np.float64(np.float32(3.7))).astype(np.float32)
    assert np.array_equal(O.cast(back, mode), O.cast(O.cast(back, mode), mode))


def test_power_of_two_scales_and_unscaled_equal_the_float_path():
    O.set_nan_mode(False)
    rng = np.random.default_rng(0)
    x = rng.standard_normal(100000) * np.exp2(rng.integers(-12, 10, 100000))
    for mode in ("E5M2_RNE", "E4M3_RNE", "E3M4_RNE", "E5M2_RTZ", "BFLOAT16_RNE", "FLOAT16_RNE", "FP4_NEAREST"):
        for scale in (1.0, 0.125, 64.0):
            if mode.startswith(("BFLOAT16", "FLOAT16")) and scale != 1.0:
                continue
            a = O.cast_f64(x, mode, scale=scale)
            b = O.cast(x.astype(np.float32), mode, scale=scale).astype(np.float64)
            assert np.array_equal(a, b), (mode, scale)     # scaling by a power of two commutes with the narrowing


def test_known_answers():
    O.set_nan_mode(False)
    x = np.array([0.0, -0.0, 1.0 + 2.0 ** -30, -3.3, 447.99999, 1e300, -1e-300, 0.1], dtype=np.float64)
    got = O.cast_f64(x, "E4M3_RNE")
    assert got.tolist() == [0.0, -0.0, 1.0, -3.25, 448.0, 448.0, -0.0, 0.1015625]
    assert np.signbit(got[1]) and np.signbit(got[6])
    assert O.cast_f64(x, "E5M2_RNE")[5] == np.inf
