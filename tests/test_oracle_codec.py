"""8-bit codes (oracle/fpemu_oracle.c: fpemu_oracle_encode / _decode) pinned two ways, CPU-only:
  * the reference's own format table (README.md:17-27) as known answers;
  * the reference's rounded VALUES (tests/golden/fp16_patterns.npz, produced by the unmodified reference CPU op):
    decode(encode(x)) must reproduce them bit for bit wherever the value is a member of the format, and be the
    nearest member (ties to the even code) where the reference keeps more precision than the format has.
The nearest-member search below is an independent numpy implementation (table look-up, no bit tricks)."""
import os

import numpy as np
import pytest

from oracle import fpemu_oracle as O

FORMATS = ["E5M2", "E4M3", "E3M4"]

# README.md:17-27: (max normal, min normal, max subnormal, min subnormal) as (code, value)
TABLE = {
    "E5M2": ((0x7B, 57344.0), (0x04, 2.0 ** -14), (0x03, 0.75 * 2.0 ** -14), (0x01, 2.0 ** -16)),
    "E4M3": ((0x7E, 448.0), (0x08, 2.0 ** -6), (0x07, 0.875 * 2.0 ** -6), (0x01, 2.0 ** -9)),
    "E3M4": ((0x7E, 30.0), (0x10, 2.0 ** -2), (0x0F, 0.9375 * 2.0 ** -2), (0x01, 2.0 ** -6)),
}


def decode_table(fmt):
    return O.decode(np.arange(256, dtype=np.uint8), fmt)


@pytest.mark.parametrize("fmt", FORMATS)
def test_format_table_known_answers(fmt):
    t = decode_table(fmt)
    for code, value in TABLE[fmt]:
        assert t[code] == np.float32(value), "%s code 0x%02x" % (fmt, code)
        assert t[code | 0x80] == np.float32(-value)
        assert O.encode(np.array([value, -value], dtype=np.float32), fmt + "_RNE").tolist() == [code, code | 0x80]
    assert t[0x00] == 0.0 and not np.signbit(t[0x00])
    assert t[0x80] == 0.0 and np.signbit(t[0x80])
    if fmt == "E5M2":                                   # S.11111.00 = Inf, S.11111.{01,10,11} = NaN
        assert t[0x7C] == np.inf and t[0xFC] == -np.inf
        assert np.isnan(t[[0x7D, 0x7E, 0x7F, 0xFD, 0xFE, 0xFF]]).all()
    else:                                               # S.1111.111 / S.111.1111 = NaN, no infinities
        assert np.isnan(t[[0x7F, 0xFF]]).all() and np.isfinite(np.delete(t, [0x7F, 0xFF])).all()
    # strictly increasing in the magnitude code: a valid sign-magnitude number line
    mags = t[:0x7C] if fmt == "E5M2" else t[:0x7F]
    assert (np.diff(mags) > 0).all()


@pytest.mark.parametrize("fmt", FORMATS)
def test_every_code_survives_a_round_trip(fmt):
    t = decode_table(fmt)
    keep = ~np.isnan(t)
    for rnd in ("RNE", "RTZ", "STOCHASTIC"):
        back = O.encode(t, "%s_%s" % (fmt, rnd), seed=5)
        assert (back[keep] == np.arange(256)[keep]).all(), rnd


def nearest_code(values, fmt):
    """Independent encoder: nearest member of the decode table, ties to the even code, sign kept."""
    t = decode_table(fmt).astype(np.float64)
    top = 0x7C if fmt == "E5M2" else 0x7F               # magnitude codes below this are finite
    mags = t[:top]
    a = np.abs(values.astype(np.float64))
    hi = np.clip(np.searchsorted(mags, a, side="left"), 0, top - 1)
    lo = np.clip(hi - 1, 0, top - 1)
    dlo, dhi = np.abs(a - mags[lo]), np.abs(mags[hi] - a)
    pick_hi = (dhi < dlo) | ((dhi == dlo) & (hi % 2 == 0))
    code = np.where(pick_hi, hi, lo).astype(np.uint8)
    return code | (np.signbit(values).astype(np.uint8) << 7)


@pytest.fixture
def x86_nan_mode():
    # the golden vectors came from the x86 build: float->half keeps the NaN sign there
    O.set_nan_mode(True)
    yield
    O.set_nan_mode(False)


@pytest.mark.parametrize("mode", ["E5M2_RNE", "E5M2_RTZ", "E5M2_RNAZ", "E5M2_DAZ_RNE", "E4M3_RNE", "E3M4_RNE"])
def test_codes_reproduce_the_reference_values(golden_dir, x86_nan_mode, mode):
    fmt = mode.split("_")[0]
    g = np.load(os.path.join(golden_dir, "fp16_patterns.npz"))
    x = np.arange(65536, dtype=np.uint16).view(np.float16).astype(np.float32)
    ref = g[mode].view(np.float16).astype(np.float32)             # the reference's rounded values
    codes = O.encode(x, mode)
    back = O.decode(codes, fmt)
    t = decode_table(fmt)
    finite = np.isfinite(ref)
    member = np.isin(ref, t[~np.isnan(t)])
    # 1. members of the format come back bit for bit (NaN stays NaN, Inf stays Inf for E5M2)
    assert O.same_bits_or_both_nan(back[member | ~finite], ref[member | ~finite]).all()
    # 2. everything finite is coded as the nearest member
    want = nearest_code(ref[finite], fmt)
    assert (codes[finite] == want).all(), "first mismatch at pattern 0x%04x" % np.nonzero(finite)[0][np.nonzero(codes[finite] != want)[0][0]]
    # 3. which values are not members: only the reference's pseudo-subnormals
    if fmt == "E5M2":
        assert member[finite].all()
    else:
        min_normal = TABLE[fmt][1][1]
        assert (np.abs(ref[finite & ~member]) < min_normal).all()
        assert (finite & ~member).sum() > 0


@pytest.mark.parametrize("fmt", FORMATS)
@pytest.mark.parametrize("scale", [1.0, 0.125, 3.7])
def test_scaled_round_trip_equals_the_cast(fmt, scale):
    rng = np.random.default_rng(11)
    x = (rng.standard_normal(20000) * np.exp2(rng.integers(-8, 6, 20000))).astype(np.float32)
    mode = fmt + "_RNE"
    for unscale_fp32 in (False, True):
        q = O.cast(x, mode, scale=scale, unscale_fp32=unscale_fp32)
        back = O.decode(O.encode(x, mode, scale=scale), fmt, scale=scale, unscale_fp32=unscale_fp32)
        same = O.same_bits_or_both_nan(back, q)
        if fmt == "E5M2":
            assert same.all()
        else:
            min_normal = TABLE[fmt][1][1]
            assert same[np.abs(x * np.float32(scale)) >= 2 * min_normal].all()
            assert same.mean() > 0.5


def test_stochastic_codes_follow_the_cast_draws():
    rng = np.random.default_rng(3)
    x = rng.standard_normal(50000).astype(np.float32)
    for fmt in FORMATS:
        mode = fmt + "_STOCHASTIC"
        q = O.cast(x, mode, seed=99, elem_offset=32)
        back = O.decode(O.encode(x, mode, seed=99, elem_offset=32), fmt)
        same = O.same_bits_or_both_nan(back, q)
        assert same[np.abs(x) >= 2 * TABLE[fmt][1][1]].all()
        if fmt == "E5M2":
            assert same.all()
