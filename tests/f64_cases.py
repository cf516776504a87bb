"""Inputs on which the reference's double path differs from narrowing first: `x * scale` computed in double lands just
beside a tie of the fp16 rounding whose two neighbours fall on different sides of the FORMAT's rounding boundary, so
(float)(x * (double)scale) and (float)x * scale end up one 8-bit grid step apart for a large share of the elements."""
import numpy as np


def adversarial_f64(n, seed, scale, dropped_bits):
    """dropped_bits: 8 (E5M2), 7 (E4M3), 6 (E3M4) low fp16 mantissa bits the format drops."""
    rng = np.random.default_rng(seed)
    hi = rng.integers(0x2400 >> dropped_bits, 0x5800 >> dropped_bits, n).astype(np.uint16) | 1      # kept LSB odd
    hb = ((hi << dropped_bits) | ((1 << (dropped_bits - 1)) - 1)).astype(np.uint16)                  # dropped bits 0111..1
    h = hb.view(np.float16).astype(np.float64)
    tie = h + np.exp2(np.floor(np.log2(h)) - 11)                                                      # an fp16 tie
    delta = rng.choice([-1.0, 1.0], n) * np.exp2(-rng.integers(26, 34, n).astype(np.float64))
    x = tie / np.float64(np.float32(scale)) * (1 + delta)
    return x * rng.choice([-1.0, 1.0], n)
