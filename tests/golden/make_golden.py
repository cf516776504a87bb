#!/usr/bin/env python
"""Generate tests/golden/*.npz by RUNNING THE REFERENCE ITSELF (unmodified CPU builds in oracle/_ref/,
compiled by oracle/build_ref.py from /root/reference/mpemu/pytquant/cpp/fpemu_impl.cpp).

Run in the build container only (it needs oracle/_ref/, i.e. /root/reference):
    python oracle/build_ref.py cpu && python tests/golden/make_golden.py

Which build produced which vector:
  * scalar build  -> every E5M2 / E5M2_DAZ / E4M3 / E3M4 / FLOAT16 mode (it is textually the algorithm of
                     the CUDA kernels and honours every rounding mode; the AVX-512 build sends all
                     deterministic modes to RNE, cpp/fpemu_impl.cpp:818-824).
  * AVX-512 build -> BFLOAT16_RNE on inputs without fp32 denormals (the scalar bf16 path mis-rounds,
                     cpp/fpemu_impl.cpp:321-324; the CUDA text flushes denormals, cuda/fpemu_kernels.cu:296-299).
All outputs use the CPU op's un-scale semantics (fp32 multiply by 1/scale).

Files:
  fp16_patterns.npz   outputs for all 65 536 fp16 bit patterns (as fp32 inputs), scale 1.0, stored as fp16 bits
  wide.npz            4096 wide-range fp32 inputs (specials included) x modes x scales {1, 0.125, 3.7, 94.09}
  blocked.npz         blocked (block_norm) outputs for block sizes {32, 64, 1024}
  stochastic.npz      round-up counts of repeated draws for a few probe values (distribution pin)
  modules.npz         per-tensor / per-channel quantize_tensor outputs of the reference qutils (CPU)
  fixture64.npz       the reference's own 64-value test vector (mpemu/pytquant/test.py:120-135) x modes x scales {1, 3.7}
                      (`python tests/golden/make_golden.py fixture64` regenerates only this file)
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ref_loader  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))

DET_MODES = ["E5M2_RTZ", "E5M2_RNE", "E5M2_RNAZ", "E5M2_RNTZ", "E5M2_RPINF", "E5M2_RNINF",
             "E5M2_DAZ_RNE", "E5M2_DAZ_RNAZ", "E5M2_DAZ_RNTZ", "E4M3_RNE", "E3M4_RNE",
             "FLOAT16_RNE", "FLOAT16_DAZ_RNE"]
SCALES = [1.0, 0.125, 3.7, 94.09]


def boundary_values():
    """Values straddling every threshold of the formats (max, saturation, flush, pseudo-subnormal edges)."""
    v = []
    for base in (57344.0, 61440.0, 65504.0, 65520.0, 448.0, 464.0, 480.0, 240.0, 30.0, 31.0, 28.0, 15.5, 1.0,
                 2.0 ** -6, 2.0 ** -7, 2.0 ** -9, 2.0 ** -10, 2.0 ** -14, 2.0 ** -16, 2.0 ** -17, 2.0 ** -24, 2.0 ** -25,
                 1.5 * 2.0 ** -9, 1.5 * 2.0 ** -10, 1.75 * 2.0 ** -7, 1.0625, 1.03125, 1.046875, 0.1, 7.6505613):
        for k in (1.0, 1.0 + 2.0 ** -11, 1.0 - 2.0 ** -12, 1.0 + 2.0 ** -4, 1.0 + 2.0 ** -3 + 2.0 ** -10):
            v += [base * k, -base * k]
    v += [0.0, -0.0, np.inf, -np.inf, np.nan, 1e-40, -1e-40, 3.4e38, -3.4e38, 1e-45]
    return np.array(v, dtype=np.float32)


def wide_inputs(n=4096, seed=1234):
    rng = np.random.default_rng(seed)
    x = rng.standard_normal(n).astype(np.float32) * np.exp2(rng.integers(-30, 20, n)).astype(np.float32)
    b = boundary_values()
    x[:len(b)] = b
    return x


def reference_fixture64():
    """The input vector of the reference's only test script (mpemu/pytquant/test.py:120-135): 64 float32 values chosen
    around the formats' maxima, saturation points and flush thresholds.  That script prints input next to output and
    asserts nothing, so the expected outputs are produced here by running the reference op on it."""
    return np.array([57344.00, -61440.0, 65504.0, -500.0, 448.0, -480.0, 30.0, -31.0,
                     26.0, 15.0, -7.6505613e-00, 5.9832452e-00, 1.5625032e-02, 3.1725775e-02, 4.3268750e-02, 6.2655000e-02,
                     1.9545313e-03, 3.9045625e-03, 5.8845638e-03, 7.8089750e-03, -6.0151856e-02, 6.9784373e-03,
                     -1.6634936e+03, -7.6505613e-04, 1.9545313e-03, 3.9045625e-03, 5.8845638e-03, 7.8089750e-03,
                     1.5629950e-02, 1.5225650e-02, -3.1256500e-02, 9.3857500e-02, 7.9284608e-01, 2.8815269e-01,
                     -1.1787039e-01, 6.1035156e-05, -1.1787039e-06, -1.0749481e-01, -1.3085605e+00, 6.5981364e-01,
                     -7.0325255e-02, 2.7448297e-01, 5.5694544e-01, -2.3220782e-01, -5.9746221e-02, 15.23213444,
                     -0.00004323, 1.9767435e-04, -1.2203161e+00, 2.9099861e-01, -7.9642259e-02, 1.3200364e+00,
                     -1.5196867e+00, -1.2530587e+00, -2.0159689e-03, -1.9767643e+00, 6.0834163e-04, 7.8943473e-05,
                     7.8247029e-04, -6.4658634e-05, -2.3020705e-06, -1.5630834e-05, -7.4762434e-07, 2.1336775e-06],
                    dtype=np.float32)


def write_fixture64(scalar, avx):
    x = reference_fixture64()
    assert x.size == 64
    d = {"input": x.view(np.uint32)}
    for mode in DET_MODES:
        for s in (1.0, 3.7):
            if mode.startswith("FLOAT16") and s != 1.0:
                continue
            d["%s@%g" % (mode, s)] = run(scalar, x, mode, s).view(np.uint32)
    if avx is not None:
        d["BFLOAT16_RNE@1"] = run(avx, x, "BFLOAT16_RNE").view(np.uint32)    # 64 = 4 AVX-512 vectors, no denormals
    for mode in ("E5M2_RNE", "E4M3_RNE", "E3M4_RNE"):
        d["%s@block16" % mode] = run(scalar, x, mode, 1.0, True, 16).view(np.uint32)
    np.savez_compressed(os.path.join(OUT, "fixture64.npz"), **d)


def run(ref, x, mode, scale=1.0, block_norm=False, block_size=1):
    t = torch.from_numpy(np.ascontiguousarray(x).copy())
    return ref.forward(t, mode, t.numel(), False, float(scale), block_norm, int(block_size))[0].numpy()


def main():
    torch.set_num_threads(1)
    scalar = ref_loader.load("fpemu_cpp_ref_scalar")
    avx = ref_loader.load("fpemu_cpp_ref_avx512")
    assert scalar is not None, "run oracle/build_ref.py cpu first"
    write_fixture64(scalar, avx)
    if sys.argv[1:] == ["fixture64"]:
        return

    # 1. exhaustive fp16 patterns, scale 1: outputs are fp16-representable -> keep 16 bits
    pat = np.arange(65536, dtype=np.uint16).view(np.float16).astype(np.float32)
    d = {}
    for mode in DET_MODES:
        o = run(scalar, pat, mode)
        o16 = o.astype(np.float16)
        back = o16.astype(np.float32)
        assert np.array_equal(back.view(np.uint32)[~np.isnan(o)], o.view(np.uint32)[~np.isnan(o)]), mode
        d[mode] = o16.view(np.uint16)
    np.savez_compressed(os.path.join(OUT, "fp16_patterns.npz"), **d)

    # 2. wide-range fp32 inputs x scales
    x = wide_inputs()
    d = {"input": x.view(np.uint32)}
    for mode in DET_MODES:
        for s in SCALES:
            if mode.startswith("FLOAT16") and s != 1.0:
                continue
            d["%s@%g" % (mode, s)] = run(scalar, x, mode, s).view(np.uint32)
    if avx is not None:
        xn = x.copy()
        den = (np.abs(xn) < np.float32(2.0 ** -126)) & (xn != 0)
        xn[den] = 0.0
        n16 = (xn.size // 16) * 16           # AVX-512 body only (the tail goes to the mis-rounding scalar path)
        d["bf16_input"] = xn[:n16].view(np.uint32)
        d["BFLOAT16_RNE@1"] = run(avx, xn[:n16], "BFLOAT16_RNE").view(np.uint32)
    np.savez_compressed(os.path.join(OUT, "wide.npz"), **d)

    # 3. blocked mode (scalar build; block sizes not multiples of 32 would also take the scalar path on AVX builds)
    rng = np.random.default_rng(7)
    xb = (rng.standard_normal(4096).astype(np.float32) *
          np.repeat(np.exp2(rng.integers(-8, 8, 128)).astype(np.float32), 32))
    d = {"input": xb.view(np.uint32)}
    for mode in ("E5M2_RNE", "E4M3_RNE", "E3M4_RNE", "E5M2_RTZ"):
        for bs in (32, 64, 1024):
            d["%s@%d" % (mode, bs)] = run(scalar, xb, mode, 1.0, True, bs).view(np.uint32)
    np.savez_compressed(os.path.join(OUT, "blocked.npz"), **d)

    # 4. stochastic distribution: N draws of single probe values (first use of the RNG state per process is
    #    irrelevant: only the round-up frequency is pinned)
    N = 1 << 18
    probes = {"E5M2_STOCHASTIC": [1.0625, 1.25 + 2.0 ** -9, -3.0 - 2.0 ** -7, 2.0 ** -15 * 1.3],
              "E4M3_STOCHASTIC": [1.03125, -1.0 - 2.0 ** -8, 300.5, 2.0 ** -8 * 1.1],
              "E3M4_STOCHASTIC": [1.046875, -1.0 - 2.0 ** -9, 25.3, 2.0 ** -5 * 1.01]}
    d = {}
    for mode, vals in probes.items():
        for v in vals:
            xin = np.full(N, v, dtype=np.float32)
            o = run(scalar, xin, mode)
            u, c = np.unique(o.view(np.uint32), return_counts=True)
            d["%s@%r" % (mode, float(np.float32(v)))] = np.stack([u.astype(np.uint64), c.astype(np.uint64)])
    d["N"] = np.array([N], dtype=np.uint64)
    np.savez_compressed(os.path.join(OUT, "stochastic.npz"), **d)
    print("golden fixtures written to", OUT)


if __name__ == "__main__":
    main()
