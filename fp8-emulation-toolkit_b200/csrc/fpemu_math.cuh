// fpemu_math.cuh — register-level arithmetic of the B200 fpemu cast.
//
// Two implementations of the same rounding rules live here and are tested against each other and
// against the CPU restatement used by the tests:
//   * round_pair_*  : two fp16 values packed in one 32-bit register, branch-free, built from LOP3 /
//                     IADD3 / HSET2 / HMNMX2 so that one warp instruction serves 64 elements.  These are
//                     the hot path (RNE / RTZ / stochastic for E5M2, E5M2_DAZ, E4M3, E3M4).
//   * round_one     : one element, every format and rounding mode of the reference's mode grammar.
//                     Used for tails, unaligned buffers and the rarely used modes.
//
// Semantics follow the reference CUDA kernels (mpemu/pytquant/cuda/fpemu_kernels.cu): fp32 -> fp16 with
// round-to-nearest-even first, then integer rounding of the fp16 encoding:
//   E5M2      :375-463   keep 2 mantissa bits; carry may reach Inf; NaN/Inf only masked
//   E5M2_DAZ  :465-541   same, fp16 subnormals (and zeros) become +0
//   E4M3      :543-654   |h| >= 448 (Inf, NaN too) -> +-448; |h| < 2^-9 -> +-0; for 2^-9 <= |h| < 2^-6 the
//                        low (9 - e) bits are cleared first, i.e. |h| is truncated to a multiple of 2^-16
//   E3M4      :873-984   |h| >= 30 -> +-30; |h| < 2^-6 -> +-0; |h| truncated to a multiple of 2^-12
//   BFLOAT16  :243-305   on fp32 bits; exponent field 0 -> +0
//   FLOAT16   :307-373
//   E4M3_IEEE :656-767, FP4 :986-1024
#pragma once

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>

#include "../../include/fpemu_b200.h"

namespace fpemu {

// ------------------------------------------------------------------------------------------ bit casts
__device__ __forceinline__ __half2 as_h2(uint32_t u) { return *reinterpret_cast<__half2*>(&u); }
__device__ __forceinline__ uint32_t as_u32(__half2 h) { return *reinterpret_cast<uint32_t*>(&h); }
__device__ __forceinline__ uint32_t as_u32(__nv_bfloat162 h) { return *reinterpret_cast<uint32_t*>(&h); }

__device__ __forceinline__ uint32_t pack_f16x2_rn(float lo, float hi) { return as_u32(__floats2half2_rn(lo, hi)); }
__device__ __forceinline__ uint32_t pack_bf16x2_rn(float lo, float hi) { return as_u32(__floats2bfloat162_rn(lo, hi)); }
__device__ __forceinline__ float2 unpack_f16x2(uint32_t h2) { return __half22float2(as_h2(h2)); }
__device__ __forceinline__ float2 unpack_bf16x2(uint32_t b2) {
  return make_float2(__uint_as_float(b2 << 16), __uint_as_float(b2 & 0xFFFF0000u));
}

// ------------------------------------------------------------------------------------------ Philox4x32-10
// Counter-based generator (Salmon et al., SC'11).  key = 64-bit seed, counter = (ctr, 0, 0).

__device__ __forceinline__ uint4 philox4x32_10(uint32_t seed_lo, uint32_t seed_hi, uint64_t ctr) {
  uint32_t c0 = (uint32_t)ctr, c1 = (uint32_t)(ctr >> 32), c2 = 0u, c3 = 0u;
  uint32_t k0 = seed_lo, k1 = seed_hi;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    uint64_t p0 = (uint64_t)0xD2511F53u * c0;
    uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
    uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
    uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
    c1 = (uint32_t)p1; c3 = (uint32_t)p0; c0 = n0; c2 = n2;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  return make_uint4(c0, c1, c2, c3);
}

// Random bits of ONE element with global index g (slow path; the vector path draws 16 / 8 elements per call).
// 8-bit formats: byte (g % 16) of Philox(g / 16);  BF16 / FP16: 16-bit lane (g % 8) of Philox(g / 8).
__device__ __forceinline__ uint32_t rand_bits_one(uint32_t seed_lo, uint32_t seed_hi, uint64_t g, bool wide) {
  uint4 w = philox4x32_10(seed_lo, seed_hi, wide ? (g >> 3) : (g >> 4));
  uint32_t words[4] = {w.x, w.y, w.z, w.w};
  if (wide) {
    uint32_t lane = (uint32_t)g & 7u;
    return (words[lane >> 1] >> ((lane & 1u) * 16u)) & 0xFFFFu;
  }
  uint32_t lane = (uint32_t)g & 15u;
  return (words[lane >> 2] >> ((lane & 3u) * 8u)) & 0xFFu;
}

// ------------------------------------------------------------------------------------------ packed pairs
// All functions take/return two fp16 encodings in one register.  `r2` carries the random bits of the
// two lanes in bits [0,8) and [16,24) (already masked to the format's dropped-bit count).

constexpr uint32_t kAbs2 = 0x7FFF7FFFu;
constexpr uint32_t kSign2 = 0x80008000u;

__device__ __forceinline__ uint32_t h2c(uint16_t bits) { return (uint32_t)bits | ((uint32_t)bits << 16); }

template <int RND, bool DAZ>
__device__ __forceinline__ uint32_t round_pair_e5m2(uint32_t h, uint32_t r2) {
  constexpr uint32_t kKeep = 0xFF00FF00u;
  const __half2 hv = as_h2(h);
  const __half2 av = __habs2(as_h2(h));                     // |h| as an operand modifier of its consumers
  const __half2 min_normal = as_h2(h2c(0x0400));
  if (RND == FPEMU_RND_RTZ && !DAZ) return h & kKeep;
  if (!DAZ) {
    if (RND == FPEMU_RND_RNE) {
      // h + 0x7F + lsb, NaN lanes untouched (Inf + 0x7F stays Inf after the mask)
      uint32_t ok = __heq2_mask(hv, hv);
      uint32_t lsb = (h >> 8) & 0x00010001u & ok;
      return (h + (0x007F007Fu & ok) + lsb) & kKeep;
    }
    if (RND == FPEMU_RND_RNAZ || RND == FPEMU_RND_RNTZ || RND == FPEMU_RND_RPINF || RND == FPEMU_RND_RNINF) {
      // round_up_bit(): RNAZ = dropped >= half, RNTZ = dropped > half, RPINF / RNINF = dropped >= half on positive /
      // negative lanes only — a constant added below the kept byte, fp16 subnormals included; NaN lanes (which
      // compare false with everything) are only masked, Inf + 0x80 stays Inf after the mask
      const __half2 zero = as_h2(0u);
      const uint32_t lanes = (RND == FPEMU_RND_RPINF) ? __hgt2_mask(hv, zero)
                           : (RND == FPEMU_RND_RNINF) ? __hlt2_mask(hv, zero) : __heq2_mask(hv, hv);
      const uint32_t inc = (RND == FPEMU_RND_RNTZ) ? 0x007F007Fu : 0x00800080u;
      return (h + (inc & lanes)) & kKeep;
    }
    // stochastic: fp16-normal lanes add 8 random bits, fp16-subnormal lanes round to nearest even
    uint32_t is_sub = __hlt2_mask(av, min_normal);
    uint32_t is_nrm = __hlt2_mask(av, as_h2(h2c(0x7C00))) & ~is_sub;
    uint32_t rne = 0x007F007Fu + ((h >> 8) & 0x00010001u);
    uint32_t inc = (r2 & is_nrm) | (rne & is_sub);
    return (h + inc) & kKeep;
  }
  // DAZ: only fp16-normal lanes are rounded; exponent-field-0 lanes (zeros included) become +0
  uint32_t is_sub = __hlt2_mask(av, min_normal);            // false for NaN: NaN lanes are only masked
  uint32_t is_nrm = __hge2_mask(av, min_normal);            // true for Inf: harmless, low bits are 0
  uint32_t inc;
  if (RND == FPEMU_RND_RNE) inc = (0x007F007Fu + ((h >> 8) & 0x00010001u)) & is_nrm;
  else if (RND == FPEMU_RND_STOCHASTIC) inc = r2 & is_nrm & __hlt2_mask(av, as_h2(h2c(0x7C00)));
  else if (RND == FPEMU_RND_RNAZ) inc = 0x00800080u & is_nrm;
  else if (RND == FPEMU_RND_RNTZ) inc = 0x007F007Fu & is_nrm;
  else inc = 0u;
  return (h + inc) & kKeep & ~is_sub;
}

// clamp(h, -c, +c) per lane: min first so that NaN lanes become +c (min / max return the non-NaN operand).
__device__ __forceinline__ uint32_t clamp_h2(uint32_t h, uint32_t c) {
  uint32_t r;
  asm("{.reg .b32 t, n; min.f16x2 t, %1, %2; neg.f16x2 n, %2; max.f16x2 %0, t, n;}" : "=r"(r) : "r"(h), "r"(c));
  return r;
}

// E4M3 (LSHIFT 7, SAT 0x5F00 = 448, FLUSH 2^-9, pseudo-subnormals below 2^-6) and
// E3M4 (LSHIFT 6, SAT 0x4F80 = 30,  FLUSH 2^-6, pseudo-subnormals below 2^-2) share one shape.
//
// The reference clears the low d = G - e mantissa bits of a value in binade e < G (G = 9 for E4M3, 13 for
// E3M4) before rounding, i.e. truncates |h| to a multiple of 2^(G-25), and flushes binades e < G - 3 (E4M3) /
// e < G - 4 (E3M4).  Both per-lane decisions are derived from p = 2^(G-e), which as an fp16 number is one
// integer subtraction on the exponent field:
//   mask   2^d - 1 = relu(p - 1), moved into integer bits by adding 1024 (HFMA2.RELU + HADD2, FMA pipe);
//   flush  factor sat(B - p) in {0, 1} with B = 9 (E4M3: keep p <= 8) / 17 (E3M4: keep p <= 16), applied with a
//          final HMUL2 so the sign of a flushed value survives (HFMA2.SAT + HMUL2, FMA pipe).
// The stochastic and truncating casts work on the signed encoding (sign-magnitude: adding to the low bits never reaches
// bit 15), so they need no abs / sign-merge.  Per pair: 5 ALU-pipe + 7 FMA-pipe instructions for RNE (the rounding itself
// is two HFMA2: see below); the ALU pipe, which issues every other cycle, is what bounds the 16-bit tensor variants.
// MAG = 1 (the encode kernels, which take the sign from h itself): the result is the rounded MAGNITUDE, and the clamp is a
// single min(|h|, max) instead of min + max.  MAG = 2: the same with h's sign put back before the flush (the RNE cast: its
// bit-select for the sign of zero becomes that merge, so the cheaper clamp is a net saving; the stochastic and truncating
// casts would pay an extra merge and keep the signed form).
template <int FMT, int RND, int MAG = 0>
__device__ __forceinline__ uint32_t round_pair_e4m3_like(uint32_t h, uint32_t r2) {
  constexpr bool E4 = (FMT == FPEMU_FMT_E4M3);
  constexpr uint32_t kSat = E4 ? 0x5F005F00u : 0x4F804F80u;
  constexpr uint32_t kKeep = E4 ? 0xFF80FF80u : 0xFFC0FFC0u;
  constexpr uint32_t kGridMul = E4 ? 0x20002000u : 0x30003000u;   // 2^(G - 16): 2^-7 / 2^-3
  constexpr uint32_t kGridNeg = kGridMul | 0x80008000u;
  constexpr uint32_t kKeepBias = E4 ? 0x48804880u : 0x4C404C40u;  // 9.0 / 17.0
  // saturate; NaN and Inf lanes become the maximum
  const uint32_t c = MAG ? as_u32(__hmin2(__habs2(as_h2(h)), as_h2(kSat))) : clamp_h2(h, kSat);
  // p = 2^(G - e) per lane, never materialised: the inverted exponent field IS the fp16 number x = 2^(16 - E) (one
  // LOP3; +Inf for a zero / subnormal lane, which makes p infinite and the lane flushed, as it must be), and the
  // factor 2^(G - 16) rides in the two HFMA2 that consume p
  const __half2 x = as_h2(~c & 0x7C007C00u);
  const __half2 keepf = __hfma2_sat(x, as_h2(kGridNeg), as_h2(kKeepBias));      // sat(B - p): 1.0 keep, 0.0 flush
  uint32_t r;
  if (RND == FPEMU_RND_RTZ) {
    r = c & kKeep;                                                // clearing <= 4 low bits first changes nothing
  } else {
    const __half2 m = __hfma2_relu(x, as_h2(kGridMul), as_h2(0xBC00BC00u));     // relu(p - 1) = 2^d - 1 (0 when d <= 0)
    const uint32_t q = as_u32(__hadd2(m, as_h2(0x64006400u)));    // 1024 + (2^d - 1): the mask sits in the mantissa
    const uint32_t ac = c & ~(q ^ 0x64006400u);                   // clear the low d bits
    if (RND == FPEMU_RND_RNE) {
      // round to nearest even at the kept mantissa position by letting the FMA pipe do it: with M = +-2^(e + 10 - kept
      // bits) (the lane's own power of two, same sign), ac + M has exactly the wanted ulp and the hardware rounds the
      // sum to nearest even; subtracting M again is exact.  x - x is +0 whatever the signs, so the sign of a zero lane
      // is taken from the sum.  A zero / subnormal lane has M = +-0 and passes through (it is flushed below).
      const __half2 pw = as_h2(c & 0xFC00FC00u);                                 // +-2^e
      const __half2 t = __hfma2(pw, as_h2(E4 ? 0x58005800u : 0x54005400u), as_h2(ac));          // ac + M
      const uint32_t d = as_u32(__hfma2(pw, as_h2(E4 ? 0xD800D800u : 0xD400D400u), t));         // (ac + M) - M
      r = MAG == 1 ? d : MAG == 2 ? (d | (h & kSign2)) : ((d & kAbs2) | (as_u32(t) & kSign2));
    } else {
      r = (ac + r2) & kKeep;                                      // a saturated lane has zero low bits: unchanged
    }
  }
  return as_u32(__hmul2(as_h2(r), keepf));
}

// E4M3_IEEE (fpemu_kernels.cu:656-767): E4M3's flush and pseudo-subnormal handling, but |h| >= 256 (NaN included)
// becomes +-Inf instead of saturating, and the rounding increment is only applied below 15.0 (`can_round`): larger
// magnitudes are truncated.  Same p = 2^(G-e) machinery as round_pair_e4m3_like, no clamp; lanes at or above 256
// compute garbage that the final select replaces.
template <int RND>
__device__ __forceinline__ uint32_t round_pair_e4m3_ieee(uint32_t h, uint32_t r2) {
  const __half2 av = __habs2(as_h2(h));                     // |h| as an operand modifier of its consumers
  const uint32_t big = ~__hlt2_mask(av, as_h2(0x5C005C00u));                    // >= 256, or NaN
  const uint32_t rounds = __hlt2_mask(av, as_h2(0x4B804B80u));                   // < 15.0
  const __half2 x = as_h2(~h & 0x7C007C00u);                                    // 2^(16 - E); p = 2^(9 - e) = x * 2^-7
  const __half2 keepf = __hfma2_sat(x, as_h2(0xA000A000u), as_h2(0x48804880u)); // 1.0 keep, 0.0 flush (|h| < 2^-9)
  uint32_t r;
  if (RND == FPEMU_RND_RTZ) {
    r = h & 0xFF80FF80u;
  } else {
    const __half2 m = __hfma2_relu(x, as_h2(0x20002000u), as_h2(0xBC00BC00u));  // 2^d - 1
    const uint32_t q = as_u32(__hadd2(m, as_h2(0x64006400u)));
    const uint32_t ac = h & ~(q ^ 0x64006400u);                                 // clear the low d bits
    uint32_t inc;
    if (RND == FPEMU_RND_RNE) inc = 0x003F003Fu + (__umulhi(ac, 1u << 25) & 0x00010001u);
    else inc = r2;
    r = (ac + (inc & rounds)) & 0xFF80FF80u;
  }
  r = as_u32(__hmul2(as_h2(r), keepf));
  return (r & ~big) | (((h & kSign2) | 0x7C007C00u) & big);
}

__device__ __forceinline__ uint32_t bf16x2_nan_lanes(uint32_t r) {
  const __nv_bfloat162 v = *reinterpret_cast<__nv_bfloat162*>(&r);
  return __hneu2_mask(v, v);
}

// BFLOAT16 target, two fp32 inputs -> packed bf16x2 encoding (reference BFLOAT16_Kernel, :266-304).
// FMUL.FTZ by 1.0 flushes fp32 denormals (the reference zeroes every input whose exponent field is 0),
// F2FP.BF16 rounds to nearest even / PRMT truncates, and lanes that came out as -0 become +0 (`uf.u = 0`).
// NaN inputs are NOT handled here (the multiply canonicalises them, the reference keeps the payload's high
// half): NaN lanes are reported through `nan_lanes` and the caller re-does such vectors element by element.
// `r2` holds the two 16-bit random fields for stochastic rounding.
template <int RND>
__device__ __forceinline__ uint32_t round_pair_bf16(float x0, float x1, uint32_t r2, uint32_t& nan_lanes) {
  float y0, y1;
  asm("mul.ftz.f32 %0, %1, 0f3F800000;" : "=f"(y0) : "f"(x0));
  asm("mul.ftz.f32 %0, %1, 0f3F800000;" : "=f"(y1) : "f"(x1));
  uint32_t r;
  if (RND == FPEMU_RND_RNE) {
    r = pack_bf16x2_rn(y0, y1);
    nan_lanes |= bf16x2_nan_lanes(r);
  } else {
    const uint32_t u0 = __float_as_uint(y0), u1 = __float_as_uint(y1);
    r = __byte_perm(u0, u1, 0x7632);                                            // high halves of both words
    nan_lanes |= bf16x2_nan_lanes(r);                                           // before the random add can wrap a NaN
    if (RND == FPEMU_RND_STOCHASTIC)                                            // Inf + bits stays Inf after the cut
      r = __byte_perm(u0 + (r2 & 0xFFFFu), u1 + (r2 >> 16), 0x7632);
  }
  const uint32_t zero = __heq2_mask(*reinterpret_cast<__nv_bfloat162*>(&r), __float2bfloat162_rn(0.0f));
  return r & ~zero;
}

template <int FMT, int RND, bool DAZ>
__device__ __forceinline__ uint32_t round_pair(uint32_t h, uint32_t r2) {
  if (FMT == FPEMU_FMT_E5M2) return round_pair_e5m2<RND, DAZ>(h, r2);
  if (FMT == FPEMU_FMT_E4M3_IEEE) return round_pair_e4m3_ieee<RND>(h, r2);
  return round_pair_e4m3_like<FMT, RND, RND == FPEMU_RND_RNE ? 2 : 0>(h, r2);
}

// ------------------------------------------------------------------------------------------ 8-bit codes
// Bit layouts are the reference's format table (README.md:17-27): E5M2 = S.EEEEE.MM bias 15 (the top byte of the
// fp16 encoding), E4M3 = S.EEEE.MMM bias 7, E3M4 = S.EEE.MMMM bias 3, gradual underflow, S.1111.111 / S.111.1111
// = NaN.  The reference itself never stores these bytes; it only produces the rounded fp16 value q.
//
// E4M3 / E3M4: the reference keeps full mantissa precision below the smallest normal (its "pseudo-subnormals",
// fpemu_kernels.cu:594-606, 924-936), so q can lie between two points of the format's subnormal grid
// (multiples of 2^-9 / 2^-6).  The code is the nearest grid point, ties to the even code.  In fp16 arithmetic:
// |q| + M with M = 2.0 / 16.0 has an ulp equal to the subnormal spacing (and is exact for every q at or above
// the smallest normal), then (t - M) * 2^-(15 - bias) is one exact HFMA2 whose fp16 ENCODING, because fp16's own
// denormals are gradual too, is the code's magnitude shifted left by 7 / 6.
// Returns the two codes in the low byte of each 16-bit lane; the high bytes are unspecified.
template <int FMT>
__device__ __forceinline__ uint32_t encode_pair(uint32_t q) {
  if (FMT == FPEMU_FMT_E5M2) return q >> 8;
  constexpr bool E4 = (FMT == FPEMU_FMT_E4M3);
  const __half2 t = __hadd2(__habs2(as_h2(q)), as_h2(E4 ? 0x40004000u : 0x4C004C00u));         // |q| + 2.0 / 16.0
  const uint32_t r = as_u32(__hfma2(t, as_h2(E4 ? 0x1C001C00u : 0x0C000C00u),                   // * 2^-8 / 2^-12
                                    as_h2(E4 ? 0xA000A000u : 0x9C009C00u)));                     // - 2^-7 / 2^-8
  const uint32_t mag = __umulhi(r, 1u << (E4 ? 25 : 26));                                        // r >> 7 / 6
  const uint32_t sgn = __umulhi(q, 1u << 24);                                                    // sign -> bit 7
  uint32_t code;                                                                                 // bits 0-6 of mag, rest of sgn
  asm("lop3.b32 %0, %1, %2, 0x007F007F, 0xE4;" : "=r"(code) : "r"(mag), "r"(sgn));
  return code;
}

// The same from a rounded MAGNITUDE pair `qm` (round_pair_e4m3_like<..., MAG = true>) and the pair `h` it came from, whose
// sign bits are the codes'.
template <int FMT>
__device__ __forceinline__ uint32_t encode_pair_mag(uint32_t qm, uint32_t h) {
  constexpr bool E4 = (FMT == FPEMU_FMT_E4M3);
  const __half2 t = __hadd2(as_h2(qm), as_h2(E4 ? 0x40004000u : 0x4C004C00u));
  const uint32_t r = as_u32(__hfma2(t, as_h2(E4 ? 0x1C001C00u : 0x0C000C00u), as_h2(E4 ? 0xA000A000u : 0x9C009C00u)));
  const uint32_t mag = __umulhi(r, 1u << (E4 ? 25 : 26));
  const uint32_t sgn = __umulhi(h, 1u << 24);
  uint32_t code;
  asm("lop3.b32 %0, %1, %2, 0x007F007F, 0xE4;" : "=r"(code) : "r"(mag), "r"(sgn));
  return code;
}

// Two codes (low byte of each 16-bit lane, high bytes zero) -> the fp16x2 value they denote.
template <int FMT>
__device__ __forceinline__ uint32_t decode_pair(uint32_t c2) {
  if (FMT == FPEMU_FMT_E5M2) return c2 << 8;
  constexpr bool E4 = (FMT == FPEMU_FMT_E4M3);
  const uint32_t bits = (c2 << (E4 ? 7 : 6)) & (E4 ? 0x3F803F80u : 0x1FC01FC0u);
  const uint32_t nan = __heq2_mask(as_h2(bits), as_h2(E4 ? 0x3F803F80u : 0x1FC01FC0u));         // S.1111.111
  const uint32_t mag = as_u32(__hmul2(as_h2(bits), as_h2(E4 ? 0x5C005C00u : 0x6C006C00u)));     // * 2^8 / 2^12, exact
  return mag | (nan & kAbs2) | ((c2 << 8) & kSign2);
}

// Element-wise forms of the same two maps (generic kernels, ragged tails).
__device__ __forceinline__ uint32_t encode_one(uint32_t q, int fmt) {
  if (fmt == FPEMU_FMT_E5M2) return (q >> 8) & 0xFFu;
  const bool e4 = (fmt == FPEMU_FMT_E4M3);
  const __half a = __ushort_as_half((unsigned short)(q & 0x7FFFu));
  const __half t = __hadd(a, __ushort_as_half(e4 ? 0x4000 : 0x4C00));
  const __half r = __hfma(t, __ushort_as_half(e4 ? 0x1C00 : 0x0C00), __ushort_as_half(e4 ? 0xA000 : 0x9C00));
  return (((uint32_t)__half_as_ushort(r) >> (e4 ? 7 : 6)) & 0x7Fu) | ((q >> 8) & 0x80u);
}
__device__ __forceinline__ uint32_t decode_one(uint32_t c, int fmt) {
  if (fmt == FPEMU_FMT_E5M2) return (c & 0xFFu) << 8;
  const bool e4 = (fmt == FPEMU_FMT_E4M3);
  const uint32_t m = c & 0x7Fu, sign = (c & 0x80u) << 8;
  if (m == 0x7Fu) return sign | 0x7FFFu;
  const __half v = __hmul(__ushort_as_half((unsigned short)(m << (e4 ? 7 : 6))), __ushort_as_half(e4 ? 0x5C00 : 0x6C00));
  return sign | __half_as_ushort(v);
}

// ------------------------------------------------------------------------------------------ one element
// Rounding increment decision shared by every format: `v` is the integer whose low L bits get dropped.
__device__ __forceinline__ uint32_t round_up_bit(uint32_t v, int L, int rnd, bool pos, bool neg) {
  const uint32_t grs = v & ((1u << L) - 1u);
  const uint32_t half = 1u << (L - 1);
  switch (rnd) {
    case FPEMU_RND_RNE:   return (grs > half) || (grs == half && ((v >> L) & 1u));
    case FPEMU_RND_RNAZ:  return grs >= half;
    case FPEMU_RND_RNTZ:  return grs > half;
    case FPEMU_RND_RPINF: return pos && grs >= half;
    case FPEMU_RND_RNINF: return neg && grs >= half;
    default:              return 0u;
  }
}

// fp16 encoding -> rounded fp16 encoding, formats E5M2 / E4M3 / E3M4 / E4M3_IEEE.
__device__ __forceinline__ uint32_t round_one_h16(uint32_t h, int fmt, int rnd, bool daz, uint32_t rbits) {
  const uint32_t a = h & 0x7FFFu, sign = h & 0x8000u;
  const bool pos = (sign == 0u) && a != 0u && a <= 0x7C00u;   // `h.f > 0` (false for NaN)
  const bool neg = (sign != 0u) && a != 0u && a <= 0x7C00u;
  if (fmt == FPEMU_FMT_E5M2) {
    if (a >= 0x7C00u) return h & 0xFF00u;
    if (a < 0x0400u) {
      if (daz) return 0u;
      const int r = (rnd == FPEMU_RND_STOCHASTIC) ? FPEMU_RND_RNE : rnd;
      return (h + (round_up_bit(h, 8, r, pos, neg) << 8)) & 0xFF00u;
    }
    const uint32_t inc = (rnd == FPEMU_RND_STOCHASTIC) ? (rbits & 0xFFu) : (round_up_bit(h, 8, rnd, pos, neg) << 8);
    return (h + inc) & 0xFF00u;
  }
  int L, grid_e;
  uint32_t sat, sat_out, flush, round_below;
  if (fmt == FPEMU_FMT_E3M4) { L = 6; sat = 0x4F80u; sat_out = 0x4F80u; flush = 0x2400u; grid_e = 13; round_below = sat; }
  else if (fmt == FPEMU_FMT_E4M3) { L = 7; sat = 0x5F00u; sat_out = 0x5F00u; flush = 0x1800u; grid_e = 9; round_below = sat; }
  else { L = 7; sat = 0x5C00u; sat_out = 0x7C00u; flush = 0x1800u; grid_e = 9; round_below = 0x4B80u; }  // E4M3_IEEE
  if (a >= sat) return sign | sat_out;
  if (a < flush) return sign;
  const int d = grid_e - (int)(a >> 10);
  const uint32_t ac = (d > 0) ? (a & ~((1u << d) - 1u)) : a;
  uint32_t inc = 0u;
  if (a < round_below)
    inc = (rnd == FPEMU_RND_STOCHASTIC) ? (rbits & ((1u << L) - 1u)) : (round_up_bit(ac, L, rnd, pos, neg) << L);
  return sign | ((ac + inc) & ~((1u << L) - 1u));
}

// BFLOAT16 on fp32 bits (fpemu_kernels.cu:266-304).
__device__ __forceinline__ uint32_t round_one_bf16(uint32_t u, int rnd, uint32_t rbits) {
  const uint32_t e = u & 0x7F800000u;
  if (e == 0u) return 0u;                                       // zeros and denormals -> +0
  if (e != 0x7F800000u) {
    if (rnd == FPEMU_RND_STOCHASTIC) u += rbits & 0xFFFFu;
    else if (rnd == FPEMU_RND_RNE || rnd == FPEMU_RND_RNAZ || rnd == FPEMU_RND_RNTZ)
      u += round_up_bit(u, 16, rnd, false, false) << 16;
  }
  return u & 0xFFFF0000u;
}

// FLOAT16 (fpemu_kernels.cu:318-372): returns the fp16 encoding.
__device__ __forceinline__ uint32_t round_one_fp16(float x, int rnd, bool daz, bool input_is_half, uint32_t rbits) {
  if (input_is_half || rnd != FPEMU_RND_STOCHASTIC) {
    uint32_t h = __half_as_ushort(__float2half_rn(x));
    if (daz && (h & 0x7C00u) == 0u) h = 0u;
    return h;
  }
  const uint32_t u = __float_as_uint(x);
  const int e = (int)((u >> 23) & 0xFFu) - 127;
  const uint32_t sign16 = (u >> 16) & 0x8000u;
  uint32_t m = u & 0x7FFFFFu;
  if (e == 128) {                                               // Inf / NaN (signalling NaNs get quieted)
    if (m != 0u) m |= 0x400000u;
    return sign16 | 0x7C00u | (m >> 13);
  }
  if (e >= 16) return sign16 | 0x7C00u;
  if (e < -14) return __half_as_ushort(__float2half_rn(x));     // fp16 subnormal range: plain RNE
  m = (m | ((uint32_t)(e + 15) << 23)) + (rbits & 0x1FFFu);     // 13 random bits, carry may bump the exponent
  return (sign16 | (m >> 13)) & 0xFFFFu;
}

// FP4 "nearest" on an fp32 value (fpemu_kernels.cu:1002-1023): powers of 4 between 4^-6 and 1, everything above
// 1 becomes 1, everything below 2^-12 becomes +0.  The reference bumps odd negative exponents with `inval * 1.6`
// in double precision and masks the mantissa away; only the exponent of that product survives, and it is e + 1
// exactly when the mantissa is >= 1.25 (1.25 * 1.6 = 2; rounding cannot reach 2.0 from 1.25 - 2^-23), so the whole
// function is integer arithmetic on the exponent field f = e + 127 ("e < 0 and odd" <=> f < 127 and f even).
__device__ __forceinline__ float round_one_fp4(float v) {
  const uint32_t u = __float_as_uint(v);
  const uint32_t sign = u & 0x80000000u;
  uint32_t f = (u >> 23) & 0xFFu;
  if (f > 127u) f = 127u;                                                        // e > 0 (Inf / NaN included): 1.0
  else if (f < 127u && !(f & 1u) && (u & 0x7FFFFFu) >= 0x200000u) f += 1u;       // 1.6x bump of an odd exponent
  if (f < 115u) return 0.0f;                                                     // e < -12 (tested before the step below)
  if (f < 127u && !(f & 1u)) f -= 1u;                                            // still odd: down to the even exponent
  return __uint_as_float(sign | (f << 23));
}

// FP4, two values at a time.  round_one_fp4() looks only at the sign, the exponent and the top two mantissa bits of
// the fp32 value, all of which sit in the upper 16 bits — the bfloat16 encoding — so two values are rounded as one
// 32-bit word of two bf16 encodings, and the result (a power of 4, or 0) is again a bf16 encoding, exactly.
// With n = |b| >> 5 (exponent field f and two mantissa bits) the rule "an even f goes up when the mantissa is >= 1.25,
// down otherwise; an odd f stays" is f' = 2 * ((n + 7) >> 3) - 1, i.e. (f' << 7) = ((|b| - 0x20) & 0xFF00) | 0x80.
//   clamp to [-1, 1]    |v| >= 1, Inf -> 1; NaN -> +1 (the reference multiplies by the scale first, which makes every
//                       NaN the positive canonical one)
//   bit 14              is 0 after the clamp: set, it keeps the per-lane subtraction from borrowing across lanes
//   keep                |v| >= 1.25 * 2^-13 (f' >= 115), everything smaller is +0
__device__ __forceinline__ uint32_t round_pair_fp4(uint32_t b) {
  uint32_t c;
  asm("{.reg .b32 t, n; min.bf16x2 t, %1, %2; neg.bf16x2 n, %2; max.bf16x2 %0, t, n;}" : "=r"(c) : "r"(b), "r"(0x3F803F80u));
  const uint32_t aabs = c & kAbs2;
  const uint32_t thr = 0x39203920u;
  const uint32_t keep = __hge2_mask(*reinterpret_cast<const __nv_bfloat162*>(&aabs), *reinterpret_cast<const __nv_bfloat162*>(&thr));
  const uint32_t t = (c | 0x40004000u) - 0x00200020u;
  return ((t & 0xBF00BF00u) | 0x00800080u) & keep;
}

// Full QDQ of one fp32 value (already widened from the tensor's dtype).  `scale`/`rscale` are what the
// kernel uses (for fp16 tensors the scale has been rounded to half, see cast kernels).
__device__ __forceinline__ float cast_one(float x, int fmt, int rnd, int flags, bool scaled, float scale,
                                          float rscale, bool input_is_half, uint32_t rbits) {
  const bool daz = (flags & FPEMU_FLAG_DAZ) != 0;
  if (fmt == FPEMU_FMT_BF16) return __uint_as_float(round_one_bf16(__float_as_uint(x), rnd, rbits));
  if (fmt == FPEMU_FMT_FP16)
    return __half2float(__ushort_as_half((unsigned short)round_one_fp16(x, rnd, daz, input_is_half, rbits)));
  // FP4 always multiplies (`scale * in[gid]`, fpemu_kernels.cu:1005), which also canonicalises NaN inputs
  if (fmt == FPEMU_FMT_FP4) return round_one_fp4((scaled ? scale : 1.0f) * x) * (scaled ? rscale : 1.0f);
  const uint32_t h = __half_as_ushort(__float2half_rn(scaled ? x * scale : x));
  const uint32_t q = round_one_h16(h, fmt, rnd, daz, rbits);
  float f = __half2float(__ushort_as_half((unsigned short)q));
  if (scaled) {
    f = f * rscale;
    if (!(flags & FPEMU_FLAG_UNSCALE_FP32)) f = __half2float(__float2half_rn(f));
  }
  return f;
}

}  // namespace fpemu
