/*
 * fpemu_oracle.c — CPU restatement of the FP8-Emulation-Toolkit fpemu cast.
 *
 * TEST INFRASTRUCTURE ONLY.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this file's library; the product (fp8-emulation-toolkit_b200/) never
 * does and has no CPU fallback.
 *
 * Every function restates, in plain scalar C with software fp16 conversion, the algorithm of the
 * reference at the file:line it cites (paths relative to /root/reference/mpemu/pytquant/).
 *
 * PARITY PIN: the restatement is checked bit-for-bit against outputs of the reference's own CPU
 * implementation (scalar build of cpp/fpemu_impl.cpp, see oracle/build_ref.py) on the exhaustive set
 * of 65 536 fp16 patterns, the reference's 64-value fixture (test.py:120-135; tests/golden/fixture64.npz) and
 * wide-range random fp32 inputs — fixtures in tests/golden/, generator tests/golden/make_golden.py — and, on the GPU
 * box, against the reference CUDA kernels compiled unmodified for sm_100a (oracle/_ref/fpemu_cuda_ref.so).
 * The reference's own tests hold no golden vectors (test.py only prints).
 *
 * Two "un-scale" semantics exist in the reference and both are restated:
 *   unscale_fp32 = 1 : out = float(h) * (1/scale) in fp32            (CPU:  cpp/fpemu_impl.cpp:761,1131,1716)
 *   unscale_fp32 = 0 : out = float(half_rn(float(h) * (1/scale)))    (CUDA: cuda/fpemu_kernels.cu:461,539,652,982;
 *                      the float product is passed to an at::Half parameter of __half2anyfloat)
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "../include/fpemu_b200.h"

typedef union { uint32_t u; float f; } f32_t;

/* ---------------------------------------------------------------- software fp16 <-> fp32 (IEEE RNE) */

/* Same result as x86 `_cvtss_sh(x, _MM_FROUND_TO_NEAREST_INT)` (cpp/fpemu_impl.cpp:117-119) and CUDA
 * `__float2half_rn` (cuda/fpemu_kernels.cu:128) for every non-NaN input.  NaN -> 0x7FFF | sign-less
 * canonical pattern as `cvt.rn.f16.f32` produces; NaN payloads are compared as "is NaN" by the tests. */
static int g_nan_mode = 0;   /* 0: CUDA cvt.rn.f16.f32 (NaN -> 0x7FFF); 1: x86 F16C (sign + top payload kept, quieted) */

/* Tests pin against the reference CPU build with mode 1 and against the reference CUDA build with mode 0. */
void fpemu_oracle_set_nan_mode(int x86_f16c) { g_nan_mode = x86_f16c ? 1 : 0; }

uint16_t fpemu_oracle_float2half(float x) {
  f32_t v; v.f = x;
  uint32_t sign = (v.u >> 16) & 0x8000u;
  uint32_t a = v.u & 0x7FFFFFFFu;
  if (a > 0x7F800000u)                                      /* NaN                         */
    return g_nan_mode ? (uint16_t)(sign | 0x7E00u | ((a & 0x7FFFFFu) >> 13)) : (uint16_t)0x7FFF;
  if (a >= 0x477FF000u) return (uint16_t)(sign | 0x7C00u);  /* >= 65520 rounds to Inf      */
  if (a < 0x33000001u) return (uint16_t)sign;               /* <= 2^-25 rounds to zero     */
  int e = (int)(a >> 23) - 127;
  uint32_t m = (a & 0x7FFFFFu) | 0x800000u;                 /* 24-bit significand          */
  uint32_t h;
  int shift;
  if (e < -14) {                                            /* fp16 subnormal              */
    shift = 13 + (-14 - e);                                 /* 14..24                      */
    h = 0;
  } else {
    shift = 13;
    h = (uint32_t)(e + 15) << 10;
    m &= 0x7FFFFFu;
  }
  uint32_t q = m >> shift;
  uint32_t rem = m & ((1u << shift) - 1u);
  uint32_t half = 1u << (shift - 1);
  h += q;
  if (rem > half || (rem == half && (q & 1u))) h += 1;      /* carry may bump the exponent */
  return (uint16_t)(sign | h);
}

float fpemu_oracle_half2float(uint16_t h) {
  f32_t v;
  uint32_t sign = ((uint32_t)h & 0x8000u) << 16;
  uint32_t e = (h >> 10) & 0x1Fu;
  uint32_t m = h & 0x3FFu;
  if (e == 0x1F) {
    v.u = sign | 0x7F800000u | (m << 13);
  } else if (e == 0) {
    if (m == 0) {
      v.u = sign;
    } else {
      int s = 0;
      while (!(m & 0x400u)) { m <<= 1; ++s; }
      m &= 0x3FFu;
      v.u = sign | ((uint32_t)(127 - 15 - s + 1) << 23) | (m << 13);
    }
  } else {
    v.u = sign | ((e + 112u) << 23) | (m << 13);
  }
  return v.f;
}

static float bf16_to_float(uint16_t b) { f32_t v; v.u = (uint32_t)b << 16; return v.f; }

/* torch's float -> BFloat16 (round to nearest even, NaN -> 0x7FC0). */
static uint16_t float_to_bf16_rn(float x) {
  f32_t v; v.f = x;
  if ((v.u & 0x7FFFFFFFu) > 0x7F800000u) return 0x7FC0;
  uint32_t r = v.u + 0x7FFFu + ((v.u >> 16) & 1u);
  return (uint16_t)(r >> 16);
}

/* ---------------------------------------------------------------- Philox4x32-10 (Salmon et al. 2011) */

static void philox_round(uint32_t c[4], const uint32_t k[2]) {
  uint64_t p0 = (uint64_t)0xD2511F53u * c[0];
  uint64_t p1 = (uint64_t)0xCD9E8D57u * c[2];
  uint32_t n0 = (uint32_t)(p1 >> 32) ^ c[1] ^ k[0];
  uint32_t n1 = (uint32_t)p1;
  uint32_t n2 = (uint32_t)(p0 >> 32) ^ c[3] ^ k[1];
  uint32_t n3 = (uint32_t)p0;
  c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
}

/* Key = (seed lo, seed hi); counter = (ctr lo, ctr hi, 0, 0). */
void fpemu_oracle_philox4x32_10(uint64_t seed, uint64_t ctr, uint32_t out[4]) {
  uint32_t k[2] = { (uint32_t)seed, (uint32_t)(seed >> 32) };
  uint32_t c[4] = { (uint32_t)ctr, (uint32_t)(ctr >> 32), 0u, 0u };
  for (int r = 0; r < 10; ++r) {
    philox_round(c, k);
    k[0] += 0x9E3779B9u; k[1] += 0xBB67AE85u;
  }
  memcpy(out, c, sizeof(c));
}

/* The product's documented mapping element index -> random bits (DESIGN.md "Stochastic rounding"):
 *   8-bit formats (E5M2, E4M3, E4M3_IEEE, E3M4): element g uses byte (g % 16) of Philox(seed, g / 16)
 *   16-bit-field formats (BF16: 16 bits, FP16: 13 bits): element g uses 16-bit lane (g % 8) of Philox(seed, g / 8)
 * (little-endian lanes of the four output words).  out[i] holds the bits for element elem_offset + i. */
void fpemu_oracle_rand_bits(uint64_t seed, uint64_t elem_offset, int64_t n, int format, uint32_t* out) {
  int wide = (format == FPEMU_FMT_BF16 || format == FPEMU_FMT_FP16);
  uint64_t cur = ~(uint64_t)0;
  uint32_t w[4] = {0, 0, 0, 0};
  for (int64_t i = 0; i < n; ++i) {
    uint64_t g = elem_offset + (uint64_t)i;
    uint64_t ctr = wide ? (g >> 3) : (g >> 4);
    if (ctr != cur) { fpemu_oracle_philox4x32_10(seed, ctr, w); cur = ctr; }
    if (wide) {
      unsigned lane = (unsigned)(g & 7u);
      out[i] = (w[lane >> 1] >> ((lane & 1u) * 16u)) & 0xFFFFu;
    } else {
      unsigned lane = (unsigned)(g & 15u);
      out[i] = (w[lane >> 2] >> ((lane & 3u) * 8u)) & 0xFFu;
    }
  }
}

/* ---------------------------------------------------------------- per-format rounding on fp16 bits */

/* The six deterministic rounding increments share one shape in every reference kernel
 * (cuda/fpemu_kernels.cu:446-456, 634-644, 964-974; cpp/fpemu_impl.cpp:744-754, 1112-1120):
 * `field` is the value whose low `lshift` bits get dropped, `half` = 1 << (lshift-1),
 * tie mask = 3 << (lshift-1).  Returns 0 or 1 (the caller shifts it by lshift). */
static unsigned round_inc(unsigned field, int lshift, int rmode, int positive, int negative) {
  unsigned grs = field & ((1u << lshift) - 1u);
  unsigned half = 1u << (lshift - 1);
  unsigned tie_mask = 3u << (lshift - 1);
  switch (rmode) {
    case FPEMU_RND_RNE:   return (grs > half) || ((field & tie_mask) == tie_mask);
    case FPEMU_RND_RNAZ:  return grs >= half;
    case FPEMU_RND_RNTZ:  return grs > half;
    case FPEMU_RND_RPINF: return positive && (grs >= half);   /* `(h.f > 0) * (rnmask >= half)` */
    case FPEMU_RND_RNINF: return negative && (grs >= half);   /* `(h.f < 0) * (rnmask >= half)` */
    default:              return 0;                            /* RTZ: every mask is 0          */
  }
}

/* E5M2_Kernel, cuda/fpemu_kernels.cu:417-462 == cvt_fp32_e5m2_scalar, cpp/fpemu_impl.cpp:716-762.
 * daz = 1: E5M2_DAZ_Kernel, cuda :494-540 == cpp :874-918. */
static uint16_t round_e5m2(uint16_t hu, int rmode, int daz, uint32_t rnd) {
  const int lshift = 8;                                        /* 10 - (8 - (5 + 1))            */
  unsigned can_round = ((hu & 0x7F00u) <= 0x7B00u);
  unsigned is_normal = ((hu & 0x7C00u) <= 0x7800u) && ((hu & 0x7C00u) >= 0x0400u);
  unsigned is_denorm = ((hu & 0x7C00u) == 0);
  unsigned is_naninf = ((hu & 0x7C00u) == 0x7C00u);
  int positive = !(hu & 0x8000u) && (hu & 0x7FFFu) && !(is_naninf && (hu & 0x3FFu));
  int negative = (hu & 0x8000u) && (hu & 0x7FFFu) && !(is_naninf && (hu & 0x3FFu));
  unsigned h = hu;
  if (!daz) {
    if (!is_naninf) {
      if (rmode == FPEMU_RND_STOCHASTIC) {
        h += can_round * is_normal * (rnd & 0xFFu);                                   /* :441 */
        h += can_round * is_denorm * (round_inc(hu, lshift, FPEMU_RND_RNE, 0, 0) << lshift); /* :443 */
      } else {
        h += can_round * (round_inc(hu, lshift, rmode, positive, negative) << lshift);
      }
    }
  } else {
    if (!is_naninf && is_normal) {                                                    /* :511 */
      if (rmode == FPEMU_RND_STOCHASTIC) h += can_round * (rnd & 0xFFu);              /* :518 */
      else h += can_round * (round_inc(hu, lshift, rmode, positive, negative) << lshift);
    } else if (is_denorm) {
      h = 0;                                                                          /* :534 sign lost */
    }
  }
  return (uint16_t)(h & 0xFF00u);                                                     /* :460 */
}

/* Shared body of E4M3_Kernel (cuda :585-653, cpp :1063-1132) and E3M4_Kernel (cuda :915-983,
 * cpp :1649-1717).  sat_bits: |h| >= sat_bits saturates; sat_exp / sat_mant: the saturated value;
 * flush_exp: exp_h < flush_exp -> signed zero; sub_exp: exp_h < sub_exp -> pseudo-subnormal with
 * dshift = sub_exp - exp_h low mantissa bits cleared before rounding. */
static uint16_t round_e4m3_like(uint16_t hu, int lshift, unsigned sat_bits, int sat_exp, int sat_mant,
                                int flush_exp, int sub_exp, int rmode, uint32_t rnd) {
  short exp_h = (short)((hu & 0x7C00u) >> 10) - 15;
  short sign_h = (short)(hu & 0x8000u);
  short mantissa_h = (short)(hu & 0x03FFu);
  unsigned can_round = ((hu & 0x7FFFu) < sat_bits);
  unsigned is_normal = ((hu & 0x7C00u) <= 0x7800u) && ((hu & 0x7C00u) >= 0x0400u);
  unsigned is_denorm = ((hu & 0x7C00u) == 0);
  unsigned is_naninf = ((hu & 0x7C00u) == 0x7C00u);
  int positive = !(hu & 0x8000u) && (hu & 0x7FFFu) && !(is_naninf && (hu & 0x3FFu));
  int negative = (hu & 0x8000u) && (hu & 0x7FFFu) && !(is_naninf && (hu & 0x3FFu));
  int dshift = 0;
  if (exp_h > sat_exp || can_round == 0) {          /* saturate, Inf and NaN included */
    mantissa_h = (short)sat_mant;
    exp_h = (short)sat_exp;
    can_round = 0;
  } else if (exp_h < flush_exp) {                   /* flush to signed zero           */
    exp_h = -15;
    mantissa_h = 0;
  } else if (exp_h < sub_exp) {                     /* pseudo-subnormal               */
    dshift = sub_exp - exp_h;
    mantissa_h = (short)(mantissa_h >> dshift);
    mantissa_h = (short)(mantissa_h << dshift);
  }
  if (!is_naninf) {
    unsigned m = (unsigned short)mantissa_h;
    if (rmode == FPEMU_RND_STOCHASTIC) {
      mantissa_h = (short)(mantissa_h + can_round * is_normal * (rnd & ((1u << lshift) - 1u)));
      mantissa_h = (short)(mantissa_h + can_round * is_denorm *
                           (round_inc(m, lshift, FPEMU_RND_RNE, 0, 0) << lshift));
    } else {
      mantissa_h = (short)(mantissa_h + can_round * (round_inc(m, lshift, rmode, positive, negative) << lshift));
    }
  }
  mantissa_h = (short)(mantissa_h & (short)(0xFFFFu << lshift));
  mantissa_h = (short)(mantissa_h + ((exp_h + 15) << 10));
  mantissa_h = (short)(mantissa_h | sign_h);
  return (uint16_t)mantissa_h;
}

static uint16_t round_e4m3(uint16_t hu, int rmode, uint32_t rnd) {
  /* lshift 7; |h| >= 0x5F00 (448) or exp > 8 -> 448 = exp 8, mant 0x300; flush exp < -9; subnormal exp < -6 */
  return round_e4m3_like(hu, 7, 0x5F00u, 8, 0x0300, -9, -6, rmode, rnd);
}

static uint16_t round_e3m4(uint16_t hu, int rmode, uint32_t rnd) {
  /* lshift 6; |h| >= 0x4F80 (30) or exp > 4 -> 30 = exp 4, mant 0x380; flush exp < -6; subnormal exp < -2 */
  return round_e4m3_like(hu, 6, 0x4F80u, 4, 0x0380, -6, -2, rmode, rnd);
}

/* E4M3_IEEE_Kernel, cuda/fpemu_kernels.cu:699-766 (CUDA text; the CPU scalar twin differs). */
static uint16_t round_e4m3_ieee(uint16_t hu, int rmode, uint32_t rnd) {
  const int lshift = 7;
  short exp_h = (short)((hu & 0x7C00u) >> 10) - 15;
  short sign_h = (short)(hu & 0x8000u);
  short mantissa_h = (short)(hu & 0x03FFu);
  unsigned can_round = ((hu & 0x7FFFu) < 0x4B80u);
  unsigned is_normal = ((hu & 0x7C00u) <= 0x7800u) && ((hu & 0x7C00u) >= 0x0400u);
  unsigned is_denorm = ((hu & 0x7C00u) == 0);
  unsigned is_naninf = ((hu & 0x7C00u) == 0x7C00u);
  int positive = !(hu & 0x8000u) && (hu & 0x7FFFu) && !(is_naninf && (hu & 0x3FFu));
  int negative = (hu & 0x8000u) && (hu & 0x7FFFu) && !(is_naninf && (hu & 0x3FFu));
  int dshift = 0;
  if (exp_h > 7) {
    mantissa_h = 0; exp_h = 16; is_naninf = 1;
  } else if (exp_h < -9) {
    exp_h = -15; mantissa_h = 0;
  } else if (exp_h < -6) {
    dshift = -6 - exp_h;
    mantissa_h = (short)(mantissa_h >> dshift);
    mantissa_h = (short)(mantissa_h << dshift);
  }
  if (!is_naninf) {
    unsigned m = (unsigned short)mantissa_h;
    if (rmode == FPEMU_RND_STOCHASTIC) {
      mantissa_h = (short)(mantissa_h + can_round * is_normal * (rnd & 0x7Fu));
      mantissa_h = (short)(mantissa_h + can_round * is_denorm * (round_inc(m, lshift, FPEMU_RND_RNE, 0, 0) << lshift));
    } else {
      mantissa_h = (short)(mantissa_h + can_round * (round_inc(m, lshift, rmode, positive, negative) << lshift));
    }
  }
  mantissa_h = (short)(mantissa_h & (short)(0xFFFFu << lshift));
  mantissa_h = (short)(mantissa_h + ((exp_h + 15) << 10));
  mantissa_h = (short)(mantissa_h | sign_h);
  return (uint16_t)mantissa_h;
}

/* BFLOAT16_Kernel, cuda/fpemu_kernels.cu:266-304 (the CUDA text is authoritative: the CPU scalar twin
 * mis-rounds and the AVX-512 twin keeps fp32 denormals). */
static float round_bf16(float x, int rmode, uint32_t rnd) {
  f32_t uf; uf.f = x;
  unsigned is_normal = ((uf.u & 0x7F800000u) <= 0x7F000000u) && ((uf.u & 0x7F800000u) >= 0x00800000u);
  unsigned is_denorm = ((uf.u & 0x7F800000u) == 0);
  unsigned is_naninf = ((uf.u & 0x7F800000u) == 0x7F800000u);
  if (!is_naninf && is_normal) {
    if (rmode == FPEMU_RND_STOCHASTIC) uf.u += (rnd & 0xFFFFu);
    else uf.u += round_inc(uf.u, 16, rmode, 0, 0) << 16;   /* RPINF/RNINF have no mask here: add 0 */
  } else if (is_denorm) {
    uf.u = 0;
  }
  uf.u &= 0xFFFF0000u;
  return uf.f;
}

/* FLOAT16_Kernel, cuda/fpemu_kernels.cu:318-372 == cpp/fpemu_impl.cpp:369-421.  Returns fp16 bits.
 * `from_f32_bits` is the raw fp32 input (the stochastic branch works on it directly). */
static uint16_t round_fp16(float x, int rmode, int no_denorm, int input_is_half, uint32_t rnd) {
  if (input_is_half || rmode != FPEMU_RND_STOCHASTIC) {
    uint16_t h = fpemu_oracle_float2half(x);
    unsigned is_denorm = (((h & 0x7FFFu) >> 10) & 0x1Fu) == 0;
    if (is_denorm && no_denorm) h = 0;
    return h;
  }
  f32_t v; v.f = x;
  int exp_h = (int)((v.u & 0x7F800000u) >> 23) - 127;
  uint32_t mantissa_h = v.u & 0x7FFFFFu;
  uint32_t sign_h = v.u & 0x80000000u;
  if (exp_h == 128) {
    if (mantissa_h && ((mantissa_h & 0x400000u) == 0)) mantissa_h |= 0x400000u;
    mantissa_h |= (0x1Fu << 23);
    mantissa_h |= (sign_h >> 3);
    return (uint16_t)(mantissa_h >> 13);
  } else if (exp_h >= 16) {
    mantissa_h = (0x1Fu << 23) | (sign_h >> 3);
    return (uint16_t)(mantissa_h >> 13);
  } else if (exp_h < -14) {
    return fpemu_oracle_float2half(x);
  }
  mantissa_h |= ((uint32_t)(exp_h + 15) << 23);
  mantissa_h |= (sign_h >> 3);
  mantissa_h += (rnd & 0x1FFFu);
  return (uint16_t)(mantissa_h >> 13);
}

/* FP4_Nearest_Kernel, cuda/fpemu_kernels.cu:1002-1023 == cvt_fp32_fp4_nearest_scalar cpp :1803-1824. */
static float round_fp4(float inval) {
  f32_t f; f.f = inval;
  int exp_f = (int)((f.u & 0x7F800000u) >> 23) - 127;
  uint32_t sign_f = f.u & 0x80000000u;
  if (exp_f < 0 && (exp_f % 2)) f.f = (float)((double)f.f * 1.6);
  if (exp_f > 0) f.u = sign_f | (127u << 23);
  f.u &= 0xFF800000u;
  exp_f = (int)((f.u & 0x7F800000u) >> 23) - 127;
  if (exp_f < 0 && (exp_f % 2)) f.u = sign_f | ((uint32_t)(exp_f + 126) << 23);
  if (exp_f < -12) f.u = 0;
  return f.f;
}

/* ---------------------------------------------------------------- element I/O */

static float load_elem(const void* p, int64_t i, int dtype) {
  switch (dtype) {
    case FPEMU_DT_F32:  return ((const float*)p)[i];
    case FPEMU_DT_F16: {
      /* GPU half->float conversion returns the canonical NaN 0x7FFFFFFF for every NaN input (x86 keeps
       * sign and payload); observable in the BFLOAT16 kernel, which masks instead of converting back. */
      f32_t v; v.f = fpemu_oracle_half2float(((const uint16_t*)p)[i]);
      if (!g_nan_mode && (v.u & 0x7FFFFFFFu) > 0x7F800000u) v.u = 0x7FFFFFFFu;
      return v.f;
    }
    default:            return bf16_to_float(((const uint16_t*)p)[i]);
  }
}

static void store_elem(void* p, int64_t i, int dtype, float v) {
  switch (dtype) {
    case FPEMU_DT_F32:  ((float*)p)[i] = v; break;
    case FPEMU_DT_F16:  ((uint16_t*)p)[i] = fpemu_oracle_float2half(v); break;
    default:            ((uint16_t*)p)[i] = float_to_bf16_rn(v); break;
  }
}

/* fp32 multiply as the GPU does it: a NaN result is the canonical 0x7FFFFFFF (sign and payload dropped);
 * x86 propagates the operand's NaN.  Only observable where no float->half conversion follows (FP4). */
static float mul_f32(float a, float b) {
  f32_t r; r.f = a * b;
  if (!g_nan_mode && (r.u & 0x7FFFFFFFu) > 0x7F800000u) r.u = 0x7FFFFFFFu;
  return r.f;
}

static float recip_f32(float s) { return (float)(1.0 / (double)s); }  /* `float r = 1.0 / scale;` */

/* One element through the scaled fp16-domain formats.  `scale` is what the kernel sees as a float,
 * `scale_h16` says the E5M2 kernels hold it (and multiply) in half precision — the at::Half
 * instantiation declares `scalar_t scale` (cuda/fpemu_kernels.cu:405,420,469,497) while E4M3/E3M4
 * declare `float scale = in_scale` (:573, :903). */
static float kernel_scale(float scale, int dtype);

static uint16_t quantize_h16(float x, int dtype, int format, int rmode, int flags, float scale, uint32_t rnd) {
  float inval;
  if (dtype == FPEMU_DT_F16 && format == FPEMU_FMT_E5M2)
    inval = fpemu_oracle_half2float(fpemu_oracle_float2half(x * scale));   /* Half * Half -> Half */
  else
    inval = x * scale;
  uint16_t h = fpemu_oracle_float2half(inval);
  switch (format) {
    case FPEMU_FMT_E5M2:      h = round_e5m2(h, rmode, (flags & FPEMU_FLAG_DAZ) != 0, rnd); break;
    case FPEMU_FMT_E4M3:      h = round_e4m3(h, rmode, rnd); break;
    case FPEMU_FMT_E3M4:      h = round_e3m4(h, rmode, rnd); break;
    default:                  h = round_e4m3_ieee(h, rmode, rnd); break;
  }
  return h;
}

static float unscale_h16(uint16_t h, int flags, float rscale) {
  float q = fpemu_oracle_half2float(h) * rscale;
  if (!(flags & FPEMU_FLAG_UNSCALE_FP32))
    q = fpemu_oracle_half2float(fpemu_oracle_float2half(q));
  return q;
}

static float cast_one_h16(float x, int dtype, int format, int rmode, int flags, float scale, float rscale,
                          uint32_t rnd) {
  return unscale_h16(quantize_h16(x, dtype, format, rmode, flags, scale, rnd), flags, rscale);
}

/* ---- 8-bit codes.  Bit layouts: the reference's format table, README.md:17-27 (E5M2 S.EEEEE.MM bias 15,
 * E4M3 S.EEEE.MMM bias 7, E3M4 S.EEE.MMMM bias 3; S.1111.111 / S.111.1111 = NaN; gradual underflow).  The
 * reference never stores the byte, so these two maps have no reference code to follow; they are written with
 * integer arithmetic only, independently of the CUDA kernels' fp16-FMA formulation.  A rounded value the
 * reference leaves between two subnormal grid points (its pseudo-subnormals keep full mantissa precision,
 * cuda/fpemu_kernels.cu:594-606, 924-936) goes to the nearest grid point, ties to the even code. */
static unsigned rne_shift(unsigned v, int s) {
  if (s <= 0) return v << (-s);
  if (s > 31) return 0u;
  unsigned q = v >> s, rem = v & ((1u << s) - 1u), half = 1u << (s - 1);
  if (rem > half || (rem == half && (q & 1u))) ++q;
  return q;
}

uint8_t fpemu_oracle_code_from_h16(uint16_t q, int format) {
  if (format == FPEMU_FMT_E5M2) return (uint8_t)(q >> 8);
  const int mbits = (format == FPEMU_FMT_E4M3) ? 3 : 4;          /* stored mantissa bits          */
  const int bias = (format == FPEMU_FMT_E4M3) ? 7 : 3;
  const unsigned a = q & 0x7FFFu, sign = (q >> 8) & 0x80u;
  const int e5 = (int)(a >> 10);
  const int min_normal_e5 = 15 + (1 - bias);                      /* fp16 exponent field of 2^(1-bias) */
  unsigned mag;
  if (e5 >= min_normal_e5) {
    mag = (a >> (10 - mbits)) - ((unsigned)(15 - bias) << mbits);
  } else {
    /* value = mant * 2^(e - 25); subnormal grid spacing = 2^(1 - bias - mbits) */
    const unsigned mant = e5 ? (0x400u | (a & 0x3FFu)) : (a & 0x3FFu);
    const int e = e5 ? e5 : 1;
    mag = rne_shift(mant, (1 - bias - mbits) - (e - 25));
  }
  return (uint8_t)(sign | mag);
}

uint16_t fpemu_oracle_h16_from_code(uint8_t c, int format) {
  if (format == FPEMU_FMT_E5M2) return (uint16_t)(c << 8);
  const int mbits = (format == FPEMU_FMT_E4M3) ? 3 : 4;
  const int bias = (format == FPEMU_FMT_E4M3) ? 7 : 3;
  const unsigned m = c & 0x7Fu, sign = (unsigned)(c & 0x80u) << 8;
  if (m == 0x7Fu) return (uint16_t)(sign | 0x7FFFu);
  const unsigned E = m >> mbits, M = m & ((1u << mbits) - 1u);
  if (E == 0) {                                                   /* M * 2^(1 - bias - mbits), exact in fp16 */
    f32_t v; v.f = (float)M;
    if (M) v.u += (uint32_t)(1 - bias - mbits) << 23;
    return (uint16_t)(sign | fpemu_oracle_float2half(v.f));
  }
  return (uint16_t)(sign | ((E + 15 - bias) << 10) | (M << (10 - mbits)));
}

int fpemu_oracle_encode(const void* in, uint8_t* codes, int64_t n, int dtype, int format, int rounding, int flags,
                        float scale, const uint32_t* rnd) {
  if (n < 0 || dtype < 0 || dtype >= FPEMU_DT_COUNT) return FPEMU_ERR_ARG;
  if (format != FPEMU_FMT_E5M2 && format != FPEMU_FMT_E4M3 && format != FPEMU_FMT_E3M4) return FPEMU_ERR_MODE;
  if (scale > 65504.0f) return FPEMU_ERR_SCALE;
  const float ks = kernel_scale(scale, dtype);
  for (int64_t i = 0; i < n; ++i)
    codes[i] = fpemu_oracle_code_from_h16(
        quantize_h16(load_elem(in, i, dtype), dtype, format, rounding, flags, ks, rnd ? rnd[i] : 0u), format);
  return FPEMU_OK;
}

int fpemu_oracle_decode(const uint8_t* codes, void* out, int64_t n, int dtype, int format, int flags, float scale) {
  if (n < 0 || dtype < 0 || dtype >= FPEMU_DT_COUNT) return FPEMU_ERR_ARG;
  if (format != FPEMU_FMT_E5M2 && format != FPEMU_FMT_E4M3 && format != FPEMU_FMT_E3M4) return FPEMU_ERR_MODE;
  const float rs = recip_f32(kernel_scale(scale, dtype));
  for (int64_t i = 0; i < n; ++i)
    store_elem(out, i, dtype, unscale_h16(fpemu_oracle_h16_from_code(codes[i], format), flags, rs));
  return FPEMU_OK;
}

/* The scale a kernel instantiated for `dtype` actually sees: `const scalar_t in_scale` rounds the
 * host float to half for at::Half tensors (cuda/fpemu_kernels.cu:379, 469, 547, 877). */
static float kernel_scale(float scale, int dtype) {
  if (dtype == FPEMU_DT_F16) return fpemu_oracle_half2float(fpemu_oracle_float2half(scale));
  return scale;
}

int fpemu_oracle_cast(const void* in, void* out, int64_t n, int dtype, int format, int rounding, int flags,
                      float scale, const uint32_t* rnd) {
  if (n < 0 || dtype < 0 || dtype >= FPEMU_DT_COUNT) return FPEMU_ERR_ARG;
  if (format < 0 || format >= FPEMU_FMT_COUNT) return FPEMU_ERR_MODE;
  if (scale > 65504.0f) return FPEMU_ERR_SCALE;            /* cuda/fpemu_kernels.cu:1037-1041 */
  float ks = kernel_scale(scale, dtype);
  float rs = recip_f32(ks);
  for (int64_t i = 0; i < n; ++i) {
    float x = load_elem(in, i, dtype);
    uint32_t r = rnd ? rnd[i] : 0u;
    float q;
    switch (format) {
      case FPEMU_FMT_BF16:
        q = round_bf16(x, rounding, r);
        break;
      case FPEMU_FMT_FP16:
        q = fpemu_oracle_half2float(round_fp16(x, rounding, (flags & FPEMU_FLAG_DAZ) != 0,
                                               dtype == FPEMU_DT_F16, r));
        break;
      case FPEMU_FMT_FP4: {
        float v = mul_f32(round_fp4(mul_f32(ks, x)), rs);
        q = v;
        break;
      }
      default:
        q = cast_one_h16(x, dtype, format, rounding, flags, ks, rs, r);
        break;
    }
    store_elem(out, i, dtype, q);
  }
  return FPEMU_OK;
}

/* Blocked mode: absmax_block + scale derivation, cuda/fpemu_kernels.cu:150-170, :407-414, :575-582,
 * :905-912 (CPU twins cpp/fpemu_impl.cpp:778-799, 1149-1171, 1734-1755).  The reference's smem tree
 * over-reads for non-power-of-two blocks; this restates the intended maximum. */
int fpemu_oracle_cast_blocked(const void* in, void* out, int64_t n, int64_t block_size, int dtype, int format,
                              int rounding, int flags, const uint32_t* rnd) {
  if (n < 0 || block_size <= 0 || (n % block_size) != 0) return FPEMU_ERR_ARG;
  if (format == FPEMU_FMT_BF16 || format == FPEMU_FMT_FP16) return FPEMU_ERR_MODE;
  for (int64_t b0 = 0; b0 < n; b0 += block_size) {
    float maxval = 0.0f;
    for (int64_t i = b0; i < b0 + block_size; ++i) {
      float a = fabsf(load_elem(in, i, dtype));
      maxval = (maxval < a) ? a : maxval;                    /* NaN never wins, as fmaxf / `<` */
    }
    f32_t f; f.f = maxval; f.u &= 0x7F800000u;
    float scale;
    switch (format) {
      case FPEMU_FMT_E5M2:
        if (dtype == FPEMU_DT_F16) {                         /* `scalar_t scale` is a half here */
          scale = fpemu_oracle_half2float(fpemu_oracle_float2half((float)(2.0 * (double)f.f)));
          scale = fpemu_oracle_half2float(fpemu_oracle_float2half((float)((double)scale / 16384.0)));
        } else {
          scale = (float)(2.0 * (double)f.f); scale = (float)((double)scale / 16384.0);
        }
        break;
      case FPEMU_FMT_E4M3: case FPEMU_FMT_E4M3_IEEE:
        scale = 2 * f.f; scale = (float)((double)scale / 8.0); break;
      case FPEMU_FMT_E3M4:
        scale = 2 * f.f; break;
      default: /* FP4 */
        scale = (float)(1.0 / (double)maxval); break;
    }
    float rs = recip_f32(scale);
    for (int64_t i = b0; i < b0 + block_size; ++i) {
      float x = load_elem(in, i, dtype);
      uint32_t r = rnd ? rnd[i] : 0u;
      float q = (format == FPEMU_FMT_FP4) ? mul_f32(round_fp4(mul_f32(scale, x)), rs)
                                          : cast_one_h16(x, dtype, format, rounding, flags, scale, rs, r);
      store_elem(out, i, dtype, q);
    }
  }
  return FPEMU_OK;
}

/* max |x| over n elements the way `torch.max(torch.abs(t.flatten()))` computes it (mpemu/qutils.py:283, :303, :327):
 * a NaN anywhere makes the result NaN (torch.max propagates NaN).  The blocked mode's own maximum
 * (absmax_block, cuda/fpemu_kernels.cu:150-170: fmaxf, NaN ignored) lives in fpemu_oracle_cast_blocked above. */
float fpemu_oracle_amax(const void* in, int64_t n, int dtype) {
  float m = 0.0f;
  for (int64_t i = 0; i < n; ++i) {
    const float a = fabsf(load_elem(in, i, dtype));
    if (a != a) return a;
    m = fmaxf(m, a);
  }
  return m;
}

/* The scale quantize_tensor derives from the maximum (mpemu/qutils.py:283-285; per channel :303-305):
 *     scale = flt_max / vmax;  scale = 6.55e+04 if scale > 3.275e+04 else scale
 * `vmax` is a 0-dim tensor, so the division is Tensor.__rtruediv__, which torch implements as
 * `vmax.reciprocal() * flt_max` (torch/_tensor.py, `__rdiv__`): the reciprocal is rounded to the tensor's dtype, the
 * product again.  For a half tensor both roundings and the `>` comparison happen in half (3.275e4 is 32752 there);
 * the clamp value travels as the Python float 6.55e4 and is rounded to half (65504) when the kernel takes its
 * `scalar_t scale` argument (kernel_scale()).  bfloat16 is not dispatched by the reference: float arithmetic on the
 * widened values.  `clamp` = 0 restates the fine-grained branch, which has no second line (:327-328). */
float fpemu_oracle_scale_from_amax_dtype(float amax, float flt_max, int dtype, int clamp) {
  if (dtype == FPEMU_DT_F16) {
    const float r = fpemu_oracle_half2float(fpemu_oracle_float2half((float)(1.0 / (double)amax)));
    const float s = fpemu_oracle_half2float(fpemu_oracle_float2half(mul_f32(r, flt_max)));
    return (clamp && s > 32752.0f) ? 6.55e4f : s;
  }
  const float s = mul_f32((float)(1.0 / (double)amax), flt_max);
  return (clamp && s > 3.275e4f) ? 6.55e4f : s;
}

float fpemu_oracle_scale_from_amax(float amax, float flt_max) {
  return fpemu_oracle_scale_from_amax_dtype(amax, flt_max, FPEMU_DT_F32, 1);
}
