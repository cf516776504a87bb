// fpemu_cast_core.cuh — per-vector body of the cast, shared by the plain, blocked and row-wise kernels.
#pragma once

#include "fpemu_common.cuh"

namespace fpemu {

struct CastArgs {
  const void* in;
  void* out;
  int64_t n;
  float scale;
  const float* d_scale;
  uint32_t seed_lo, seed_hi;
  uint64_t elem_offset;
  int fmt, rnd, flags;
  // optional: derive the scale from a device-resident absolute maximum inside the cast kernel
  // (scale = clamp(flt_max / *d_amax), mpemu/qutils.py:283-285); the scale used is written to d_scale_out
  const float* d_amax;
  float flt_max;
  float* d_scale_out;
  // optional: a device-resident 64-bit word added to the seed when the kernel starts (fpemu_set_seed_source), so that
  // a captured CUDA graph draws fresh random bits on every replay
  const unsigned long long* d_seed;
};

// The seed a stochastic kernel uses: the host-side seed of the call plus the device-side word, if one is installed.
__device__ __forceinline__ CastArgs with_device_seed(const CastArgs& in) {
  CastArgs a = in;
  if (a.rnd == FPEMU_RND_STOCHASTIC && a.d_seed != nullptr) {
    const unsigned long long s = (((unsigned long long)a.seed_hi << 32) | a.seed_lo) + __ldg(a.d_seed);
    a.seed_lo = (uint32_t)s;
    a.seed_hi = (uint32_t)(s >> 32);
  }
  return a;
}

// Host side: the device word installed for the current device (nullptr when none), fpemu_api.cu.
const unsigned long long* seed_source();

constexpr int kThreads = 256;
// 32-byte vectors in flight per thread (tuned on B200, profiles/tune_r1.txt): fp32 is HBM-bound with 2;
// 16-bit tensors carry twice the elements per byte and want 4.
template <int DT> struct UnrollFor { static constexpr int value = (DT == FPEMU_DT_F32) ? 2 : 4; };

struct ScaleCtx { float scale, rscale; bool unscale32; };

template <int DT>
__device__ __forceinline__ float2 get_pair(const Vec32& v, int k) {
  if (DT == FPEMU_DT_F32) return make_float2(__uint_as_float(v.w[2 * k]), __uint_as_float(v.w[2 * k + 1]));
  if (DT == FPEMU_DT_F16) return unpack_f16x2(v.w[k]);
  return unpack_bf16x2(v.w[k]);
}
template <int DT>
__device__ __forceinline__ void put_pair(Vec32& v, int k, float2 f) {
  if (DT == FPEMU_DT_F32) { v.w[2 * k] = __float_as_uint(f.x); v.w[2 * k + 1] = __float_as_uint(f.y); }
  else if (DT == FPEMU_DT_F16) v.w[k] = pack_f16x2_rn(f.x, f.y);
  else v.w[k] = pack_bf16x2_rn(f.x, f.y);
}

// BFLOAT16 target from fp32 / fp16 tensors: packed path, element-wise redo only for vectors holding a NaN.
// 16 random bits per element: element g uses 16-bit lane (g % 8) of Philox(g / 8), i.e. pair k of the vector
// uses word ((2k) % 8) / 2 of call (2k) / 8.
template <int RND, int DT>
__device__ __forceinline__ void process_vec_bf16(Vec32& v, const CastArgs& a, uint64_t g0, const uint32_t (&rl)[8]) {
  constexpr int kPairs = DtInfo<DT>::kVecElems / 2;
  constexpr bool kSR = (RND == FPEMU_RND_STOCHASTIC);
  const Vec32 orig = v;
  uint32_t nan_any = 0u;
  uint32_t packed[kPairs];
#pragma unroll
  for (int k = 0; k < kPairs; ++k) {
    const uint32_t rbits = kSR ? rl[k] : 0u;
    const float2 f = get_pair<DT>(v, k);
    packed[k] = round_pair_bf16<RND>(f.x, f.y, rbits, nan_any);
  }
  if (nan_any == 0u) {
#pragma unroll
    for (int k = 0; k < kPairs; ++k) put_pair<DT>(v, k, unpack_bf16x2(packed[k]));
  } else {
    // rare: redo the vector element by element with the reference's integer rule (keeps NaN payload high bits)
#pragma unroll
    for (int k = 0; k < kPairs; ++k) {
      const uint32_t rbits = kSR ? rl[k] : 0u;
      float2 f = get_pair<DT>(orig, k);
      f.x = __uint_as_float(round_one_bf16(__float_as_uint(f.x), RND, rbits & 0xFFFFu));
      f.y = __uint_as_float(round_one_bf16(__float_as_uint(f.y), RND, rbits >> 16));
      put_pair<DT>(v, k, f);
    }
  }
}

// Scale set-up shared by the vector kernels: host scalar, device scalar, or derived from a device-resident
// absolute maximum (scale_from_amax(), mpemu/qutils.py:283-285).
template <int DT, bool SCALED>
__device__ __forceinline__ ScaleCtx make_scale_ctx(const CastArgs& a) {
  ScaleCtx sc;
  sc.scale = 1.0f; sc.rscale = 1.0f;
  sc.unscale32 = (a.flags & FPEMU_FLAG_UNSCALE_FP32) != 0;
  if (SCALED) {
    float s = a.d_scale ? __ldg(a.d_scale) : a.scale;
    if (a.d_amax != nullptr) {                // per-tensor amax-scaled cast without a separate scale launch
      s = scale_from_amax(__ldg(a.d_amax), a.flt_max, true, DT);
      if (a.d_scale_out != nullptr && blockIdx.x == 0 && threadIdx.x == 0) *a.d_scale_out = s;
    }
    s = effective_scale(s, DT);
    sc.scale = s;
    sc.rscale = __frcp_rn(s);                 // == (float)(1.0 / (double)s): `float r = 1.0 / scale;`
  }
  return sc;
}

// Bytes [shift, shift + 16) of the 32-byte string Philox(block) | Philox(block + 1): the random field of a vector that
// does not start on a Philox-block boundary (a slab of a sharded tensor whose first global element index is not a
// multiple of 16).  Only the ANYOFF kernel variants contain this code: its two draws and the select chains cost
// 20-odd registers, which the aligned case should not pay for.
__device__ __forceinline__ uint4 rand_window16(uint32_t seed_lo, uint32_t seed_hi, uint64_t block, uint32_t shift) {
  const uint4 r0 = philox4x32_10(seed_lo, seed_hi, block);
  const uint4 r1 = philox4x32_10(seed_lo, seed_hi, block + 1);
  const uint32_t w[8] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w};
  const uint32_t ws = shift >> 2, bs = (shift & 3u) * 8u;        // shift < 16: ws in 0..3
  uint32_t o[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const uint32_t lo = ws == 0 ? w[i] : ws == 1 ? w[i + 1] : ws == 2 ? w[i + 2] : w[i + 3];
    const uint32_t hi = ws == 0 ? w[i + 1] : ws == 1 ? w[i + 2] : ws == 2 ? w[i + 3] : w[i + 4];
    o[i] = __funnelshift_r(lo, hi, bs);
  }
  return make_uint4(o[0], o[1], o[2], o[3]);
}

// Random bytes of one vector for the 8-bit formats: element g uses byte (g % 16) of Philox(g / 16).
// ANYOFF = false: g0 is a multiple of the vector length and of 8 (the launch guarantees elem_offset % 16 == 0);
// ANYOFF = true : any g0.
template <int FMT, int RND, int DT, bool ANYOFF = false>
__device__ __forceinline__ void vec_rand_bytes(const CastArgs& a, uint64_t g0, uint32_t (&rw)[4]) {
  rw[0] = rw[1] = rw[2] = rw[3] = 0u;
  if (RND == FPEMU_RND_STOCHASTIC) {
    const uint32_t m = (FMT == FPEMU_FMT_E5M2) ? 0xFFFFFFFFu
                     : ((FMT == FPEMU_FMT_E4M3 || FMT == FPEMU_FMT_E4M3_IEEE) ? 0x7F7F7F7Fu : 0x3F3F3F3Fu);
    if (ANYOFF) {
      const uint4 r = rand_window16(a.seed_lo, a.seed_hi, g0 >> 4, (uint32_t)g0 & 15u);
      rw[0] = r.x & m; rw[1] = r.y & m;
      if (DT != FPEMU_DT_F32) { rw[2] = r.z & m; rw[3] = r.w & m; }
    } else {
      const uint4 r = philox4x32_10(a.seed_lo, a.seed_hi, g0 >> 4);
      if (DT == FPEMU_DT_F32) {                       // 8 elements: half of the 16-byte draw
        const bool hi = (g0 >> 3) & 1u;
        rw[0] = (hi ? r.z : r.x) & m; rw[1] = (hi ? r.w : r.y) & m;
      } else {
        rw[0] = r.x & m; rw[1] = r.y & m; rw[2] = r.z & m; rw[3] = r.w & m;
      }
    }
  }
}

// 16-bit random lanes of one vector for the BFLOAT16 / FLOAT16 targets: element g uses lane (g % 8) of Philox(g / 8);
// word k of `rl` holds the lanes of elements g0 + 2k (low half) and g0 + 2k + 1 (high half).
template <int DT, bool ANYOFF = false>
__device__ __forceinline__ void vec_rand_lanes(const CastArgs& a, uint64_t g0, uint32_t (&rl)[8]) {
  constexpr int NW = DtInfo<DT>::kVecElems / 2;                 // 4 words (fp32 tensors) or 8 (16-bit tensors)
  rl[4] = rl[5] = rl[6] = rl[7] = 0u;
  if (!ANYOFF) {
    const uint4 r0 = philox4x32_10(a.seed_lo, a.seed_hi, g0 >> 3);
    rl[0] = r0.x; rl[1] = r0.y; rl[2] = r0.z; rl[3] = r0.w;
    if (NW > 4) {
      const uint4 r1 = philox4x32_10(a.seed_lo, a.seed_hi, (g0 >> 3) + 1);
      rl[4] = r1.x; rl[5] = r1.y; rl[6] = r1.z; rl[7] = r1.w;
    }
  } else {
    const uint32_t sh = 2u * ((uint32_t)g0 & 7u);
    const uint4 r0 = rand_window16(a.seed_lo, a.seed_hi, g0 >> 3, sh);
    rl[0] = r0.x; rl[1] = r0.y; rl[2] = r0.z; rl[3] = r0.w;
    if (NW > 4) {
      const uint4 r1 = rand_window16(a.seed_lo, a.seed_hi, (g0 >> 3) + 1, sh);
      rl[4] = r1.x; rl[5] = r1.y; rl[6] = r1.z; rl[7] = r1.w;
    }
  }
}

// Pair k of the vector as the fp16x2 encoding the 8-bit roundings start from (scale applied, fp16 RNE).
template <int DT, bool SCALED>
__device__ __forceinline__ uint32_t pair_to_h(const Vec32& v, int k, const ScaleCtx& sc) {
  if (DT == FPEMU_DT_F16 && !SCALED) {
    // the reference multiplies by scale even when it is 1 (`in[gid] * scale`, fpemu_kernels.cu:420, 588),
    // which turns every NaN encoding into the canonical 0x7FFF; HMUL2 by 1.0 does the same and is exact
    // for everything else
    return as_u32(__hmul2(as_h2(v.w[k]), as_h2(0x3C003C00u)));
  }
  float2 f = get_pair<DT>(v, k);
  if (SCALED) { f.x *= sc.scale; f.y *= sc.scale; }
  return pack_f16x2_rn(f.x, f.y);
}

// The rounded fp16x2 pair q -> un-scaled pair k of the output vector.
template <int DT, bool SCALED>
__device__ __forceinline__ void finish_pair(Vec32& v, int k, uint32_t q, const ScaleCtx& sc) {
  if (DT == FPEMU_DT_F16) {
    if (SCALED) {
      float2 g = unpack_f16x2(q);
      q = pack_f16x2_rn(g.x * sc.rscale, g.y * sc.rscale);
    }
    v.w[k] = q;
  } else {
    float2 g = unpack_f16x2(q);
    if (SCALED) {
      g.x *= sc.rscale; g.y *= sc.rscale;
      if (!sc.unscale32) g = unpack_f16x2(pack_f16x2_rn(g.x, g.y));   // reference CUDA op: through fp16
    }
    put_pair<DT>(v, k, g);
  }
}

// One 32-byte vector whose first element has global index g0: a multiple of the vector length on a launch whose
// elem_offset is a multiple of 16, or any value in the ANYOFF instantiations.
template <int FMT, int RND, bool DAZ, int DT, bool SCALED, bool ANYOFF = false>
__device__ __forceinline__ void process_vec(Vec32& v, const ScaleCtx& sc, const CastArgs& a, uint64_t g0) {
  constexpr int kPairs = DtInfo<DT>::kVecElems / 2;
  constexpr bool kSR = (RND == FPEMU_RND_STOCHASTIC);

  if (FMT == FPEMU_FMT_E5M2 || FMT == FPEMU_FMT_E4M3 || FMT == FPEMU_FMT_E3M4 || FMT == FPEMU_FMT_E4M3_IEEE) {
    uint32_t rw[4];
    vec_rand_bytes<FMT, RND, DT, ANYOFF>(a, g0, rw);
#pragma unroll
    for (int k = 0; k < kPairs; ++k) {
      const uint32_t h = pair_to_h<DT, SCALED>(v, k, sc);
      const uint32_t r2 = kSR ? __byte_perm(rw[k >> 1], 0u, (k & 1) ? 0x4342u : 0x4140u) : 0u;
      finish_pair<DT, SCALED>(v, k, round_pair<FMT, RND, DAZ>(h, r2), sc);
    }
  } else if (FMT == FPEMU_FMT_FP4) {
    // FP4_NEAREST (fpemu_kernels.cu:986-1024): `scale * in` in fp32, rounded on the upper halves of two products at a
    // time (round_pair_fp4), un-scaled in fp32.  An unscaled bf16 tensor never leaves its own encoding.
#pragma unroll
    for (int k = 0; k < kPairs; ++k) {
      if (DT == FPEMU_DT_BF16 && !SCALED) {
        v.w[k] = round_pair_fp4(v.w[k]);
      } else {
        float2 f = get_pair<DT>(v, k);
        // always a multiplication, as in the reference (by 1 when unscaled): it is what turns a NaN into +NaN
        f.x = __fmul_rn(f.x, SCALED ? sc.scale : 1.0f);
        f.y = __fmul_rn(f.y, SCALED ? sc.scale : 1.0f);
        const uint32_t q = round_pair_fp4(__byte_perm(__float_as_uint(f.x), __float_as_uint(f.y), 0x7632));
        float2 g = unpack_bf16x2(q);
        if (SCALED) { g.x *= sc.rscale; g.y *= sc.rscale; }
        put_pair<DT>(v, k, g);
      }
    }
  } else if (FMT == FPEMU_FMT_BF16 && DT == FPEMU_DT_BF16) {
    // a bf16 tensor cast to BFLOAT16: every rounding mode is the identity (stochastic bits are added below the
    // kept bits and never carry); only the flush of exponent-field-0 values to +0 remains
#pragma unroll
    for (int k = 0; k < kPairs; ++k) {
      const uint32_t aabs = v.w[k] & kAbs2;
      const uint32_t tiny = __hlt2_mask(*reinterpret_cast<const __nv_bfloat162*>(&aabs),
                                        __float2bfloat162_rn(1.17549435e-38f));      // NaN lanes compare false
      v.w[k] &= ~tiny;
    }
  } else if (FMT == FPEMU_FMT_BF16) {
    uint32_t rl[8] = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};
    if (kSR) vec_rand_lanes<DT, ANYOFF>(a, g0, rl);
    if (kSR && DT == FPEMU_DT_F16) {
      // fp16 tensor, stochastic BFLOAT16: for fp16-normal values only the 3 lowest fp16 mantissa bits are dropped and
      // only the top 3 of the 16 random bits can carry into the kept part ((m3 << 13) + r16 >= 2^16  <=>  m3 + (r16 >> 13)
      // >= 8), so the rounding is `(h + (r16 >> 13)) & ~7` on the fp16 encoding itself; a carry into the exponent
      // reaches Inf exactly where the fp32-domain result (65536) converts back to fp16 Inf.  Zeros become +0.
      // Vectors with a NaN or a non-zero fp16 subnormal (normalised in fp32: different cut position) take the
      // general path below.
      uint32_t special = 0u, q[kPairs];
#pragma unroll
      for (int k = 0; k < kPairs; ++k) {
        const uint32_t rbits = rl[k];
        const uint32_t h = v.w[k];
        const __half2 av = as_h2(h & kAbs2);
        const uint32_t nrm = __hge2_mask(av, as_h2(h2c(0x0400)));          // normal or Inf (false for NaN)
        const uint32_t zer = __heq2_mask(av, as_h2(0u));
        special |= ~(nrm | zer);
        const uint32_t r3 = __umulhi(rbits, 1u << 19) & 0x00070007u;       // (r16 >> 13) per lane
        q[k] = (h + r3) & 0xFFF8FFF8u & ~zer;
      }
      if (special == 0u) {
#pragma unroll
        for (int k = 0; k < kPairs; ++k) v.w[k] = q[k];
        return;
      }
    }
    process_vec_bf16<RND, DT>(v, a, g0, rl);
  } else {
    // FLOAT16 target: per element; 13 random bits per element from the same 16-bit lanes as BFLOAT16
    const bool daz = (a.flags & FPEMU_FLAG_DAZ) != 0;
    uint32_t rl[8] = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};
    if (kSR) vec_rand_lanes<DT, ANYOFF>(a, g0, rl);
    if (kSR && DT != FPEMU_DT_F16) {
      // Stochastic FLOAT16 of fp32 / bf16 values (fpemu_kernels.cu:340-372): inside the fp16 normal range
      // [2^-14, 2^16) the reference adds 13 random bits under the kept mantissa and truncates — done here on the fp32
      // encoding, after which the conversion to fp16 is exact (a carry up to 2^16 converts to Inf exactly as the
      // reference's exponent carry does); everything else (fp16-subnormal range: plain RNE; >= 2^16 and Inf: Inf)
      // is what cvt.rn.f16.f32 does by itself.  NaN keeps sign and payload in the reference but not in the
      // hardware conversion: vectors holding a NaN take the element-wise code below.
      uint32_t nan_any = 0u, hp[kPairs];
#pragma unroll
      for (int k = 0; k < kPairs; ++k) {
        const uint32_t rbits = rl[k];
        const float2 f = get_pair<DT>(v, k);
        uint32_t u0 = __float_as_uint(f.x), u1 = __float_as_uint(f.y);
        if (((u0 & 0x7FFFFFFFu) - 0x38800000u) < 0x0F000000u) u0 = (u0 + (rbits & 0x1FFFu)) & 0xFFFFE000u;
        if (((u1 & 0x7FFFFFFFu) - 0x38800000u) < 0x0F000000u) u1 = (u1 + ((rbits >> 16) & 0x1FFFu)) & 0xFFFFE000u;
        hp[k] = pack_f16x2_rn(__uint_as_float(u0), __uint_as_float(u1));
        nan_any |= __hneu2_mask(as_h2(hp[k]), as_h2(hp[k]));
      }
      if (nan_any == 0u) {
#pragma unroll
        for (int k = 0; k < kPairs; ++k) put_pair<DT>(v, k, unpack_f16x2(hp[k]));
        return;
      }
    }
#pragma unroll
    for (int k = 0; k < kPairs; ++k) {
      const uint32_t rbits = kSR ? rl[k] : 0u;
      float2 f = get_pair<DT>(v, k);
      const bool ih = (DT == FPEMU_DT_F16);
      f.x = __half2float(__ushort_as_half((unsigned short)round_one_fp16(f.x, RND, daz, ih, rbits & 0xFFFFu)));
      f.y = __half2float(__ushort_as_half((unsigned short)round_one_fp16(f.y, RND, daz, ih, rbits >> 16)));
      put_pair<DT>(v, k, f);
    }
  }
}

}  // namespace fpemu
