// fpemu_codec_core.cuh — per-vector and per-CTA code of the 8-bit encode, shared by fpemu_codec.cu and fpemu_multi.cu.
#pragma once

#include "fpemu_cast_core.cuh"

namespace fpemu {

// Vectors in flight per thread of the single-tensor encode / decode kernels (reads are 4/5 of the encode traffic).
#ifndef FPEMU_CODEC_U32
#define FPEMU_CODEC_U32 4
#endif
#ifndef FPEMU_CODEC_U16
#define FPEMU_CODEC_U16 4
#endif
template <int DT> struct CodecUnroll { static constexpr int value = (DT == FPEMU_DT_F32) ? FPEMU_CODEC_U32 : FPEMU_CODEC_U16; };

__device__ __forceinline__ void stg_codes(void* p, const uint32_t (&w)[4], int words) {
  if (words == 2) asm volatile("st.global.L1::no_allocate.v2.u32 [%0], {%1,%2};" :: "l"(p), "r"(w[0]), "r"(w[1]) : "memory");
  else asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" :: "l"(p), "r"(w[0]), "r"(w[1]), "r"(w[2]), "r"(w[3]) : "memory");
}
__device__ __forceinline__ void ldg_codes(const void* p, uint32_t (&w)[4], int words) {
  if (words == 2) {
    asm volatile("ld.global.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(w[0]), "=r"(w[1]) : "l"(p));
    w[2] = w[3] = 0u;
  } else {
    asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]) : "l"(p));
  }
}

// fp16x2 -> two E4M3 codes with the hardware conversion (F2FP.SATFINITE.E4M3.F16): round to nearest even at 3
// mantissa bits, saturate at 448.  For |h| >= 2^-6 (the format's normal range) and for zeros this is bit for bit the
// reference's E4M3 RNE followed by the code extraction: same rounding position, Inf and overflow -> 0x7E.  It differs
// where the reference does: NaN (hardware 0x7F, reference 448 = 0x7E) and 0 < |h| < 2^-6 (hardware rounds once onto the
// subnormal grid; the reference flushes below 2^-9 and rounds its pseudo-subnormals twice).  Those lanes are
// reported in `other` and their vector is redone with the exact sequence.
__device__ __forceinline__ uint32_t cvt_e4m3x2_rn(uint32_t h2) {
  unsigned short r;
  asm("cvt.rn.satfinite.e4m3x2.f16x2 %0, %1;" : "=h"(r) : "r"(h2));
  return r;
}

// One input vector -> its codes, 4 per word.  `rw` = the vector's random bytes (vec_rand_bytes layout).
template <int FMT, int RND, bool DAZ, int DT, bool SCALED>
__device__ __forceinline__ void encode_vec(const Vec32& v, const ScaleCtx& sc, const uint32_t (&rw)[4], uint32_t (&w)[4],
                                           const bool expect_normals) {
  constexpr int kPairs = DtInfo<DT>::kVecElems / 2;
  constexpr bool kSR = (RND == FPEMU_RND_STOCHASTIC);
  // The hardware path pays off only when almost no warp meets a lane below the smallest normal (a thread that does
  // redoes its vector, and its whole warp waits): the caller says so with FPEMU_FLAG_EXPECT_NORMALS, typically
  // because the scale was derived from the tensor's maximum.  The codes are the same either way.
  if (FMT == FPEMU_FMT_E4M3 && RND == FPEMU_RND_RNE && expect_normals) {
    uint32_t other = 0u, c[kPairs];
#pragma unroll
    for (int k = 0; k < kPairs; ++k) {
      const uint32_t h = pair_to_h<DT, SCALED>(v, k, sc);
      const __half2 av = __habs2(as_h2(h));
      // (|h| < 2^-6 or NaN) and not zero
      other |= __hltu2_mask(av, as_h2(0x24002400u)) & __hneu2_mask(av, as_h2(0u));
      c[k] = cvt_e4m3x2_rn(h);
    }
    if (other == 0u) {
      w[0] = w[1] = w[2] = w[3] = 0u;
#pragma unroll
      for (int k = 0; k < kPairs; k += 2) w[k >> 1] = c[k] | (c[k + 1] << 16);
      return;
    }
  }
  uint32_t e[kPairs];
#pragma unroll
  for (int k = 0; k < kPairs; ++k) {
    const uint32_t h = pair_to_h<DT, SCALED>(v, k, sc);
    const uint32_t r2 = kSR ? __byte_perm(rw[k >> 1], 0u, (k & 1) ? 0x4342u : 0x4140u) : 0u;
    if (FMT == FPEMU_FMT_E4M3 || FMT == FPEMU_FMT_E3M4) e[k] = encode_pair_mag<FMT>(round_pair_e4m3_like<FMT, RND, 1>(h, r2), h);
    else e[k] = encode_pair<FMT>(round_pair<FMT, RND, DAZ>(h, r2));
  }
  w[0] = w[1] = w[2] = w[3] = 0u;
#pragma unroll
  for (int k = 0; k < kPairs; k += 2) w[k >> 1] = __byte_perm(e[k], e[k + 1], 0x6420);
}

// The work of CTA `local_cta` of a tensor: U vectors per thread, and the tensor's ragged tail when `last_cta`.
// Shared by the single-tensor kernel (fpemu_codec.cu) and the multi-tensor one (fpemu_multi.cu).
template <int FMT, int RND, bool DAZ, int DT, bool SCALED, int U>
__device__ __forceinline__ void encode_cta(const CastArgs& a, const ScaleCtx& sc, int64_t local_cta, bool last_cta) {
  constexpr int VE = DtInfo<DT>::kVecElems;
  constexpr bool kSR = (RND == FPEMU_RND_STOCHASTIC);
  const int64_t nvec = a.n / VE;
  const char* in = static_cast<const char*>(a.in);
  char* out = static_cast<char*>(a.out);
  const int64_t base = local_cta * (kThreads * U) + threadIdx.x;
  const bool expect_normals = (a.flags & FPEMU_FLAG_EXPECT_NORMALS) != 0;
  Vec32 v[U];
#pragma unroll
  for (int j = 0; j < U; ++j) {
    const int64_t vi = base + j * kThreads;
    if (vi < nvec) v[j] = ldg256(in + vi * 32);
  }
  uint32_t rw[U][4];
  if (kSR && DT == FPEMU_DT_F32) {
    // An fp32 vector is 8 elements = half of a 16-byte Philox draw, and lanes 2m / 2m+1 hold the two halves of the
    // same draw in every slot j (vector index parity == lane parity, elem_offset % 16 == 0).  Each lane therefore
    // draws for every other slot only and the pair swaps halves with one shuffle per word: half the Philox work.
    // All 32 lanes run this (no bounds guard before the shuffles).
    static_assert(U % 2 == 0, "slot pairing");
    const bool odd = (threadIdx.x & 1) != 0;
    constexpr uint32_t m = (FMT == FPEMU_FMT_E5M2) ? 0xFFFFFFFFu : (FMT == FPEMU_FMT_E4M3 ? 0x7F7F7F7Fu : 0x3F3F3F3Fu);
#pragma unroll
    for (int j = 0; j < U; j += 2) {
      const int64_t vi = base + (int64_t)(j + (odd ? 1 : 0)) * kThreads;          // the slot this lane draws for
      const uint4 r = philox4x32_10(a.seed_lo, a.seed_hi, (a.elem_offset + (uint64_t)vi * VE) >> 4);
      const uint32_t mine0 = (odd ? r.z : r.x) & m, mine1 = (odd ? r.w : r.y) & m;   // my half: odd lanes use z, w
      const uint32_t got0 = __shfl_xor_sync(0xFFFFFFFFu, (odd ? r.x : r.z) & m, 1);
      const uint32_t got1 = __shfl_xor_sync(0xFFFFFFFFu, (odd ? r.y : r.w) & m, 1);
      rw[j][0] = odd ? got0 : mine0;     rw[j][1] = odd ? got1 : mine1;
      rw[j + 1][0] = odd ? mine0 : got0; rw[j + 1][1] = odd ? mine1 : got1;
      rw[j][2] = rw[j][3] = rw[j + 1][2] = rw[j + 1][3] = 0u;
    }
  } else {
#pragma unroll
    for (int j = 0; j < U; ++j)
      vec_rand_bytes<FMT, RND, DT>(a, a.elem_offset + (uint64_t)(base + j * kThreads) * VE, rw[j]);
  }
#pragma unroll
  for (int j = 0; j < U; ++j) {
    const int64_t vi = base + j * kThreads;
    if (vi < nvec) {
      uint32_t w[4];
      encode_vec<FMT, RND, DAZ, DT, SCALED>(v[j], sc, rw[j], w, expect_normals);
      stg_codes(out + vi * VE, w, VE / 4);
    }
  }
  // ragged tail (< one vector): zero-padded register vector through the same code, byte stores
  const int tail = (int)(a.n - nvec * VE);
  if (tail > 0 && last_cta && threadIdx.x == 0) {
    Vec32 t;
#pragma unroll
    for (int i = 0; i < 8; ++i) t.w[i] = 0u;
    const int64_t e0 = nvec * VE;
    if (DT == FPEMU_DT_F32) {
      const uint32_t* p = reinterpret_cast<const uint32_t*>(a.in) + e0;
      for (int i = 0; i < tail; ++i) t.w[i] = p[i];
    } else {
      const unsigned short* p = reinterpret_cast<const unsigned short*>(a.in) + e0;
      for (int i = 0; i < tail; ++i) t.w[i >> 1] |= (uint32_t)p[i] << ((i & 1) * 16);
    }
    uint32_t w[4], rt[4];
    vec_rand_bytes<FMT, RND, DT>(a, a.elem_offset + (uint64_t)e0, rt);
    encode_vec<FMT, RND, DAZ, DT, SCALED>(t, sc, rt, w, expect_normals);
    unsigned char* c = reinterpret_cast<unsigned char*>(a.out) + e0;
    for (int i = 0; i < tail; ++i) c[i] = (unsigned char)(w[i >> 2] >> ((i & 3) * 8));
  }
}


}  // namespace fpemu
