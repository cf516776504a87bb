// fpemu_cast.cu — the elementwise quantize-dequantize cast (C ABI: fpemu_cast).
//
// Replaces the kernel launches of the reference's fpemu_cuda_forward
// (mpemu/pytquant/cuda/fpemu_kernels.cu:1028-1381) for block_norm == false.
//
// Data movement: one 256-bit LDG + one 256-bit STG per thread per vector (8 fp32 / 16 half elements),
// 2 (fp32) or 4 (16-bit) vectors in flight per thread, warp-contiguous (thread t of a CTA touches vector
// base + t), one pass per thread.  No shared memory, no scratch memory.
// Algorithmic traffic: read 1 + write 1 element = 8 B/elem (fp32) or 4 B/elem (fp16 / bf16).
#include "fpemu_cast_core.cuh"

namespace fpemu {

// ANYOFF: stochastic rounding of a slab whose elem_offset is not a multiple of 16 (the random field of a vector then
// straddles Philox blocks); a separate instantiation so that the aligned case keeps its register count.
template <int FMT, int RND, bool DAZ, int DT, bool SCALED, bool ANYOFF = false>
__global__ void __launch_bounds__(kThreads) cast_vec_kernel(const CastArgs a_) {
  const CastArgs a = with_device_seed(a_);
  constexpr int VE = DtInfo<DT>::kVecElems;
  const ScaleCtx sc = make_scale_ctx<DT, SCALED>(a);
  const int64_t nvec = a.n / VE;
  const char* in = static_cast<const char*>(a.in);
  char* out = static_cast<char*>(a.out);
  // One pass per thread, no grid-stride loop: on B200 a grid of ceil(nvec / (kThreads * U)) short-lived CTAs
  // streams at ~6.9 TB/s while a persistent 148 x k grid of looping CTAs stays at 6.0-6.6 TB/s
  // (profiles/tune_r1.txt) — the block scheduler overlaps one CTA's stores with the next one's loads.
  constexpr int U = UnrollFor<DT>::value;
  const int64_t base = (int64_t)blockIdx.x * (kThreads * U) + threadIdx.x;
  Vec32 v[U];
#pragma unroll
  for (int j = 0; j < U; ++j) {
    const int64_t vi = base + j * kThreads;
    if (vi < nvec) v[j] = ldg256(in + vi * 32);
  }
#pragma unroll
  for (int j = 0; j < U; ++j) {
    const int64_t vi = base + j * kThreads;
    if (vi < nvec) {
      process_vec<FMT, RND, DAZ, DT, SCALED, ANYOFF>(v[j], sc, a, a.elem_offset + (uint64_t)vi * VE);
      stg256(out + vi * 32, v[j]);
    }
  }
  // ragged tail (< one vector): staged through a zero-padded register vector so it runs the same code
  const int tail = (int)(a.n - nvec * VE);
  if (tail > 0 && blockIdx.x == gridDim.x - 1 && threadIdx.x == 0) {
    Vec32 v;
#pragma unroll
    for (int i = 0; i < 8; ++i) v.w[i] = 0u;
    const int64_t e0 = nvec * VE;
    if (DT == FPEMU_DT_F32) {
      const uint32_t* p = reinterpret_cast<const uint32_t*>(a.in) + e0;
      for (int i = 0; i < tail; ++i) v.w[i] = p[i];
    } else {
      const unsigned short* p = reinterpret_cast<const unsigned short*>(a.in) + e0;
      for (int i = 0; i < tail; ++i) v.w[i >> 1] |= (uint32_t)p[i] << ((i & 1) * 16);
    }
    process_vec<FMT, RND, DAZ, DT, SCALED, ANYOFF>(v, sc, a, a.elem_offset + (uint64_t)e0);
    if (DT == FPEMU_DT_F32) {
      uint32_t* p = reinterpret_cast<uint32_t*>(a.out) + e0;
      for (int i = 0; i < tail; ++i) p[i] = v.w[i];
    } else {
      unsigned short* p = reinterpret_cast<unsigned short*>(a.out) + e0;
      for (int i = 0; i < tail; ++i) p[i] = (unsigned short)(v.w[i >> 1] >> ((i & 1) * 16));
    }
  }
}

// Input and output misaligned by DIFFERENT amounts (a direct C-ABI caller casting from one view into another): vectors
// are cut on the OUTPUT's 32-byte boundaries, so stores stay 256-bit; the matching 32 input bytes start `d` bytes into
// an aligned 32-byte block and are assembled from the two aligned blocks they straddle (the second one is the next
// thread's first: an L1 hit, DRAM traffic is unchanged) with a uniform funnel shift.  The upper block of the last
// vector is read in full: it holds at least one valid byte (d > 0), so the read never leaves the buffer's last sector.
template <int FMT, int RND, int DT, bool SCALED>
__global__ void __launch_bounds__(kThreads) cast_shift_kernel(const CastArgs a_, const int d) {
  const CastArgs a = with_device_seed(a_);
  constexpr int VE = DtInfo<DT>::kVecElems;
  constexpr int U = 2;
  constexpr bool kAny = (RND == FPEMU_RND_STOCHASTIC);
  const ScaleCtx sc = make_scale_ctx<DT, SCALED>(a);
  const int64_t nvec = a.n / VE;
  const char* in = static_cast<const char*>(a.in) - d;          // the aligned block the first vector starts in
  char* out = static_cast<char*>(a.out);
  const int ws = d >> 2, bs = (d & 3) * 8;
  const int64_t base = (int64_t)blockIdx.x * (kThreads * U) + threadIdx.x;
  Vec32 lo[U], hi[U];
#pragma unroll
  for (int j = 0; j < U; ++j) {
    const int64_t vi = base + j * kThreads;
    if (vi < nvec) {
      lo[j] = *reinterpret_cast<const Vec32*>(in + vi * 32);     // default caching: the neighbour re-reads this block
      hi[j] = *reinterpret_cast<const Vec32*>(in + vi * 32 + 32);
    }
  }
#pragma unroll
  for (int j = 0; j < U; ++j) {
    const int64_t vi = base + j * kThreads;
    if (vi >= nvec) continue;
    Vec32 v;
    switch (ws) {
#define FPEMU_SHIFT_CASE(K)                                                                              \
      case K:                                                                                            \
        _Pragma("unroll") for (int i = 0; i < 8; ++i) {                                                  \
          const uint32_t x0 = (K + i < 8) ? lo[j].w[(K + i) & 7] : hi[j].w[(K + i) & 7];                 \
          const uint32_t x1 = (K + i + 1 < 8) ? lo[j].w[(K + i + 1) & 7] : hi[j].w[(K + i + 1) & 7];     \
          v.w[i] = __funnelshift_r(x0, x1, bs);                                                          \
        }                                                                                                \
        break;
      FPEMU_SHIFT_CASE(0) FPEMU_SHIFT_CASE(1) FPEMU_SHIFT_CASE(2) FPEMU_SHIFT_CASE(3)
      FPEMU_SHIFT_CASE(4) FPEMU_SHIFT_CASE(5) FPEMU_SHIFT_CASE(6)
      default: FPEMU_SHIFT_CASE(7)
#undef FPEMU_SHIFT_CASE
    }
    process_vec<FMT, RND, false, DT, SCALED, kAny>(v, sc, a, a.elem_offset + (uint64_t)vi * VE);
    stg256(out + vi * 32, v);
  }
  // ragged tail (< one vector): element-wise loads, same register path
  const int tail = (int)(a.n - nvec * VE);
  if (tail > 0 && blockIdx.x == gridDim.x - 1 && threadIdx.x == 0) {
    Vec32 v;
#pragma unroll
    for (int i = 0; i < 8; ++i) v.w[i] = 0u;
    const int64_t e0 = nvec * VE;
    if (DT == FPEMU_DT_F32) {
      const uint32_t* p = reinterpret_cast<const uint32_t*>(a.in) + e0;
      for (int i = 0; i < tail; ++i) v.w[i] = p[i];
    } else {
      const unsigned short* p = reinterpret_cast<const unsigned short*>(a.in) + e0;
      for (int i = 0; i < tail; ++i) v.w[i >> 1] |= (uint32_t)p[i] << ((i & 1) * 16);
    }
    process_vec<FMT, RND, false, DT, SCALED, kAny>(v, sc, a, a.elem_offset + (uint64_t)e0);
    if (DT == FPEMU_DT_F32) {
      uint32_t* p = reinterpret_cast<uint32_t*>(a.out) + e0;
      for (int i = 0; i < tail; ++i) p[i] = v.w[i];
    } else {
      unsigned short* p = reinterpret_cast<unsigned short*>(a.out) + e0;
      for (int i = 0; i < tail; ++i) p[i] = (unsigned short)(v.w[i >> 1] >> ((i & 1) * 16));
    }
  }
}

// Every format x rounding mode, any alignment, per-element random draws.  Correctness path, not the fast one.
__global__ void __launch_bounds__(256) cast_generic_kernel(const CastArgs a_, const int dt) {
  const CastArgs a = with_device_seed(a_);
  const bool scaled = (a.d_scale != nullptr) || (a.d_amax != nullptr) || (a.scale != 1.0f);
  float s = 1.0f, rs = 1.0f;
  if (scaled) {
    s = a.d_scale ? __ldg(a.d_scale) : a.scale;
    if (a.d_amax != nullptr) {
      s = scale_from_amax(__ldg(a.d_amax), a.flt_max, true, dt);
      if (a.d_scale_out != nullptr && blockIdx.x == 0 && threadIdx.x == 0) *a.d_scale_out = s;
    }
    s = effective_scale(s, dt);
    rs = __frcp_rn(s);
  }
  const bool wide = (a.fmt == FPEMU_FMT_BF16 || a.fmt == FPEMU_FMT_FP16);
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < a.n; i += stride) {
    const float x = load_elem(a.in, i, dt);
    uint32_t rb = 0u;
    if (a.rnd == FPEMU_RND_STOCHASTIC) rb = rand_bits_one(a.seed_lo, a.seed_hi, a.elem_offset + (uint64_t)i, wide);
    store_elem(a.out, i, dt, cast_one(x, a.fmt, a.rnd, a.flags, scaled, s, rs, dt == FPEMU_DT_F16, rb));
  }
}

// Same rare modes, but aligned buffers: 256-bit vector I/O, one Philox draw per vector, element-wise rounding.
// (E4M3_IEEE, FP4; anything else without a packed kernel runs the fully run-time variant.)
// CFMT / CRND >= 0 fix the format / rounding at compile time (the element-wise code then folds to a few dozen
// instructions); -1 reads them from the arguments.
template <int DT, int CFMT, int CRND>
__global__ void __launch_bounds__(kThreads) cast_vec_generic_kernel(const CastArgs a_) {
  constexpr int VE = DtInfo<DT>::kVecElems;
  constexpr int U = 2;
  CastArgs a = a_;
  if (CFMT >= 0) a.fmt = CFMT;
  if (CRND >= 0) a.rnd = CRND;
  a = with_device_seed(a);
  const bool scaled = (a.d_scale != nullptr) || (a.d_amax != nullptr) || (a.scale != 1.0f);
  float s = 1.0f, rs = 1.0f;
  if (scaled) {
    s = a.d_scale ? __ldg(a.d_scale) : a.scale;
    if (a.d_amax != nullptr) {
      s = scale_from_amax(__ldg(a.d_amax), a.flt_max, true, DT);
      if (a.d_scale_out != nullptr && blockIdx.x == 0 && threadIdx.x == 0) *a.d_scale_out = s;
    }
    s = effective_scale(s, DT);
    rs = __frcp_rn(s);
  }
  const bool sr = a.rnd == FPEMU_RND_STOCHASTIC;
  const bool wide = (a.fmt == FPEMU_FMT_BF16 || a.fmt == FPEMU_FMT_FP16);
  const int64_t nvec = a.n / VE;
  const char* in = static_cast<const char*>(a.in);
  char* out = static_cast<char*>(a.out);
  const int64_t base = (int64_t)blockIdx.x * (kThreads * U) + threadIdx.x;
  Vec32 v[U];
#pragma unroll
  for (int j = 0; j < U; ++j) {
    const int64_t vi = base + j * kThreads;
    if (vi < nvec) v[j] = ldg256(in + vi * 32);
  }
#pragma unroll
  for (int j = 0; j < U; ++j) {
    const int64_t vi = base + j * kThreads;
    if (vi >= nvec) continue;
    const uint64_t g0 = a.elem_offset + (uint64_t)vi * VE;
    // random fields of this vector: 8-bit formats use byte (g % 16) of Philox(g / 16), wide ones 16-bit lane
    // (g % 8) of Philox(g / 8); a vector never straddles more than two draws
    uint32_t rw[8] = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};
    if (sr) {
      const uint4 r0 = philox4x32_10(a.seed_lo, a.seed_hi, wide ? (g0 >> 3) : (g0 >> 4));
      rw[0] = r0.x; rw[1] = r0.y; rw[2] = r0.z; rw[3] = r0.w;
      if (wide && VE == 16) {
        const uint4 r1 = philox4x32_10(a.seed_lo, a.seed_hi, (g0 >> 3) + 1);
        rw[4] = r1.x; rw[5] = r1.y; rw[6] = r1.z; rw[7] = r1.w;
      }
    }
#pragma unroll
    for (int k = 0; k < VE / 2; ++k) {
      float2 f = get_pair<DT>(v[j], k);
      uint32_t rb0 = 0u, rb1 = 0u;
      if (sr) {
        if (wide) {
          rb0 = rw[k] & 0xFFFFu; rb1 = rw[k] >> 16;
        } else {
          const int e0 = (int)(g0 & 15u) + 2 * k;              // byte index of element 2k inside the 16-byte draw
          rb0 = (rw[e0 >> 2] >> ((e0 & 3) * 8)) & 0xFFu;
          rb1 = (rw[(e0 + 1) >> 2] >> (((e0 + 1) & 3) * 8)) & 0xFFu;
        }
      }
      f.x = cast_one(f.x, a.fmt, a.rnd, a.flags, scaled, s, rs, DT == FPEMU_DT_F16, rb0);
      f.y = cast_one(f.y, a.fmt, a.rnd, a.flags, scaled, s, rs, DT == FPEMU_DT_F16, rb1);
      put_pair<DT>(v[j], k, f);
    }
    stg256(out + vi * 32, v[j]);
  }
  // ragged tail: element-wise
  const int64_t e0 = nvec * VE;
  if (blockIdx.x == gridDim.x - 1 && threadIdx.x < (int)(a.n - e0)) {
    const int64_t i = e0 + threadIdx.x;
    const float x = load_elem(a.in, i, DT);
    const uint32_t rb = sr ? rand_bits_one(a.seed_lo, a.seed_hi, a.elem_offset + (uint64_t)i, wide) : 0u;
    store_elem(a.out, i, DT, cast_one(x, a.fmt, a.rnd, a.flags, scaled, s, rs, DT == FPEMU_DT_F16, rb));
  }
}

// ------------------------------------------------------------------------------------------ dispatch
typedef void (*CastKernel)(const CastArgs);

template <int FMT, int RND, bool DAZ, int DT>
static CastKernel pick_scaled(bool scaled, bool anyoff) {
  constexpr bool kSR = (RND == FPEMU_RND_STOCHASTIC);
  if (FMT == FPEMU_FMT_BF16 || FMT == FPEMU_FMT_FP16)
    return (kSR && anyoff) ? cast_vec_kernel<FMT, RND, DAZ, DT, false, kSR> : cast_vec_kernel<FMT, RND, DAZ, DT, false>;
  if (kSR && anyoff) return scaled ? cast_vec_kernel<FMT, RND, DAZ, DT, true, kSR> : cast_vec_kernel<FMT, RND, DAZ, DT, false, kSR>;
  return scaled ? cast_vec_kernel<FMT, RND, DAZ, DT, true> : cast_vec_kernel<FMT, RND, DAZ, DT, false>;
}

template <int FMT, int RND, bool DAZ>
static CastKernel pick_dt(int dt, bool scaled, bool anyoff) {
  switch (dt) {
    case FPEMU_DT_F32:  return pick_scaled<FMT, RND, DAZ, FPEMU_DT_F32>(scaled, anyoff);
    case FPEMU_DT_F16:  return pick_scaled<FMT, RND, DAZ, FPEMU_DT_F16>(scaled, anyoff);
    default:            return pick_scaled<FMT, RND, DAZ, FPEMU_DT_BF16>(scaled, anyoff);
  }
}

template <int FMT>
static CastKernel pick_rnd(int rnd, bool daz, int dt, bool scaled, bool anyoff) {
  if (FMT == FPEMU_FMT_E5M2 && daz) {
    if (rnd == FPEMU_RND_RNE) return pick_dt<FPEMU_FMT_E5M2, FPEMU_RND_RNE, true>(dt, scaled, anyoff);
    if (rnd == FPEMU_RND_STOCHASTIC) return pick_dt<FPEMU_FMT_E5M2, FPEMU_RND_STOCHASTIC, true>(dt, scaled, anyoff);
    if (rnd == FPEMU_RND_RNAZ) return pick_dt<FPEMU_FMT_E5M2, FPEMU_RND_RNAZ, true>(dt, scaled, anyoff);
    if (rnd == FPEMU_RND_RNTZ) return pick_dt<FPEMU_FMT_E5M2, FPEMU_RND_RNTZ, true>(dt, scaled, anyoff);
    return nullptr;
  }
  if (FMT == FPEMU_FMT_E5M2) {                    // the reference's extra E5M2 rounding modes (:1058-1158)
    if (rnd == FPEMU_RND_RNAZ) return pick_dt<FPEMU_FMT_E5M2, FPEMU_RND_RNAZ, false>(dt, scaled, anyoff);
    if (rnd == FPEMU_RND_RNTZ) return pick_dt<FPEMU_FMT_E5M2, FPEMU_RND_RNTZ, false>(dt, scaled, anyoff);
    if (rnd == FPEMU_RND_RPINF) return pick_dt<FPEMU_FMT_E5M2, FPEMU_RND_RPINF, false>(dt, scaled, anyoff);
    if (rnd == FPEMU_RND_RNINF) return pick_dt<FPEMU_FMT_E5M2, FPEMU_RND_RNINF, false>(dt, scaled, anyoff);
  }
  switch (rnd) {
    case FPEMU_RND_RNE:        return pick_dt<FMT, FPEMU_RND_RNE, false>(dt, scaled, anyoff);
    case FPEMU_RND_RTZ:        return (FMT == FPEMU_FMT_FP16) ? nullptr : pick_dt<FMT, FPEMU_RND_RTZ, false>(dt, scaled, anyoff);
    case FPEMU_RND_STOCHASTIC: return pick_dt<FMT, FPEMU_RND_STOCHASTIC, false>(dt, scaled, anyoff);
    default:                   return nullptr;
  }
}

template <int CFMT, int CRND>
static CastKernel pick_generic_dt(int dt) {
  switch (dt) {
    case FPEMU_DT_F32:  return cast_vec_generic_kernel<FPEMU_DT_F32, CFMT, CRND>;
    case FPEMU_DT_F16:  return cast_vec_generic_kernel<FPEMU_DT_F16, CFMT, CRND>;
    default:            return cast_vec_generic_kernel<FPEMU_DT_BF16, CFMT, CRND>;
  }
}

// Compile-time specialisations for the rare modes of the reference's grammar; anything else runs the
// fully run-time variant.
static CastKernel pick_generic(int fmt, int rnd, int dt) {
  if (fmt == FPEMU_FMT_FP4) return pick_generic_dt<FPEMU_FMT_FP4, FPEMU_RND_RNE>(dt);
  return pick_generic_dt<-1, -1>(dt);
}

static CastKernel pick_kernel(int fmt, int rnd, bool daz, int dt, bool scaled, bool anyoff) {
  switch (fmt) {
    case FPEMU_FMT_E5M2: return pick_rnd<FPEMU_FMT_E5M2>(rnd, daz, dt, scaled, anyoff);
    case FPEMU_FMT_E4M3: return pick_rnd<FPEMU_FMT_E4M3>(rnd, daz, dt, scaled, anyoff);
    case FPEMU_FMT_E3M4: return pick_rnd<FPEMU_FMT_E3M4>(rnd, daz, dt, scaled, anyoff);
    case FPEMU_FMT_BF16: return pick_rnd<FPEMU_FMT_BF16>(rnd, false, dt, false, anyoff);
    case FPEMU_FMT_FP16: return pick_rnd<FPEMU_FMT_FP16>(rnd, false, dt, false, anyoff);
    case FPEMU_FMT_E4M3_IEEE: return daz ? nullptr : pick_rnd<FPEMU_FMT_E4M3_IEEE>(rnd, false, dt, scaled, anyoff);
    case FPEMU_FMT_FP4:  return (rnd == FPEMU_RND_RNE && !daz) ? pick_dt<FPEMU_FMT_FP4, FPEMU_RND_RNE, false>(dt, scaled, false) : nullptr;
    default:             return nullptr;   // anything without a packed kernel: generic kernel
  }
}

// cast_shift_kernel instances: the four north_star formats x {RNE, STOCHASTIC, RTZ}; other modes keep the element-wise
// kernel for this layout.
typedef void (*ShiftKernel)(const CastArgs, const int);

template <int FMT, int RND>
static ShiftKernel shift_dt(int dt, bool scaled) {
  constexpr bool kTakesScale = (FMT != FPEMU_FMT_BF16);
  switch (dt) {
    case FPEMU_DT_F32:
      return (kTakesScale && scaled) ? cast_shift_kernel<FMT, RND, FPEMU_DT_F32, kTakesScale> : cast_shift_kernel<FMT, RND, FPEMU_DT_F32, false>;
    case FPEMU_DT_F16:
      return (kTakesScale && scaled) ? cast_shift_kernel<FMT, RND, FPEMU_DT_F16, kTakesScale> : cast_shift_kernel<FMT, RND, FPEMU_DT_F16, false>;
    default:
      return (kTakesScale && scaled) ? cast_shift_kernel<FMT, RND, FPEMU_DT_BF16, kTakesScale> : cast_shift_kernel<FMT, RND, FPEMU_DT_BF16, false>;
  }
}
template <int FMT>
static ShiftKernel shift_rnd(int rnd, int dt, bool scaled) {
  switch (rnd) {
    case FPEMU_RND_RNE:        return shift_dt<FMT, FPEMU_RND_RNE>(dt, scaled);
    case FPEMU_RND_RTZ:        return shift_dt<FMT, FPEMU_RND_RTZ>(dt, scaled);
    case FPEMU_RND_STOCHASTIC: return shift_dt<FMT, FPEMU_RND_STOCHASTIC>(dt, scaled);
    default:                   return nullptr;
  }
}
static ShiftKernel pick_shift(int fmt, int rnd, bool daz, int dt, bool scaled) {
  if (daz) return nullptr;
  switch (fmt) {
    case FPEMU_FMT_E5M2: return shift_rnd<FPEMU_FMT_E5M2>(rnd, dt, scaled);
    case FPEMU_FMT_E4M3: return shift_rnd<FPEMU_FMT_E4M3>(rnd, dt, scaled);
    case FPEMU_FMT_E3M4: return shift_rnd<FPEMU_FMT_E3M4>(rnd, dt, scaled);
    case FPEMU_FMT_BF16: return shift_rnd<FPEMU_FMT_BF16>(rnd, dt, false);
    default:             return nullptr;
  }
}

int validate_cast(int64_t n, int dtype, int format, int rounding, float scale, const void* in, const void* out) {
  if (n < 0) return FPEMU_ERR_ARG;
  if (dtype < 0 || dtype >= FPEMU_DT_COUNT) return FPEMU_ERR_DTYPE;
  if (format < 0 || format >= FPEMU_FMT_COUNT || rounding < 0 || rounding >= FPEMU_RND_COUNT) return FPEMU_ERR_MODE;
  if (format == FPEMU_FMT_FP16 && rounding != FPEMU_RND_RNE && rounding != FPEMU_RND_STOCHASTIC) return FPEMU_ERR_MODE;
  if (scale > 65504.0f) return FPEMU_ERR_SCALE;                 // `if (scale > 65504)`, :1037: a NaN scale passes, as there
  if (n > 0 && (in == nullptr || out == nullptr)) return FPEMU_ERR_ARG;
  const uintptr_t esz = (uintptr_t)dtype_bytes(dtype);
  if (((uintptr_t)in % esz) || ((uintptr_t)out % esz)) return FPEMU_ERR_ALIGN;
  return FPEMU_OK;
}

int launch_cast(const CastArgs& a, int dtype, cudaStream_t stream) {
  if (a.n == 0) return FPEMU_OK;
  const bool scaled = (a.d_scale != nullptr) || (a.d_amax != nullptr) || (a.scale != 1.0f);
  const bool daz = (a.flags & FPEMU_FLAG_DAZ) != 0;
  const bool aligned = (((uintptr_t)a.in | (uintptr_t)a.out) & 31u) == 0;
  // the packed kernels gather their random fields per element index (any elem_offset); only the element-wise rounding
  // inside cast_vec_generic_kernel still assumes that a vector does not straddle Philox blocks
  const bool rng_ok = (a.rnd != FPEMU_RND_STOCHASTIC) || ((a.elem_offset & 15u) == 0);
  CastKernel k = aligned ? pick_kernel(a.fmt, a.rnd, daz, dtype, scaled, !rng_ok) : nullptr;
  const int sms = device_info().sm_count;
  // A view that starts off a 32-byte boundary (a slice of a flat parameter buffer, x[1:]), input and output off by the
  // same amount: the few elements up to the boundary go through the element-wise kernel, the rest is an aligned
  // problem for the vector kernels.
  const uintptr_t mis = (uintptr_t)a.in & 31u;
  if (!aligned && mis == ((uintptr_t)a.out & 31u)) {
    const int esz = dtype_bytes(dtype);
    const int64_t head = (int64_t)((32u - mis) / (uintptr_t)esz);
    if (head < a.n && pick_kernel(a.fmt, a.rnd, daz, dtype, scaled, false) != nullptr) {
      CastArgs h = a;
      h.n = head;
      cast_generic_kernel<<<1, 32, 0, stream>>>(h, dtype);
      CastArgs b = a;
      b.in = static_cast<const char*>(a.in) + head * esz;
      b.out = static_cast<char*>(a.out) + head * esz;
      b.n = a.n - head;
      b.elem_offset = a.elem_offset + (uint64_t)head;
      return launch_cast(b, dtype, stream);
    }
  }
  // Input and output off a 32-byte boundary by different amounts: element-wise head up to the OUTPUT's boundary, then
  // cast_shift_kernel (aligned 256-bit stores, input assembled from two aligned blocks).
  if (!aligned && mis != ((uintptr_t)a.out & 31u)) {
    ShiftKernel sk = pick_shift(a.fmt, a.rnd, daz, dtype, scaled);
    const int esz = dtype_bytes(dtype);
    const uintptr_t mis_out = (uintptr_t)a.out & 31u;
    const int64_t head = mis_out ? (int64_t)((32u - mis_out) / (uintptr_t)esz) : 0;
    if (sk != nullptr && head < a.n) {
      if (head > 0) {
        CastArgs h = a;
        h.n = head;
        cast_generic_kernel<<<1, 32, 0, stream>>>(h, dtype);
      }
      CastArgs b = a;
      b.in = static_cast<const char*>(a.in) + head * esz;
      b.out = static_cast<char*>(a.out) + head * esz;
      b.n = a.n - head;
      b.elem_offset = a.elem_offset + (uint64_t)head;
      const int d = (int)((uintptr_t)b.in & 31u);                // != 0: the two buffers are misaligned differently
      const int ve = dtype == FPEMU_DT_F32 ? 8 : 16;
      int64_t want = (b.n / ve + kThreads * 2 - 1) / (kThreads * 2);
      if (want < 1) want = 1;
      if (want > 0x7FFFFFFFLL) return FPEMU_ERR_TOO_LARGE;
      sk<<<(unsigned)want, kThreads, 0, stream>>>(b, d);
      return (int)cudaGetLastError();
    }
  }
  if (k != nullptr) {
    const int ve = dtype == FPEMU_DT_F32 ? 8 : 16;
    const int u = dtype == FPEMU_DT_F32 ? UnrollFor<FPEMU_DT_F32>::value : UnrollFor<FPEMU_DT_F16>::value;
    const int64_t nvec = a.n / ve;
    const int64_t per_cta = (int64_t)kThreads * u;
    int64_t want = (nvec + per_cta - 1) / per_cta;
    if (want < 1) want = 1;                                  // a launch with only the ragged tail
    if (want > 0x7FFFFFFFLL) return FPEMU_ERR_TOO_LARGE;     // > 2^43 fp32 elements
    k<<<(unsigned)want, kThreads, 0, stream>>>(a);
  } else if (aligned && rng_ok) {
    const int ve = dtype == FPEMU_DT_F32 ? 8 : 16;
    int64_t want = (a.n / ve + kThreads * 2 - 1) / (kThreads * 2);
    if (want < 1) want = 1;
    if (want > 0x7FFFFFFFLL) return FPEMU_ERR_TOO_LARGE;
    CastKernel g = pick_generic(a.fmt, a.rnd, dtype);
    g<<<(unsigned)want, kThreads, 0, stream>>>(a);
  } else {
    int64_t want = (a.n + 255) / 256;
    const int64_t cap = (int64_t)sms * 8;
    const int grid = (int)(want < cap ? want : cap);
    cast_generic_kernel<<<grid, 256, 0, stream>>>(a, dtype);
  }
  return (int)cudaGetLastError();
}

}  // namespace fpemu

extern "C" int fpemu_cast(const void* in, void* out, int64_t n, int dtype, int format, int rounding, int flags,
                          float scale, const float* d_scale, uint64_t seed, uint64_t elem_offset, void* stream) {
  int st = fpemu::validate_cast(n, dtype, format, rounding, d_scale ? 1.0f : scale, in, out);
  if (st != FPEMU_OK) return st;
  fpemu::CastArgs a;
  a.in = in; a.out = out; a.n = n;
  a.scale = scale; a.d_scale = d_scale;
  a.seed_lo = (uint32_t)seed; a.seed_hi = (uint32_t)(seed >> 32);
  a.elem_offset = elem_offset;
  a.fmt = format; a.rnd = rounding; a.flags = flags;
  a.d_amax = nullptr; a.flt_max = 0.0f; a.d_scale_out = nullptr; a.d_seed = fpemu::seed_source();
  if (format == FPEMU_FMT_BF16 || format == FPEMU_FMT_FP16) { a.scale = 1.0f; a.d_scale = nullptr; }  // no scale in these kernels
  return fpemu::launch_cast(a, dtype, static_cast<cudaStream_t>(stream));
}
