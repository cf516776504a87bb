// fpemu_codec.cu — the 8-bit wire form of the E5M2 / E4M3 / E3M4 cast (C ABI: fpemu_encode, fpemu_decode).
//
// fpemu_encode runs exactly the rounding of fpemu_cast (same kernels' per-pair code, same random fields) and stores
// the format's one-byte code (reference format table README.md:17-27) instead of the de-quantised value;
// fpemu_decode turns codes back into values and applies the cast's un-scale.  For every input whose rounded value
// is a member of the format,  decode(encode(x)) == cast(x)  bit for bit; E5M2 always is, E4M3 / E3M4 values the
// reference leaves between two subnormal grid points (its pseudo-subnormals) go to the nearest code.
//
// Data movement per element: encode reads 4 B (fp32) or 2 B and writes 1 B; decode the reverse — 5 B / 3 B of
// algorithmic traffic against 8 B / 4 B for the value-to-value cast.  256-bit accesses on the wide side, 64- or
// 128-bit on the code side, warp-contiguous, one pass per thread.
#include "fpemu_codec_core.cuh"

namespace fpemu {

template <int FMT, int RND, bool DAZ, int DT, bool SCALED>
__global__ void __launch_bounds__(kThreads) encode_vec_kernel(const CastArgs a_) {
  const CastArgs a = with_device_seed(a_);
  const ScaleCtx sc = make_scale_ctx<DT, SCALED>(a);
  encode_cta<FMT, RND, DAZ, DT, SCALED, CodecUnroll<DT>::value>(a, sc, (int64_t)blockIdx.x, blockIdx.x == gridDim.x - 1);
}

// Codes of one output vector -> values, un-scaled like the cast.
template <int FMT, int DT, bool SCALED>
__device__ __forceinline__ void decode_vec(const uint32_t (&w)[4], const ScaleCtx& sc, Vec32& v) {
  constexpr int kPairs = DtInfo<DT>::kVecElems / 2;
#pragma unroll
  for (int k = 0; k < kPairs; ++k) {
    const uint32_t c2 = __byte_perm(w[k >> 1], 0u, (k & 1) ? 0x4342u : 0x4140u);
    finish_pair<DT, SCALED>(v, k, decode_pair<FMT>(c2), sc);
  }
}

template <int FMT, int DT, bool SCALED>
__global__ void __launch_bounds__(kThreads) decode_vec_kernel(const CastArgs a) {
  constexpr int VE = DtInfo<DT>::kVecElems;
  constexpr int U = CodecUnroll<DT>::value;
  const ScaleCtx sc = make_scale_ctx<DT, SCALED>(a);
  const int64_t nvec = a.n / VE;
  const char* in = static_cast<const char*>(a.in);
  char* out = static_cast<char*>(a.out);
  const int64_t base = (int64_t)blockIdx.x * (kThreads * U) + threadIdx.x;
  uint32_t w[U][4];
#pragma unroll
  for (int j = 0; j < U; ++j) {
    const int64_t vi = base + j * kThreads;
    if (vi < nvec) ldg_codes(in + vi * VE, w[j], VE / 4);
  }
#pragma unroll
  for (int j = 0; j < U; ++j) {
    const int64_t vi = base + j * kThreads;
    if (vi < nvec) {
      Vec32 v;
      decode_vec<FMT, DT, SCALED>(w[j], sc, v);
      stg256(out + vi * 32, v);
    }
  }
  const int tail = (int)(a.n - nvec * VE);
  if (tail > 0 && blockIdx.x == gridDim.x - 1 && threadIdx.x == 0) {
    const int64_t e0 = nvec * VE;
    const unsigned char* c = reinterpret_cast<const unsigned char*>(a.in) + e0;
    uint32_t t[4] = {0u, 0u, 0u, 0u};
    for (int i = 0; i < tail; ++i) t[i >> 2] |= (uint32_t)c[i] << ((i & 3) * 8);
    Vec32 v;
    decode_vec<FMT, DT, SCALED>(t, sc, v);
    if (DT == FPEMU_DT_F32) {
      uint32_t* p = reinterpret_cast<uint32_t*>(a.out) + e0;
      for (int i = 0; i < tail; ++i) p[i] = v.w[i];
    } else {
      unsigned short* p = reinterpret_cast<unsigned short*>(a.out) + e0;
      for (int i = 0; i < tail; ++i) p[i] = (unsigned short)(v.w[i >> 1] >> ((i & 1) * 16));
    }
  }
}

// Any alignment, any rounding mode of the three formats: one element per thread-iteration.
__global__ void __launch_bounds__(256) encode_generic_kernel(const CastArgs a_, const int dt) {
  const CastArgs a = with_device_seed(a_);
  const bool scaled = (a.d_scale != nullptr) || (a.scale != 1.0f);
  float s = 1.0f;
  if (scaled) s = effective_scale(a.d_scale ? __ldg(a.d_scale) : a.scale, dt);
  const bool daz = (a.flags & FPEMU_FLAG_DAZ) != 0;
  unsigned char* out = static_cast<unsigned char*>(a.out);
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < a.n; i += stride) {
    const float x = load_elem(a.in, i, dt);
    uint32_t rb = 0u;
    if (a.rnd == FPEMU_RND_STOCHASTIC) rb = rand_bits_one(a.seed_lo, a.seed_hi, a.elem_offset + (uint64_t)i, false);
    const uint32_t h = __half_as_ushort(__float2half_rn(scaled ? x * s : x));
    out[i] = (unsigned char)encode_one(round_one_h16(h, a.fmt, a.rnd, daz, rb), a.fmt);
  }
}

__global__ void __launch_bounds__(256) decode_generic_kernel(const CastArgs a, const int dt) {
  const bool scaled = (a.d_scale != nullptr) || (a.scale != 1.0f);
  float rs = 1.0f;
  if (scaled) rs = __frcp_rn(effective_scale(a.d_scale ? __ldg(a.d_scale) : a.scale, dt));
  const unsigned char* in = static_cast<const unsigned char*>(a.in);
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < a.n; i += stride) {
    float f = __half2float(__ushort_as_half((unsigned short)decode_one(in[i], a.fmt)));
    if (scaled) {
      f = f * rs;
      if (!(a.flags & FPEMU_FLAG_UNSCALE_FP32) || dt == FPEMU_DT_F16) f = __half2float(__float2half_rn(f));
    }
    store_elem(a.out, i, dt, f);
  }
}

// ------------------------------------------------------------------------------------------ dispatch
typedef void (*CodecKernel)(const CastArgs);

template <int FMT, int RND, bool DAZ>
static CodecKernel enc_dt(int dt, bool scaled) {
  switch (dt) {
    case FPEMU_DT_F32:
      return scaled ? encode_vec_kernel<FMT, RND, DAZ, FPEMU_DT_F32, true> : encode_vec_kernel<FMT, RND, DAZ, FPEMU_DT_F32, false>;
    case FPEMU_DT_F16:
      return scaled ? encode_vec_kernel<FMT, RND, DAZ, FPEMU_DT_F16, true> : encode_vec_kernel<FMT, RND, DAZ, FPEMU_DT_F16, false>;
    default:
      return scaled ? encode_vec_kernel<FMT, RND, DAZ, FPEMU_DT_BF16, true> : encode_vec_kernel<FMT, RND, DAZ, FPEMU_DT_BF16, false>;
  }
}

template <int FMT>
static CodecKernel enc_rnd(int rnd, bool daz, int dt, bool scaled) {
  if (FMT == FPEMU_FMT_E5M2 && daz) {
    if (rnd == FPEMU_RND_RNE) return enc_dt<FPEMU_FMT_E5M2, FPEMU_RND_RNE, true>(dt, scaled);
    if (rnd == FPEMU_RND_STOCHASTIC) return enc_dt<FPEMU_FMT_E5M2, FPEMU_RND_STOCHASTIC, true>(dt, scaled);
    return nullptr;
  }
  switch (rnd) {
    case FPEMU_RND_RNE:        return enc_dt<FMT, FPEMU_RND_RNE, false>(dt, scaled);
    case FPEMU_RND_RTZ:        return enc_dt<FMT, FPEMU_RND_RTZ, false>(dt, scaled);
    case FPEMU_RND_STOCHASTIC: return enc_dt<FMT, FPEMU_RND_STOCHASTIC, false>(dt, scaled);
    default:                   return nullptr;
  }
}

template <int FMT>
static CodecKernel dec_dt(int dt, bool scaled) {
  switch (dt) {
    case FPEMU_DT_F32:  return scaled ? decode_vec_kernel<FMT, FPEMU_DT_F32, true> : decode_vec_kernel<FMT, FPEMU_DT_F32, false>;
    case FPEMU_DT_F16:  return scaled ? decode_vec_kernel<FMT, FPEMU_DT_F16, true> : decode_vec_kernel<FMT, FPEMU_DT_F16, false>;
    default:            return scaled ? decode_vec_kernel<FMT, FPEMU_DT_BF16, true> : decode_vec_kernel<FMT, FPEMU_DT_BF16, false>;
  }
}

static int validate_codec(const void* wide, const void* codes, int64_t n, int dtype, int format, int rounding, float scale) {
  if (n < 0) return FPEMU_ERR_ARG;
  if (dtype < 0 || dtype >= FPEMU_DT_COUNT) return FPEMU_ERR_DTYPE;
  if (format != FPEMU_FMT_E5M2 && format != FPEMU_FMT_E4M3 && format != FPEMU_FMT_E3M4) return FPEMU_ERR_MODE;
  if (rounding < 0 || rounding >= FPEMU_RND_COUNT) return FPEMU_ERR_MODE;
  if (scale > 65504.0f) return FPEMU_ERR_SCALE;
  if (n > 0 && (wide == nullptr || codes == nullptr)) return FPEMU_ERR_ARG;
  if ((uintptr_t)wide % (uintptr_t)dtype_bytes(dtype)) return FPEMU_ERR_ALIGN;
  return FPEMU_OK;
}

static int launch_codec(CodecKernel k, void (*generic)(const CastArgs, int), const CastArgs& a, int dtype, cudaStream_t stream) {
  if (a.n == 0) return FPEMU_OK;
  if (k != nullptr) {
    const int ve = dtype == FPEMU_DT_F32 ? 8 : 16;
    const int64_t per_cta = (int64_t)kThreads * (dtype == FPEMU_DT_F32 ? CodecUnroll<FPEMU_DT_F32>::value : CodecUnroll<FPEMU_DT_F16>::value);
    int64_t want = (a.n / ve + per_cta - 1) / per_cta;
    if (want < 1) want = 1;
    if (want > 0x7FFFFFFFLL) return FPEMU_ERR_TOO_LARGE;
    k<<<(unsigned)want, kThreads, 0, stream>>>(a);
  } else {
    int64_t want = (a.n + 255) / 256;
    const int64_t cap = (int64_t)device_info().sm_count * 8;
    generic<<<(int)(want < cap ? want : cap), 256, 0, stream>>>(a, dtype);
  }
  return (int)cudaGetLastError();
}

static CastArgs codec_args(const void* in, void* out, int64_t n, int format, int rounding, int flags, float scale,
                           const float* d_scale, uint64_t seed, uint64_t elem_offset) {
  CastArgs a;
  a.in = in; a.out = out; a.n = n;
  a.scale = scale; a.d_scale = d_scale;
  a.seed_lo = (uint32_t)seed; a.seed_hi = (uint32_t)(seed >> 32);
  a.elem_offset = elem_offset;
  a.fmt = format; a.rnd = rounding; a.flags = flags;
  a.d_amax = nullptr; a.flt_max = 0.0f; a.d_scale_out = nullptr; a.d_seed = seed_source();
  return a;
}

}  // namespace fpemu

extern "C" int fpemu_encode(const void* in, void* codes, int64_t n, int dtype, int format, int rounding, int flags,
                            float scale, const float* d_scale, uint64_t seed, uint64_t elem_offset, void* stream) {
  using namespace fpemu;
  int st = validate_codec(in, codes, n, dtype, format, rounding, d_scale ? 1.0f : scale);
  if (st != FPEMU_OK) return st;
  const CastArgs a = codec_args(in, codes, n, format, rounding, flags, scale, d_scale, seed, elem_offset);
  const bool scaled = (d_scale != nullptr) || (scale != 1.0f);
  const bool daz = (flags & FPEMU_FLAG_DAZ) != 0;
  const uintptr_t code_align = dtype == FPEMU_DT_F32 ? 8u : 16u;
  const bool aligned = ((uintptr_t)in % 32u) == 0 && ((uintptr_t)codes % code_align) == 0;
  const bool rng_ok = (rounding != FPEMU_RND_STOCHASTIC) || ((elem_offset & 15u) == 0);
  CodecKernel k = nullptr;
  if (aligned && rng_ok) {
    if (format == FPEMU_FMT_E5M2) k = enc_rnd<FPEMU_FMT_E5M2>(rounding, daz, dtype, scaled);
    else if (format == FPEMU_FMT_E4M3) k = enc_rnd<FPEMU_FMT_E4M3>(rounding, false, dtype, scaled);
    else k = enc_rnd<FPEMU_FMT_E3M4>(rounding, false, dtype, scaled);
  }
  return launch_codec(k, encode_generic_kernel, a, dtype, static_cast<cudaStream_t>(stream));
}

extern "C" int fpemu_decode(const void* codes, void* out, int64_t n, int dtype, int format, int flags, float scale,
                            const float* d_scale, void* stream) {
  using namespace fpemu;
  int st = validate_codec(out, codes, n, dtype, format, FPEMU_RND_RNE, d_scale ? 1.0f : scale);
  if (st != FPEMU_OK) return st;
  const CastArgs a = codec_args(codes, out, n, format, FPEMU_RND_RNE, flags, scale, d_scale, 0, 0);
  const bool scaled = (d_scale != nullptr) || (scale != 1.0f);
  const uintptr_t code_align = dtype == FPEMU_DT_F32 ? 8u : 16u;
  const bool aligned = ((uintptr_t)out % 32u) == 0 && ((uintptr_t)codes % code_align) == 0;
  CodecKernel k = nullptr;
  if (aligned) {
    if (format == FPEMU_FMT_E5M2) k = dec_dt<FPEMU_FMT_E5M2>(dtype, scaled);
    else if (format == FPEMU_FMT_E4M3) k = dec_dt<FPEMU_FMT_E4M3>(dtype, scaled);
    else k = dec_dt<FPEMU_FMT_E3M4>(dtype, scaled);
  }
  return launch_codec(k, decode_generic_kernel, a, dtype, static_cast<cudaStream_t>(stream));
}
