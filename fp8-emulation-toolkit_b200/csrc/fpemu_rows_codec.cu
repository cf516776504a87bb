// fpemu_rows_codec.cu — per-channel 8-bit wire form (C ABI: fpemu_encode_rowwise, fpemu_decode_rowwise).
//
// The per-channel branch of quantize_tensor (mpemu/qutils.py:289-311) gives every row of a weight matrix its own
// scale = clamp(flt_max / max|row|).  fpemu_cast_rowwise produces the de-quantised values; this file produces what
// would be stored or sent instead: one code per element plus one float per row, and turns that pair back into
// exactly the values fpemu_cast_rowwise returns (for every element whose rounded value is a member of the format,
// see fpemu_codec.cu).
//
// Encode: same shape as the register-resident row kernel (fpemu_group.cu) — a warp or a CTA loads its row once,
// reduces |x|, derives the scale and emits codes from the registers: read 1 element + write 1 byte, single pass.
// Rows that do not fit (more than 16 384 fp32 / 32 768 16-bit elements), unaligned buffers or columns, and modes
// without a packed kernel use one warp per row, two passes, element-wise.
// Decode: one pass over the codes; each 32-byte output vector looks up its row's scale.
#include "fpemu_codec_core.cuh"
#include "fpemu_group_core.cuh"

namespace fpemu {

template <int FMT, int RND, int DT, int R>
__global__ void __launch_bounds__(256) encode_rows_kernel(const CastArgs a_, const int64_t rows, const int vpr, const int lg,
                                                         const bool clamp, const float flt_max, float* d_row_scale) {
  const CastArgs a = with_device_seed(a_);
  constexpr int VE = DtInfo<DT>::kVecElems;
  const int G = 1 << lg;
  const int tg = threadIdx.x & (G - 1);
  const int64_t row = (int64_t)blockIdx.x * (256 >> lg) + (threadIdx.x >> lg);
  const bool live = row < rows;
  const int64_t v0 = row * vpr;
  const char* in = static_cast<const char*>(a.in);
  char* out = static_cast<char*>(a.out);
  Vec32 v[R];
  float m = 0.0f;
#pragma unroll
  for (int j = 0; j < R; ++j) {
    const int vi = j * G + tg;
    if (live && vi < vpr) v[j] = ldg256(in + (v0 + vi) * 32);
    else {
#pragma unroll
      for (int i = 0; i < 8; ++i) v[j].w[i] = 0u;
    }
  }
#pragma unroll
  for (int j = 0; j < R; ++j) m = absmax_nan(m, vec_absmax<DT, true>(v[j]));
  const float s_host = scale_from_amax(group_absmax(m, lg), flt_max, clamp, DT);
  if (live && tg == 0) d_row_scale[row] = s_host;
  ScaleCtx sc;
  sc.scale = effective_scale(s_host, DT);
  sc.rscale = __frcp_rn(sc.scale);
  sc.unscale32 = false;
#pragma unroll
  for (int j = 0; j < R; ++j) {
    const int vi = j * G + tg;
    if (live && vi < vpr) {
      uint32_t rw[4], w[4];
      vec_rand_bytes<FMT, RND, DT>(a, a.elem_offset + (uint64_t)(v0 + vi) * VE, rw);
      encode_vec<FMT, RND, false, DT, true>(v[j], sc, rw, w, true);      // scaled to the row maximum: hardware path
      stg_codes(out + (v0 + vi) * VE, w, VE / 4);
    }
  }
}

// One warp per row, two passes, any alignment, any rounding mode of the three formats.
__global__ void __launch_bounds__(256) encode_rows_generic_kernel(const CastArgs a_, const int64_t rows, const int64_t cols,
                                                                 const int dt, const bool clamp, const float flt_max,
                                                                 float* d_row_scale) {
  const CastArgs a = with_device_seed(a_);
  const int lane = threadIdx.x & 31;
  const int64_t warp0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  const bool daz = (a.flags & FPEMU_FLAG_DAZ) != 0;
  unsigned char* out = static_cast<unsigned char*>(a.out);
  for (int64_t r = warp0; r < rows; r += nwarps) {
    const int64_t e0 = r * cols;
    float m = 0.0f;
    for (int64_t i = lane; i < cols; i += 32) m = absmax_nan(m, load_elem(a.in, e0 + i, dt));
    m = __uint_as_float(warp_max_u32(__float_as_uint(m)));
    const float s_host = scale_from_amax(m, flt_max, clamp, dt);
    if (lane == 0) d_row_scale[r] = s_host;
    const float s = effective_scale(s_host, dt);
    for (int64_t i = lane; i < cols; i += 32) {
      const float x = load_elem(a.in, e0 + i, dt);
      uint32_t rb = 0u;
      if (a.rnd == FPEMU_RND_STOCHASTIC) rb = rand_bits_one(a.seed_lo, a.seed_hi, a.elem_offset + (uint64_t)(e0 + i), false);
      const uint32_t h = __half_as_ushort(__float2half_rn(x * s));
      out[e0 + i] = (unsigned char)encode_one(round_one_h16(h, a.fmt, a.rnd, daz, rb), a.fmt);
    }
  }
}

// Codes -> values with the row's scale; a vector never straddles rows (cols % vector length == 0).
template <int FMT, int DT>
__global__ void __launch_bounds__(kThreads) decode_rows_kernel(const CastArgs a, const int64_t vpr, const float* d_row_scale) {
  constexpr int VE = DtInfo<DT>::kVecElems;
  constexpr int U = CodecUnroll<DT>::value;
  const int64_t nvec = a.n / VE;
  const char* in = static_cast<const char*>(a.in);
  char* out = static_cast<char*>(a.out);
  const int64_t base = (int64_t)blockIdx.x * (kThreads * U) + threadIdx.x;
  uint32_t w[U][4];
  float s[U];
#pragma unroll
  for (int j = 0; j < U; ++j) {
    const int64_t vi = base + j * kThreads;
    s[j] = 1.0f;
    if (vi < nvec) {
      ldg_codes(in + vi * VE, w[j], VE / 4);
      s[j] = __ldg(d_row_scale + vi / vpr);
    }
  }
#pragma unroll
  for (int j = 0; j < U; ++j) {
    const int64_t vi = base + j * kThreads;
    if (vi < nvec) {
      ScaleCtx sc;
      sc.scale = effective_scale(s[j], DT);
      sc.rscale = __frcp_rn(sc.scale);
      sc.unscale32 = (a.flags & FPEMU_FLAG_UNSCALE_FP32) != 0;
      Vec32 v;
#pragma unroll
      for (int k = 0; k < VE / 2; ++k) {
        const uint32_t c2 = __byte_perm(w[j][k >> 1], 0u, (k & 1) ? 0x4342u : 0x4140u);
        finish_pair<DT, true>(v, k, decode_pair<FMT>(c2), sc);
      }
      stg256(out + vi * 32, v);
    }
  }
}

__global__ void __launch_bounds__(256) decode_rows_generic_kernel(const CastArgs a, const int64_t cols, const int dt,
                                                                 const float* d_row_scale) {
  const unsigned char* in = static_cast<const unsigned char*>(a.in);
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < a.n; i += stride) {
    const float rs = __frcp_rn(effective_scale(__ldg(d_row_scale + i / cols), dt));
    float f = __half2float(__ushort_as_half((unsigned short)decode_one(in[i], a.fmt))) * rs;
    if (!(a.flags & FPEMU_FLAG_UNSCALE_FP32) || dt == FPEMU_DT_F16) f = __half2float(__float2half_rn(f));
    store_elem(a.out, i, dt, f);
  }
}

// Row-wise cast with scales the caller supplies (one float per row): the value-to-value counterpart of decode_rows_kernel.
// Used where the reference derives a per-channel scale from something other than the maximum (its `mean` scaling
// takes torch.mode of each channel, mpemu/qutils.py:294-301).
template <int FMT, int RND, int DT>
__global__ void __launch_bounds__(kThreads) cast_rows_scaled_kernel(const CastArgs a_, const int64_t vpr, const float* d_row_scale) {
  const CastArgs a = with_device_seed(a_);
  constexpr int VE = DtInfo<DT>::kVecElems;
  constexpr int U = UnrollFor<DT>::value;
  const int64_t nvec = a.n / VE;
  const char* in = static_cast<const char*>(a.in);
  char* out = static_cast<char*>(a.out);
  const int64_t base = (int64_t)blockIdx.x * (kThreads * U) + threadIdx.x;
  Vec32 v[U];
  float s[U];
#pragma unroll
  for (int j = 0; j < U; ++j) {
    const int64_t vi = base + j * kThreads;
    s[j] = 1.0f;
    if (vi < nvec) {
      v[j] = ldg256(in + vi * 32);
      s[j] = __ldg(d_row_scale + vi / vpr);
    }
  }
#pragma unroll
  for (int j = 0; j < U; ++j) {
    const int64_t vi = base + j * kThreads;
    if (vi < nvec) {
      ScaleCtx sc;
      sc.scale = effective_scale(s[j], DT);
      sc.rscale = __frcp_rn(sc.scale);
      sc.unscale32 = (a.flags & FPEMU_FLAG_UNSCALE_FP32) != 0;
      process_vec<FMT, RND, false, DT, true>(v[j], sc, a, a.elem_offset + (uint64_t)vi * VE);
      stg256(out + vi * 32, v[j]);
    }
  }
}

__global__ void __launch_bounds__(256) cast_rows_scaled_generic_kernel(const CastArgs a_, const int64_t cols, const int dt,
                                                                      const float* d_row_scale) {
  const CastArgs a = with_device_seed(a_);
  const bool wide = (a.fmt == FPEMU_FMT_BF16 || a.fmt == FPEMU_FMT_FP16);
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < a.n; i += stride) {
    const float s = effective_scale(__ldg(d_row_scale + i / cols), dt);
    uint32_t rb = 0u;
    if (a.rnd == FPEMU_RND_STOCHASTIC) rb = rand_bits_one(a.seed_lo, a.seed_hi, a.elem_offset + (uint64_t)i, wide);
    store_elem(a.out, i, dt, cast_one(load_elem(a.in, i, dt), a.fmt, a.rnd, a.flags, true, s, __frcp_rn(s), dt == FPEMU_DT_F16, rb));
  }
}

typedef void (*RowsScaledKernel)(const CastArgs, const int64_t, const float*);
template <int FMT, int RND>
static RowsScaledKernel rs_dt(int dt) {
  switch (dt) {
    case FPEMU_DT_F32:  return cast_rows_scaled_kernel<FMT, RND, FPEMU_DT_F32>;
    case FPEMU_DT_F16:  return cast_rows_scaled_kernel<FMT, RND, FPEMU_DT_F16>;
    default:            return cast_rows_scaled_kernel<FMT, RND, FPEMU_DT_BF16>;
  }
}
template <int FMT>
static RowsScaledKernel rs_rnd(int rnd, int dt) {
  switch (rnd) {
    case FPEMU_RND_RNE:        return rs_dt<FMT, FPEMU_RND_RNE>(dt);
    case FPEMU_RND_RTZ:        return rs_dt<FMT, FPEMU_RND_RTZ>(dt);
    case FPEMU_RND_STOCHASTIC: return rs_dt<FMT, FPEMU_RND_STOCHASTIC>(dt);
    default:                   return nullptr;
  }
}

typedef void (*EncRowsKernel)(const CastArgs, const int64_t, const int, const int, const bool, const float, float*);
typedef void (*DecRowsKernel)(const CastArgs, const int64_t, const float*);

template <int FMT, int RND, int DT>
static EncRowsKernel er_shape(int r) {
  if (r == 2) return encode_rows_kernel<FMT, RND, DT, 2>;
  if (r == 3) return encode_rows_kernel<FMT, RND, DT, 3>;
  if (r == 4) return encode_rows_kernel<FMT, RND, DT, 4>;
  if (r == 6) return encode_rows_kernel<FMT, RND, DT, 6>;
  return encode_rows_kernel<FMT, RND, DT, 8>;
}
template <int FMT, int RND>
static EncRowsKernel er_dt(int dt, int shape) {
  switch (dt) {
    case FPEMU_DT_F32:  return er_shape<FMT, RND, FPEMU_DT_F32>(shape);
    case FPEMU_DT_F16:  return er_shape<FMT, RND, FPEMU_DT_F16>(shape);
    default:            return er_shape<FMT, RND, FPEMU_DT_BF16>(shape);
  }
}
template <int FMT>
static EncRowsKernel er_rnd(int rnd, int dt, int shape) {
  switch (rnd) {
    case FPEMU_RND_RNE:        return er_dt<FMT, FPEMU_RND_RNE>(dt, shape);
    case FPEMU_RND_RTZ:        return er_dt<FMT, FPEMU_RND_RTZ>(dt, shape);
    case FPEMU_RND_STOCHASTIC: return er_dt<FMT, FPEMU_RND_STOCHASTIC>(dt, shape);
    default:                   return nullptr;
  }
}
template <int FMT>
static DecRowsKernel dr_dt(int dt) {
  switch (dt) {
    case FPEMU_DT_F32:  return decode_rows_kernel<FMT, FPEMU_DT_F32>;
    case FPEMU_DT_F16:  return decode_rows_kernel<FMT, FPEMU_DT_F16>;
    default:            return decode_rows_kernel<FMT, FPEMU_DT_BF16>;
  }
}

static int validate_rows(const void* wide, const void* codes, const float* d_row_scale, int64_t rows, int64_t cols,
                         int dtype, int format, int rounding) {
  if (rows < 0 || cols < 0) return FPEMU_ERR_ARG;
  if (dtype < 0 || dtype >= FPEMU_DT_COUNT) return FPEMU_ERR_DTYPE;
  if (format != FPEMU_FMT_E5M2 && format != FPEMU_FMT_E4M3 && format != FPEMU_FMT_E3M4) return FPEMU_ERR_MODE;
  if (rounding < 0 || rounding >= FPEMU_RND_COUNT) return FPEMU_ERR_MODE;
  if (rows * cols > 0 && (wide == nullptr || codes == nullptr || d_row_scale == nullptr)) return FPEMU_ERR_ARG;
  if ((uintptr_t)wide % (uintptr_t)dtype_bytes(dtype)) return FPEMU_ERR_ALIGN;
  return FPEMU_OK;
}

static CastArgs rows_args(const void* in, void* out, int64_t n, int format, int rounding, int flags, uint64_t seed,
                          uint64_t elem_offset) {
  CastArgs a;
  a.in = in; a.out = out; a.n = n;
  a.scale = 1.0f; a.d_scale = nullptr;
  a.seed_lo = (uint32_t)seed; a.seed_hi = (uint32_t)(seed >> 32);
  a.elem_offset = elem_offset;
  a.fmt = format; a.rnd = rounding; a.flags = flags;
  a.d_amax = nullptr; a.flt_max = 0.0f; a.d_scale_out = nullptr; a.d_seed = seed_source();
  return a;
}

}  // namespace fpemu

extern "C" int fpemu_encode_rowwise(const void* in, void* codes, int64_t rows, int64_t cols, int dtype, int format,
                                    int rounding, int flags, float flt_max, float* d_row_scale, uint64_t seed,
                                    uint64_t elem_offset, void* stream_) {
  using namespace fpemu;
  int st = validate_rows(in, codes, d_row_scale, rows, cols, dtype, format, rounding);
  if (st != FPEMU_OK) return st;
  const int64_t n = rows * cols;
  if (n == 0) return FPEMU_OK;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  const bool clamp = (flags & FPEMU_FLAG_NO_SCALE_CLAMP) == 0;
  const CastArgs a = rows_args(in, codes, n, format, rounding, flags, seed, elem_offset);
  const int ve = dtype == FPEMU_DT_F32 ? 8 : 16;
  const uintptr_t code_align = dtype == FPEMU_DT_F32 ? 8u : 16u;
  const bool aligned = ((uintptr_t)in % 32u) == 0 && ((uintptr_t)codes % code_align) == 0 && cols % ve == 0;
  const bool rng_ok = (rounding != FPEMU_RND_STOCHASTIC) || ((elem_offset & 15u) == 0 && cols % 16 == 0);
  const int64_t vpr = cols / ve;
  int lg = 0, r = 0;
  if (aligned && rng_ok && !(flags & FPEMU_FLAG_DAZ) && group_shape_tight(vpr, &lg, &r)) {   // fewest idle vector slots
    EncRowsKernel k = nullptr;
    switch (format) {
      case FPEMU_FMT_E5M2: k = er_rnd<FPEMU_FMT_E5M2>(rounding, dtype, r); break;
      case FPEMU_FMT_E4M3: k = er_rnd<FPEMU_FMT_E4M3>(rounding, dtype, r); break;
      default:             k = er_rnd<FPEMU_FMT_E3M4>(rounding, dtype, r); break;
    }
    const int64_t per_cta = 256 >> lg;
    const int64_t grid = (rows + per_cta - 1) / per_cta;
    if (k != nullptr && grid <= 0x7FFFFFFFLL) {
      k<<<(unsigned)grid, 256, 0, stream>>>(a, rows, (int)vpr, lg, clamp, flt_max, d_row_scale);
      return (int)cudaGetLastError();
    }
  }
  int64_t want = (rows + 7) / 8;
  const int64_t cap = (int64_t)device_info().sm_count * 8;
  encode_rows_generic_kernel<<<(int)(want < cap ? want : cap), 256, 0, stream>>>(a, rows, cols, dtype, clamp, flt_max, d_row_scale);
  return (int)cudaGetLastError();
}

extern "C" int fpemu_decode_rowwise(const void* codes, void* out, int64_t rows, int64_t cols, int dtype, int format,
                                    int flags, const float* d_row_scale, void* stream_) {
  using namespace fpemu;
  int st = validate_rows(out, codes, d_row_scale, rows, cols, dtype, format, FPEMU_RND_RNE);
  if (st != FPEMU_OK) return st;
  const int64_t n = rows * cols;
  if (n == 0) return FPEMU_OK;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  const CastArgs a = rows_args(codes, out, n, format, FPEMU_RND_RNE, flags, 0, 0);
  const int ve = dtype == FPEMU_DT_F32 ? 8 : 16;
  const uintptr_t code_align = dtype == FPEMU_DT_F32 ? 8u : 16u;
  if (((uintptr_t)out % 32u) == 0 && ((uintptr_t)codes % code_align) == 0 && cols % ve == 0) {
    DecRowsKernel k = format == FPEMU_FMT_E5M2 ? dr_dt<FPEMU_FMT_E5M2>(dtype)
                    : format == FPEMU_FMT_E4M3 ? dr_dt<FPEMU_FMT_E4M3>(dtype) : dr_dt<FPEMU_FMT_E3M4>(dtype);
    const int u = dtype == FPEMU_DT_F32 ? CodecUnroll<FPEMU_DT_F32>::value : CodecUnroll<FPEMU_DT_F16>::value;
    const int64_t per_cta = (int64_t)kThreads * u;
    const int64_t grid = (n / ve + per_cta - 1) / per_cta;
    if (grid > 0x7FFFFFFFLL) return FPEMU_ERR_TOO_LARGE;
    k<<<(unsigned)grid, kThreads, 0, stream>>>(a, cols / ve, d_row_scale);
    return (int)cudaGetLastError();
  }
  int64_t want = (n + 255) / 256;
  const int64_t cap = (int64_t)device_info().sm_count * 8;
  decode_rows_generic_kernel<<<(int)(want < cap ? want : cap), 256, 0, stream>>>(a, cols, dtype, d_row_scale);
  return (int)cudaGetLastError();
}

extern "C" int fpemu_cast_rows_scaled(const void* in, void* out, int64_t rows, int64_t cols, int dtype, int format,
                                      int rounding, int flags, const float* d_row_scale, uint64_t seed,
                                      uint64_t elem_offset, void* stream_) {
  using namespace fpemu;
  // every format that takes a scale: the three 8-bit formats, and E4M3_IEEE / FP4 (which have no byte code, hence are
  // not in validate_rows' list) — the reference's per-channel loop handles all of them (mpemu/qutils.py:289-311)
  const bool extra = (format == FPEMU_FMT_E4M3_IEEE || format == FPEMU_FMT_FP4);
  int st = validate_rows(in, out, d_row_scale, rows, cols, dtype, extra ? FPEMU_FMT_E4M3 : format, rounding);
  if (st != FPEMU_OK) return st;
  if ((uintptr_t)out % (uintptr_t)dtype_bytes(dtype)) return FPEMU_ERR_ALIGN;
  const int64_t n = rows * cols;
  if (n == 0) return FPEMU_OK;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  const CastArgs a = rows_args(in, out, n, format, rounding, flags, seed, elem_offset);
  const int ve = dtype == FPEMU_DT_F32 ? 8 : 16;
  const bool aligned = (((uintptr_t)in | (uintptr_t)out) & 31u) == 0 && cols % ve == 0;
  const bool rng_ok = (rounding != FPEMU_RND_STOCHASTIC) || ((elem_offset & 15u) == 0);
  RowsScaledKernel k = nullptr;
  if (aligned && rng_ok && !(flags & FPEMU_FLAG_DAZ)) {
    k = format == FPEMU_FMT_E5M2 ? rs_rnd<FPEMU_FMT_E5M2>(rounding, dtype)
      : format == FPEMU_FMT_E4M3 ? rs_rnd<FPEMU_FMT_E4M3>(rounding, dtype)
      : format == FPEMU_FMT_E3M4 ? rs_rnd<FPEMU_FMT_E3M4>(rounding, dtype)
      : format == FPEMU_FMT_E4M3_IEEE ? rs_rnd<FPEMU_FMT_E4M3_IEEE>(rounding, dtype) : nullptr;   // FP4: element-wise
  }
  if (k != nullptr) {
    const int u = dtype == FPEMU_DT_F32 ? UnrollFor<FPEMU_DT_F32>::value : UnrollFor<FPEMU_DT_F16>::value;
    const int64_t per_cta = (int64_t)kThreads * u;
    const int64_t grid = (n / ve + per_cta - 1) / per_cta;
    if (grid > 0x7FFFFFFFLL) return FPEMU_ERR_TOO_LARGE;
    k<<<(unsigned)grid, kThreads, 0, stream>>>(a, cols / ve, d_row_scale);
    return (int)cudaGetLastError();
  }
  int64_t want = (n + 255) / 256;
  const int64_t cap = (int64_t)device_info().sm_count * 8;
  cast_rows_scaled_generic_kernel<<<(int)(want < cap ? want : cap), 256, 0, stream>>>(a, cols, dtype, d_row_scale);
  return (int)cudaGetLastError();
}
