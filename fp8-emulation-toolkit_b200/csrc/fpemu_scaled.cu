// fpemu_scaled.cu — everything that derives a scale from the data before casting:
//   fpemu_amax             single-pass |x| maximum (vector loads -> warp redux -> one atomicMax per CTA)
//   fpemu_scale_from_amax  device-side `flt_max / amax` + clamp of mpemu/qutils.py:283-285
//   fpemu_cast_blocked     per-block power-of-two scale ("block_norm", fpemu_kernels.cu:150-170 + :407-414 ...)
//   fpemu_cast_rowwise     per-channel amax-scaled cast (replaces the Python loop of mpemu/qutils.py:289-311)
#include "fpemu_reduce.cuh"

namespace fpemu {

constexpr int kAmaxThreads = 256;
constexpr int kAmaxUnroll = 4;

template <int DT>
__global__ void __launch_bounds__(kAmaxThreads) amax_kernel(const void* in_, int64_t n, float* d_amax) {
  constexpr int VE = DtInfo<DT>::kVecElems;
  const char* in = static_cast<const char*>(in_);
  const int64_t nvec = n / VE;
  float m = 0.0f;
  {
    // one pass per thread (same launch shape as the cast kernel)
    const int64_t base = (int64_t)blockIdx.x * (kAmaxThreads * kAmaxUnroll) + threadIdx.x;
    Vec32 v[kAmaxUnroll];
#pragma unroll
    for (int j = 0; j < kAmaxUnroll; ++j) {
      const int64_t vi = base + j * kAmaxThreads;
      if (vi < nvec) v[j] = ldg256(in + vi * 32);
      else {
#pragma unroll
        for (int i = 0; i < 8; ++i) v[j].w[i] = 0u;
      }
    }
#pragma unroll
    for (int j = 0; j < kAmaxUnroll; ++j) m = absmax_nan(m, vec_absmax<DT, true>(v[j]));
  }
  // ragged tail
  if (blockIdx.x == 0 && threadIdx.x < (int)(n - nvec * VE))
    m = absmax_nan(m, load_elem(in_, nvec * VE + threadIdx.x, DT));
  uint32_t bits = warp_max_u32(__float_as_uint(m));
  __shared__ uint32_t warp_bits[kAmaxThreads / 32];
  if ((threadIdx.x & 31) == 0) warp_bits[threadIdx.x >> 5] = bits;
  __syncthreads();
  if (threadIdx.x < 32) {
    bits = threadIdx.x < kAmaxThreads / 32 ? warp_bits[threadIdx.x] : 0u;
    bits = warp_max_u32(bits);
    // most CTAs lose against the running maximum: look before paying for the atomic
    if (threadIdx.x == 0 && bits > *reinterpret_cast<volatile unsigned int*>(d_amax))
      atomicMax(reinterpret_cast<unsigned int*>(d_amax), bits);
  }
}

// unaligned buffers: scalar loads
__global__ void __launch_bounds__(256) amax_generic_kernel(const void* in, int64_t n, int dt, float* d_amax) {
  float m = 0.0f;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    m = absmax_nan(m, load_elem(in, i, dt));
  uint32_t bits = warp_max_u32(__float_as_uint(m));
  if ((threadIdx.x & 31) == 0 && bits != 0u) atomicMax(reinterpret_cast<unsigned int*>(d_amax), bits);
}

__global__ void scale_from_amax_kernel(const float* d_amax, int count, float flt_max, int dt, float* d_scale) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < count) d_scale[i] = scale_from_amax(d_amax[i], flt_max, true, dt);
}

// ------------------------------------------------------------------------------------------ mean scaling
// Per-tensor `mean` scaling of quantize_tensor (mpemu/qutils.py:276-281):
//     mean = torch.mean(abs(x));  mean = abs(mean) if abs(mean) > 1e-5 else flt_min
//     scale = flt_min / abs(mean)  (a 0-dim tensor: reciprocal * flt_min);  scale = 1.0 if scale < 1.0 else scale
// Two launches, deterministic: per-CTA partial sums of |x| in double (fixed grid-stride assignment) into scratch,
// then one CTA adds the partials in a fixed tree order and applies the rule.  The sum is exact to well below one
// float ulp of the mean, so the result equals torch's float reduction except where that reduction's own rounding
// decides — which only matters when mean < flt_min (otherwise the scale is 1 exactly).
constexpr int kMeanParts = 1024;

template <int DT>
__global__ void __launch_bounds__(256) abs_sum_partial_kernel(const void* in_, int64_t n, double* parts) {
  constexpr int VE = DtInfo<DT>::kVecElems;
  const char* in = static_cast<const char*>(in_);
  const int64_t nvec = n / VE;
  double acc = 0.0;
  for (int64_t vi = (int64_t)blockIdx.x * 256 + threadIdx.x; vi < nvec; vi += (int64_t)gridDim.x * 256) {
    const Vec32 v = ldg256(in + vi * 32);
    float s = 0.0f;                                            // 8 / 16 addends in float, then into the double
#pragma unroll
    for (int k = 0; k < VE / 2; ++k) {
      const float2 f = get_pair<DT>(v, k);
      s += fabsf(f.x) + fabsf(f.y);
    }
    acc += (double)s;
  }
  if (blockIdx.x == 0)
    for (int64_t i = nvec * VE + threadIdx.x; i < n; i += 256) acc += (double)fabsf(load_elem(in_, i, DT));
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xFFFFFFFFu, acc, o);
  __shared__ double warp_sum[8];
  if ((threadIdx.x & 31) == 0) warp_sum[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int w = 0; w < 8; ++w) t += warp_sum[w];
    parts[blockIdx.x] = t;
  }
}

__global__ void __launch_bounds__(256) abs_sum_generic_kernel(const void* in, int64_t n, int dt, double* parts) {
  double acc = 0.0;
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (int64_t)gridDim.x * 256)
    acc += (double)fabsf(load_elem(in, i, dt));
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xFFFFFFFFu, acc, o);
  __shared__ double warp_sum[8];
  if ((threadIdx.x & 31) == 0) warp_sum[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int w = 0; w < 8; ++w) t += warp_sum[w];
    parts[blockIdx.x] = t;
  }
}

__global__ void __launch_bounds__(kMeanParts) mean_scale_kernel(const double* parts, int nparts, int64_t n, float flt_min,
                                                              float floor_, int dt, float* d_scale) {
  __shared__ double s[kMeanParts];
  s[threadIdx.x] = threadIdx.x < nparts ? parts[threadIdx.x] : 0.0;
  __syncthreads();
  for (int o = kMeanParts / 2; o > 0; o >>= 1) {
    if (threadIdx.x < o) s[threadIdx.x] += s[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    float mean = (float)(s[0] / (double)n);
    if (dt == FPEMU_DT_F16) mean = __half2float(__float2half_rn(mean));           // torch.mean of a half tensor is a half
    float scale = 1.0f;                                       // below the floor `mean` becomes flt_min: flt_min / flt_min
    if (fabsf(mean) > floor_) {
      if (dt == FPEMU_DT_F16) {
        const float r = __half2float(__float2half_rn(__frcp_rn(fabsf(mean))));
        scale = __half2float(__float2half_rn(r * flt_min));
      } else {                                                // float32, and bfloat16 (not dispatched upstream) widened
        scale = __fmul_rn(__frcp_rn(fabsf(mean)), flt_min);
      }
    }
    if (scale < 1.0f) scale = 1.0f;
    *d_scale = scale;
  }
}

// Power-of-two block scale for ONE block spanning the whole tensor (block_size == size: the reference CPU op's reading,
// cpp/fpemu_impl.cpp:778-779), derived on the device from the reduced maximum.
__global__ void block_scale_from_amax_kernel(const float* d_amax, int fmt, int dt, float* d_scale) {
  *d_scale = block_scale(*d_amax, fmt, dt);
}

// ------------------------------------------------------------------------------------------ blocked
// Fast path: a quantization block is `vpb` (power of two, <= 32) consecutive 32-byte vectors, i.e. it lives in
// `vpb` adjacent lanes of one warp: xor-shuffle butterfly, everything stays in registers (8 B/elem fp32).
template <int FMT, int RND, int DT>
__global__ void __launch_bounds__(kThreads) blocked_vec_kernel(const CastArgs a_, const int vpb) {
  const CastArgs a = with_device_seed(a_);
  constexpr int VE = DtInfo<DT>::kVecElems;
  const int64_t nvec = a.n / VE;                            // n % block_size == 0 and block_size % VE == 0
  const char* in = static_cast<const char*>(a.in);
  char* out = static_cast<char*>(a.out);
  ScaleCtx sc;
  sc.unscale32 = (a.flags & FPEMU_FLAG_UNSCALE_FP32) != 0;
  // one vector per thread, one pass (same launch shape as the plain cast); lanes past the end stay in the
  // warp because the shuffles below need all 32 of them
  {
    const int64_t vi = (int64_t)blockIdx.x * kThreads + threadIdx.x;
    const bool live = vi < nvec;
    Vec32 v;
    if (live) v = ldg256(in + vi * 32);
    else {
#pragma unroll
      for (int i = 0; i < 8; ++i) v.w[i] = 0u;
    }
    uint32_t mb = __float_as_uint(vec_absmax<DT, false>(v));   // NaN ignored: absmax_block, :150-170
    for (int d = 1; d < vpb; d <<= 1) mb = max(mb, __shfl_xor_sync(0xFFFFFFFFu, mb, d));
    const float s = block_scale(__uint_as_float(mb), FMT, DT);
    sc.scale = s;
    sc.rscale = __frcp_rn(s);
    if (live) {
      process_vec<FMT, RND, false, DT, true>(v, sc, a, a.elem_offset + (uint64_t)vi * VE);
      stg256(out + vi * 32, v);
    }
  }
}

// General path: one warp per quantization block of any size / alignment / format; two passes over the block
// (the second one is served by L1/L2).
__global__ void __launch_bounds__(256) blocked_generic_kernel(const CastArgs a_, const int64_t block_size, const int dt) {
  const CastArgs a = with_device_seed(a_);
  const int lane = threadIdx.x & 31;
  const int64_t nblocks = a.n / block_size;
  const int64_t warp0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  const bool wide = (a.fmt == FPEMU_FMT_BF16 || a.fmt == FPEMU_FMT_FP16);
  for (int64_t b = warp0; b < nblocks; b += nwarps) {
    const int64_t e0 = b * block_size;
    float m = 0.0f;
    for (int64_t i = lane; i < block_size; i += 32) m = fmaxf(m, fabsf(load_elem(a.in, e0 + i, dt)));
    m = __uint_as_float(warp_max_u32(__float_as_uint(m)));
    const float s = block_scale(m, a.fmt, dt);
    const float rs = __frcp_rn(s);
    for (int64_t i = lane; i < block_size; i += 32) {
      const float x = load_elem(a.in, e0 + i, dt);
      uint32_t rb = 0u;
      if (a.rnd == FPEMU_RND_STOCHASTIC) rb = rand_bits_one(a.seed_lo, a.seed_hi, a.elem_offset + (uint64_t)(e0 + i), wide);
      store_elem(a.out, e0 + i, dt, cast_one(x, a.fmt, a.rnd, a.flags, true, s, rs, dt == FPEMU_DT_F16, rb));
    }
  }
}

// Blocks shorter than a warp: one THREAD per block (a warp per block would leave 31 of 32 lanes idle).  The reference's
// embedding call binds block_size = 1 (mpemu/hybrid_emu.py:119 passes the row length in the `blocknorm` slot), i.e. every
// element gets its own power-of-two scale: that case is fully coalesced here.
__global__ void __launch_bounds__(256) blocked_small_kernel(const CastArgs a_, const int block_size, const int dt) {
  const CastArgs a = with_device_seed(a_);
  const int64_t nblocks = a.n / block_size;
  const bool wide = (a.fmt == FPEMU_FMT_BF16 || a.fmt == FPEMU_FMT_FP16);
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; b < nblocks; b += stride) {
    const int64_t e0 = b * block_size;
    float m = 0.0f;
    for (int i = 0; i < block_size; ++i) m = fmaxf(m, fabsf(load_elem(a.in, e0 + i, dt)));
    const float s = block_scale(m, a.fmt, dt);
    const float rs = __frcp_rn(s);
    for (int i = 0; i < block_size; ++i) {
      const float x = load_elem(a.in, e0 + i, dt);
      uint32_t rb = 0u;
      if (a.rnd == FPEMU_RND_STOCHASTIC) rb = rand_bits_one(a.seed_lo, a.seed_hi, a.elem_offset + (uint64_t)(e0 + i), wide);
      store_elem(a.out, e0 + i, dt, cast_one(x, a.fmt, a.rnd, a.flags, true, s, rs, dt == FPEMU_DT_F16, rb));
    }
  }
}

typedef void (*BlockedKernel)(const CastArgs, const int);

template <int FMT, int RND>
static BlockedKernel pick_blocked_dt(int dt) {
  switch (dt) {
    case FPEMU_DT_F32:  return blocked_vec_kernel<FMT, RND, FPEMU_DT_F32>;
    case FPEMU_DT_F16:  return blocked_vec_kernel<FMT, RND, FPEMU_DT_F16>;
    default:            return blocked_vec_kernel<FMT, RND, FPEMU_DT_BF16>;
  }
}
template <int FMT>
static BlockedKernel pick_blocked_rnd(int rnd, int dt) {
  switch (rnd) {
    case FPEMU_RND_RNE:        return pick_blocked_dt<FMT, FPEMU_RND_RNE>(dt);
    case FPEMU_RND_RTZ:        return pick_blocked_dt<FMT, FPEMU_RND_RTZ>(dt);
    case FPEMU_RND_STOCHASTIC: return pick_blocked_dt<FMT, FPEMU_RND_STOCHASTIC>(dt);
    default:                   return nullptr;
  }
}
static BlockedKernel pick_blocked(int fmt, int rnd, int dt) {
  switch (fmt) {
    case FPEMU_FMT_E5M2: return pick_blocked_rnd<FPEMU_FMT_E5M2>(rnd, dt);
    case FPEMU_FMT_E4M3: return pick_blocked_rnd<FPEMU_FMT_E4M3>(rnd, dt);
    case FPEMU_FMT_E3M4: return pick_blocked_rnd<FPEMU_FMT_E3M4>(rnd, dt);
    default:             return nullptr;
  }
}

// ------------------------------------------------------------------------------------------ row-wise
// One warp per row: pass 1 = |x| max of the row, pass 2 = cast with scale clamp(flt_max / max).  The row is
// re-read from L1/L2 (rows of Linear/Conv weights are a few KB), so DRAM traffic stays 8 B/elem (fp32).
template <int FMT, int RND, int DT>
__global__ void __launch_bounds__(kThreads) rowwise_vec_kernel(const CastArgs a_, const int64_t rows, const int64_t cols,
                                                              const float flt_max, float* d_row_scale, const bool clamp) {
  const CastArgs a = with_device_seed(a_);
  constexpr int VE = DtInfo<DT>::kVecElems;
  const int lane = threadIdx.x & 31;
  const int64_t warp0 = ((int64_t)blockIdx.x * kThreads + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * kThreads) >> 5;
  const int64_t vpr = cols / VE;                            // cols % VE == 0 on this path
  ScaleCtx sc;
  sc.unscale32 = (a.flags & FPEMU_FLAG_UNSCALE_FP32) != 0;
  for (int64_t r = warp0; r < rows; r += nwarps) {
    const char* in = static_cast<const char*>(a.in) + r * cols * DtInfo<DT>::kBytes;
    char* out = static_cast<char*>(a.out) + r * cols * DtInfo<DT>::kBytes;
    float m = 0.0f;
    for (int64_t vi = lane; vi < vpr; vi += 32) {
      const Vec32 v = *reinterpret_cast<const Vec32*>(in + vi * 32);   // default caching: pass 2 hits L1
      m = absmax_nan(m, vec_absmax<DT, true>(v));
    }
    m = __uint_as_float(warp_max_u32(__float_as_uint(m)));
    const float s_host = scale_from_amax(m, flt_max, clamp, DT);
    if (d_row_scale != nullptr && lane == 0) d_row_scale[r] = s_host;
    const float s = effective_scale(s_host, DT);
    sc.scale = s;
    sc.rscale = __frcp_rn(s);
    for (int64_t vi = lane; vi < vpr; vi += 32) {
      Vec32 v = *reinterpret_cast<const Vec32*>(in + vi * 32);
      process_vec<FMT, RND, false, DT, true>(v, sc, a, a.elem_offset + (uint64_t)(r * cols + vi * VE));
      stg256(out + vi * 32, v);
    }
  }
}

__global__ void __launch_bounds__(256) rowwise_generic_kernel(const CastArgs a_, const int64_t rows, const int64_t cols,
                                                             const int dt, const float flt_max, float* d_row_scale,
                                                             const bool clamp) {
  const CastArgs a = with_device_seed(a_);
  const int lane = threadIdx.x & 31;
  const int64_t warp0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  const bool wide = (a.fmt == FPEMU_FMT_BF16 || a.fmt == FPEMU_FMT_FP16);
  for (int64_t r = warp0; r < rows; r += nwarps) {
    const int64_t e0 = r * cols;
    float m = 0.0f;
    for (int64_t i = lane; i < cols; i += 32) m = absmax_nan(m, load_elem(a.in, e0 + i, dt));
    m = __uint_as_float(warp_max_u32(__float_as_uint(m)));
    const float s_host = scale_from_amax(m, flt_max, clamp, dt);
    if (d_row_scale != nullptr && lane == 0) d_row_scale[r] = s_host;
    const float s = effective_scale(s_host, dt);
    const float rs = __frcp_rn(s);
    for (int64_t i = lane; i < cols; i += 32) {
      const float x = load_elem(a.in, e0 + i, dt);
      uint32_t rb = 0u;
      if (a.rnd == FPEMU_RND_STOCHASTIC) rb = rand_bits_one(a.seed_lo, a.seed_hi, a.elem_offset + (uint64_t)(e0 + i), wide);
      store_elem(a.out, e0 + i, dt, cast_one(x, a.fmt, a.rnd, a.flags, true, s, rs, dt == FPEMU_DT_F16, rb));
    }
  }
}

typedef void (*RowwiseKernel)(const CastArgs, const int64_t, const int64_t, const float, float*, const bool);

template <int FMT, int RND>
static RowwiseKernel pick_rowwise_dt(int dt) {
  switch (dt) {
    case FPEMU_DT_F32:  return rowwise_vec_kernel<FMT, RND, FPEMU_DT_F32>;
    case FPEMU_DT_F16:  return rowwise_vec_kernel<FMT, RND, FPEMU_DT_F16>;
    default:            return rowwise_vec_kernel<FMT, RND, FPEMU_DT_BF16>;
  }
}
template <int FMT>
static RowwiseKernel pick_rowwise_rnd(int rnd, int dt) {
  switch (rnd) {
    case FPEMU_RND_RNE:        return pick_rowwise_dt<FMT, FPEMU_RND_RNE>(dt);
    case FPEMU_RND_RTZ:        return pick_rowwise_dt<FMT, FPEMU_RND_RTZ>(dt);
    case FPEMU_RND_STOCHASTIC: return pick_rowwise_dt<FMT, FPEMU_RND_STOCHASTIC>(dt);
    default:                   return nullptr;
  }
}
static RowwiseKernel pick_rowwise(int fmt, int rnd, int dt) {
  switch (fmt) {
    case FPEMU_FMT_E5M2: return pick_rowwise_rnd<FPEMU_FMT_E5M2>(rnd, dt);
    case FPEMU_FMT_E4M3: return pick_rowwise_rnd<FPEMU_FMT_E4M3>(rnd, dt);
    case FPEMU_FMT_E3M4: return pick_rowwise_rnd<FPEMU_FMT_E3M4>(rnd, dt);
    default:             return nullptr;
  }
}

int validate_cast(int64_t n, int dtype, int format, int rounding, float scale, const void* in, const void* out);
int launch_cast(const CastArgs& a, int dtype, cudaStream_t stream);
bool launch_group_reg(const CastArgs& a, int dtype, int64_t groups, int64_t vectors_per_group, int rule, float flt_max,
                      float* d_scale_out, cudaStream_t stream, int* status);

static CastArgs make_args(const void* in, void* out, int64_t n, int format, int rounding, int flags, uint64_t seed,
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

using namespace fpemu;

extern "C" int fpemu_amax(const void* in, int64_t n, int dtype, float* d_amax, void* stream_) {
  if (n < 0 || d_amax == nullptr || (n > 0 && in == nullptr)) return FPEMU_ERR_ARG;
  if (dtype < 0 || dtype >= FPEMU_DT_COUNT) return FPEMU_ERR_DTYPE;
  if ((uintptr_t)in % (uintptr_t)dtype_bytes(dtype)) return FPEMU_ERR_ALIGN;
  if (n == 0) return FPEMU_OK;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (((uintptr_t)in & 31u) == 0) {
    const int ve = dtype == FPEMU_DT_F32 ? 8 : 16;
    const int64_t per_cta = (int64_t)kAmaxThreads * kAmaxUnroll;
    int64_t want = (n / ve + per_cta - 1) / per_cta;
    if (want < 1) want = 1;
    if (want > 0x7FFFFFFFLL) return FPEMU_ERR_TOO_LARGE;
    const unsigned grid = (unsigned)want;
    switch (dtype) {
      case FPEMU_DT_F32:  amax_kernel<FPEMU_DT_F32><<<grid, kAmaxThreads, 0, stream>>>(in, n, d_amax); break;
      case FPEMU_DT_F16:  amax_kernel<FPEMU_DT_F16><<<grid, kAmaxThreads, 0, stream>>>(in, n, d_amax); break;
      default:            amax_kernel<FPEMU_DT_BF16><<<grid, kAmaxThreads, 0, stream>>>(in, n, d_amax); break;
    }
  } else {
    // a view off a 32-byte boundary: the elements up to the boundary go through the scalar kernel, the rest is aligned
    const int esz = dtype_bytes(dtype);
    const int64_t head = (int64_t)((32u - ((uintptr_t)in & 31u)) / (uintptr_t)esz);
    const int64_t first = head < n ? head : n;
    amax_generic_kernel<<<1, 32, 0, stream>>>(in, first, dtype, d_amax);
    if (first < n) return fpemu_amax(static_cast<const char*>(in) + first * esz, n - first, dtype, d_amax, stream_);
  }
  return (int)cudaGetLastError();
}

extern "C" int fpemu_scale_from_amax_dtype(const float* d_amax, int count, float flt_max, int dtype, float* d_scale,
                                           void* stream_) {
  if (count < 0 || (count > 0 && (d_amax == nullptr || d_scale == nullptr))) return FPEMU_ERR_ARG;
  if (dtype < 0 || dtype >= FPEMU_DT_COUNT) return FPEMU_ERR_DTYPE;
  if (count == 0) return FPEMU_OK;
  scale_from_amax_kernel<<<(count + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream_)>>>(d_amax, count, flt_max, dtype,
                                                                                             d_scale);
  return (int)cudaGetLastError();
}

extern "C" int fpemu_scale_from_amax(const float* d_amax, int count, float flt_max, float* d_scale, void* stream_) {
  return fpemu_scale_from_amax_dtype(d_amax, count, flt_max, FPEMU_DT_F32, d_scale, stream_);
}

extern "C" int fpemu_mean_scale_work_bytes(void) { return (int)(kMeanParts * sizeof(double)); }

extern "C" int fpemu_scale_from_abs_mean(const void* in, int64_t n, int dtype, float flt_min, float mean_floor,
                                         void* d_work, float* d_scale, void* stream_) {
  if (n <= 0 || in == nullptr || d_work == nullptr || d_scale == nullptr) return FPEMU_ERR_ARG;
  if (dtype < 0 || dtype >= FPEMU_DT_COUNT) return FPEMU_ERR_DTYPE;
  if (((uintptr_t)in % (uintptr_t)dtype_bytes(dtype)) || ((uintptr_t)d_work % 8u)) return FPEMU_ERR_ALIGN;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  double* parts = static_cast<double*>(d_work);
  const int ve = dtype == FPEMU_DT_F32 ? 8 : 16;
  int64_t want = (n / ve + 256 * 4 - 1) / (256 * 4);
  if (want < 1) want = 1;
  const int grid = (int)(want < kMeanParts ? want : kMeanParts);
  if (((uintptr_t)in & 31u) == 0) {
    switch (dtype) {
      case FPEMU_DT_F32:  abs_sum_partial_kernel<FPEMU_DT_F32><<<grid, 256, 0, stream>>>(in, n, parts); break;
      case FPEMU_DT_F16:  abs_sum_partial_kernel<FPEMU_DT_F16><<<grid, 256, 0, stream>>>(in, n, parts); break;
      default:            abs_sum_partial_kernel<FPEMU_DT_BF16><<<grid, 256, 0, stream>>>(in, n, parts); break;
    }
  } else {
    abs_sum_generic_kernel<<<grid, 256, 0, stream>>>(in, n, dtype, parts);
  }
  cudaError_t err = cudaGetLastError();
  if (err != cudaSuccess) return (int)err;
  mean_scale_kernel<<<1, kMeanParts, 0, stream>>>(parts, grid, n, flt_min, mean_floor, dtype, d_scale);
  return (int)cudaGetLastError();
}

extern "C" int fpemu_block_scale_from_amax(const float* d_amax, int format, int dtype, float* d_scale, void* stream_) {
  if (d_amax == nullptr || d_scale == nullptr) return FPEMU_ERR_ARG;
  if (dtype < 0 || dtype >= FPEMU_DT_COUNT) return FPEMU_ERR_DTYPE;
  if (format != FPEMU_FMT_E5M2 && format != FPEMU_FMT_E4M3 && format != FPEMU_FMT_E3M4 && format != FPEMU_FMT_E4M3_IEEE &&
      format != FPEMU_FMT_FP4)
    return FPEMU_ERR_MODE;
  block_scale_from_amax_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream_)>>>(d_amax, format, dtype, d_scale);
  return (int)cudaGetLastError();
}

extern "C" int fpemu_cast_blocked(const void* in, void* out, int64_t n, int64_t block_size, int dtype, int format,
                                  int rounding, int flags, uint64_t seed, uint64_t elem_offset, void* stream_) {
  int st = validate_cast(n, dtype, format, rounding, 1.0f, in, out);
  if (st != FPEMU_OK) return st;
  // BFLOAT16 / FLOAT16 / E5M2_DAZ kernels of the reference take no block_norm argument
  if (format == FPEMU_FMT_BF16 || format == FPEMU_FMT_FP16 || (flags & FPEMU_FLAG_DAZ)) return FPEMU_ERR_MODE;
  if (block_size <= 0 || (n % block_size) != 0) return FPEMU_ERR_ARG;
  if (n == 0) return FPEMU_OK;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  const CastArgs a = make_args(in, out, n, format, rounding, flags, seed, elem_offset);
  const int sms = device_info().sm_count;
  const int ve = dtype == FPEMU_DT_F32 ? 8 : 16;
  const bool aligned = (((uintptr_t)in | (uintptr_t)out) & 31u) == 0;
  const bool rng_ok = (rounding != FPEMU_RND_STOCHASTIC) || ((elem_offset & 15u) == 0);
  const int64_t vpb = block_size / ve;
  // blocks that do not fit the in-warp butterfly: one warp / one CTA per block, block held in registers
  if (aligned && rng_ok && block_size % ve == 0 && !(vpb <= 32 && (vpb & (vpb - 1)) == 0)) {
    int status = 0;
    if (launch_group_reg(a, dtype, n / block_size, vpb, 2, 0.0f, nullptr, stream, &status)) return status;
  }
  BlockedKernel k = nullptr;
  if (aligned && rng_ok && block_size % ve == 0 && vpb <= 32 && (vpb & (vpb - 1)) == 0)
    k = pick_blocked(format, rounding, dtype);
  if (k != nullptr) {
    const int64_t want = (n / ve + kThreads - 1) / kThreads;
    if (want > 0x7FFFFFFFLL) return FPEMU_ERR_TOO_LARGE;
    k<<<(unsigned)want, kThreads, 0, stream>>>(a, (int)vpb);
  } else if (block_size < 32) {
    int64_t want = (n / block_size + 255) / 256;             // one thread per block
    const int64_t cap = (int64_t)sms * 16;
    blocked_small_kernel<<<(int)(want < cap ? want : cap), 256, 0, stream>>>(a, (int)block_size, dtype);
  } else {
    int64_t want = (n / block_size + 7) / 8;                 // 8 warps per CTA, one warp per block
    const int64_t cap = (int64_t)sms * 8;
    blocked_generic_kernel<<<(int)(want < cap ? want : cap), 256, 0, stream>>>(a, block_size, dtype);
  }
  return (int)cudaGetLastError();
}

extern "C" int fpemu_cast_rowwise(const void* in, void* out, int64_t rows, int64_t cols, int dtype, int format,
                                  int rounding, int flags, float flt_max, float* d_row_scale, uint64_t seed,
                                  uint64_t elem_offset, void* stream_) {
  if (rows < 0 || cols < 0) return FPEMU_ERR_ARG;
  const int64_t n = rows * cols;
  int st = validate_cast(n, dtype, format, rounding, 1.0f, in, out);
  if (st != FPEMU_OK) return st;
  if (format == FPEMU_FMT_BF16 || format == FPEMU_FMT_FP16) return FPEMU_ERR_MODE;   // these take no scale
  if (n == 0) return FPEMU_OK;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  const bool clamp = (flags & FPEMU_FLAG_NO_SCALE_CLAMP) == 0;
  const CastArgs a = make_args(in, out, n, format, rounding, flags, seed, elem_offset);
  const int sms = device_info().sm_count;
  const int ve = dtype == FPEMU_DT_F32 ? 8 : 16;
  const bool aligned = (((uintptr_t)in | (uintptr_t)out) & 31u) == 0 && cols % ve == 0;
  const bool rng_ok = (rounding != FPEMU_RND_STOCHASTIC) || ((elem_offset & 15u) == 0 && cols % 16 == 0);
  if (aligned && rng_ok && (flags & FPEMU_FLAG_DAZ) == 0) {
    int status = 0;   // rows that fit in registers: single pass
    if (launch_group_reg(a, dtype, rows, cols / ve, clamp ? 0 : 1, flt_max, d_row_scale, stream, &status)) return status;
  }
  RowwiseKernel k = (aligned && rng_ok && (flags & FPEMU_FLAG_DAZ) == 0) ? pick_rowwise(format, rounding, dtype) : nullptr;
  int64_t want = (rows + 7) / 8;                             // 8 warps per CTA, one warp per row
  const int64_t cap = (int64_t)sms * 8;
  const int grid = (int)(want < cap ? want : cap);
  if (k != nullptr) k<<<grid, kThreads, 0, stream>>>(a, rows, cols, flt_max, d_row_scale, clamp);
  else rowwise_generic_kernel<<<grid, 256, 0, stream>>>(a, rows, cols, dtype, flt_max, d_row_scale, clamp);
  return (int)cudaGetLastError();
}

// Per-tensor amax-scaled cast in ONE call: clears the caller's two-float scratch (d_work[0] = running maximum,
// d_work[1] = scale used), reduces |x|, and casts with scale = clamp(flt_max / amax) derived inside the cast kernel —
// three enqueues (memset, amax kernel, cast kernel), no host sync.  Replaces the per-tensor branch of
// quantize_tensor (mpemu/qutils.py:274-287).
extern "C" int fpemu_cast_amax_scaled(const void* in, void* out, int64_t n, int dtype, int format, int rounding,
                                      int flags, float flt_max, float* d_work, uint64_t seed, uint64_t elem_offset,
                                      void* stream_) {
  int st = validate_cast(n, dtype, format, rounding, 1.0f, in, out);
  if (st != FPEMU_OK) return st;
  if (d_work == nullptr) return FPEMU_ERR_ARG;
  if (format == FPEMU_FMT_BF16 || format == FPEMU_FMT_FP16) return FPEMU_ERR_MODE;   // these formats take no scale
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  cudaError_t err = cudaMemsetAsync(d_work, 0, 2 * sizeof(float), stream);
  if (err != cudaSuccess) return (int)err;
  if (n == 0) return FPEMU_OK;
  st = fpemu_amax(in, n, dtype, d_work, stream_);
  if (st != FPEMU_OK) return st;
  CastArgs a = make_args(in, out, n, format, rounding, flags, seed, elem_offset);
  a.d_amax = d_work;
  a.flt_max = flt_max;
  a.d_scale_out = d_work + 1;
  return launch_cast(a, dtype, stream);
}
