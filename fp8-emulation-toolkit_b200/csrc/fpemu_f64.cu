// fpemu_f64.cu — float64 tensors.  The reference dispatches its kernels for double as well
// (AT_DISPATCH_FLOATING_TYPES_AND_HALF, cuda/fpemu_kernels.cu:1059 ...): the element is narrowed to float
// (`__anyfloat2float_rn`, :93-105; for the scaled formats `float inval = in[gid] * scale` with BOTH operands double,
// :404, :420), rounded exactly like a float tensor's element, and the result widened back (`__half2anyfloat`, :136-148).
// Here: two 256-bit loads = 8 doubles per thread-vector, narrowed in registers, the fp32 per-vector code of the cast
// (process_vec), two 256-bit stores.  16 B / element of DRAM traffic, HBM-bound.
#include "fpemu_cast_core.cuh"

namespace fpemu {

union Dbl4 {
  Vec32 v;
  double d[4];
};

// The float the reference's kernels start from: (float)x for the formats without a scale, (float)(x * (double)scale)
// for the others.
template <bool SCALED>
__device__ __forceinline__ float narrow(double x, double scale) {
  return SCALED ? __double2float_rn(x * scale) : __double2float_rn(x);
}

template <int FMT, int RND, bool SCALED>
__global__ void __launch_bounds__(kThreads) cast_f64_kernel(const CastArgs a_) {
  const CastArgs a = with_device_seed(a_);
  ScaleCtx sc = make_scale_ctx<FPEMU_DT_F32, SCALED>(a);
  const double dscale = (double)sc.scale;
  sc.scale = 1.0f;                                    // the multiplication has happened in double; process_vec un-scales only
  const int64_t nvec = a.n / 8;
  const char* in = static_cast<const char*>(a.in);
  char* out = static_cast<char*>(a.out);
  const int64_t vi = (int64_t)blockIdx.x * kThreads + threadIdx.x;
  if (vi < nvec) {
    Dbl4 lo, hi;
    lo.v = ldg256(in + vi * 64);
    hi.v = ldg256(in + vi * 64 + 32);
    Vec32 f;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      f.w[i] = __float_as_uint(narrow<SCALED>(lo.d[i], dscale));
      f.w[4 + i] = __float_as_uint(narrow<SCALED>(hi.d[i], dscale));
    }
    process_vec<FMT, RND, false, FPEMU_DT_F32, SCALED>(f, sc, a, a.elem_offset + (uint64_t)vi * 8);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      lo.d[i] = (double)__uint_as_float(f.w[i]);
      hi.d[i] = (double)__uint_as_float(f.w[4 + i]);
    }
    stg256(out + vi * 64, lo.v);
    stg256(out + vi * 64 + 32, hi.v);
  }
  // ragged tail (< 8 elements): zero-padded register vector through the same code
  const int tail = (int)(a.n - nvec * 8);
  if (tail > 0 && blockIdx.x == gridDim.x - 1 && threadIdx.x == 0) {
    const double* p = static_cast<const double*>(a.in) + nvec * 8;
    double* q = static_cast<double*>(a.out) + nvec * 8;
    Vec32 f;
#pragma unroll
    for (int i = 0; i < 8; ++i) f.w[i] = 0u;
    for (int i = 0; i < tail; ++i) f.w[i] = __float_as_uint(narrow<SCALED>(p[i], dscale));
    process_vec<FMT, RND, false, FPEMU_DT_F32, SCALED>(f, sc, a, a.elem_offset + (uint64_t)nvec * 8);
    for (int i = 0; i < tail; ++i) q[i] = (double)__uint_as_float(f.w[i]);
  }
}

// Every mode, any alignment and element offset: one element per thread and step.
__global__ void __launch_bounds__(256) cast_f64_generic_kernel(const CastArgs a_) {
  const CastArgs a = with_device_seed(a_);
  const bool takes_scale = a.fmt != FPEMU_FMT_BF16 && a.fmt != FPEMU_FMT_FP16;
  const bool scaled = takes_scale && ((a.d_scale != nullptr) || (a.scale != 1.0f));
  float s = 1.0f, rs = 1.0f;
  if (scaled) {
    s = a.d_scale ? __ldg(a.d_scale) : a.scale;
    rs = __frcp_rn(s);
  }
  const bool wide = !takes_scale;
  const double* in = static_cast<const double*>(a.in);
  double* out = static_cast<double*>(a.out);
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < a.n; i += stride) {
    const float x = scaled ? __double2float_rn(in[i] * (double)s) : __double2float_rn(in[i]);
    uint32_t rb = 0u;
    if (a.rnd == FPEMU_RND_STOCHASTIC) rb = rand_bits_one(a.seed_lo, a.seed_hi, a.elem_offset + (uint64_t)i, wide);
    // the scale has been applied in double: cast_one multiplies by 1 and un-scales by rs
    out[i] = (double)cast_one(x, a.fmt, a.rnd, a.flags, scaled, 1.0f, rs, false, rb);
  }
}

typedef void (*F64Kernel)(const CastArgs);

template <int FMT>
static F64Kernel f64_rnd(int rnd, bool scaled) {
  constexpr bool kTakesScale = (FMT != FPEMU_FMT_BF16 && FMT != FPEMU_FMT_FP16);
  switch (rnd) {
    case FPEMU_RND_RNE:
      return (kTakesScale && scaled) ? cast_f64_kernel<FMT, FPEMU_RND_RNE, kTakesScale> : cast_f64_kernel<FMT, FPEMU_RND_RNE, false>;
    case FPEMU_RND_STOCHASTIC:
      return (kTakesScale && scaled) ? cast_f64_kernel<FMT, FPEMU_RND_STOCHASTIC, kTakesScale>
                                     : cast_f64_kernel<FMT, FPEMU_RND_STOCHASTIC, false>;
    case FPEMU_RND_RTZ:
      if (FMT == FPEMU_FMT_FP16) return nullptr;
      return (kTakesScale && scaled) ? cast_f64_kernel<FMT, FPEMU_RND_RTZ, kTakesScale> : cast_f64_kernel<FMT, FPEMU_RND_RTZ, false>;
    default:
      return nullptr;
  }
}

int validate_cast(int64_t n, int dtype, int format, int rounding, float scale, const void* in, const void* out);

}  // namespace fpemu

extern "C" int fpemu_cast_f64(const void* in, void* out, int64_t n, int format, int rounding, int flags, float scale,
                              const float* d_scale, uint64_t seed, uint64_t elem_offset, void* stream_) {
  using namespace fpemu;
  int st = validate_cast(n, FPEMU_DT_F32, format, rounding, d_scale ? 1.0f : scale, in, out);
  if (st != FPEMU_OK) return st;
  if (((uintptr_t)in | (uintptr_t)out) & 7u) return FPEMU_ERR_ALIGN;
  if (n == 0) return FPEMU_OK;
  CastArgs a;
  a.in = in; a.out = out; a.n = n;
  a.scale = scale; a.d_scale = d_scale;
  a.seed_lo = (uint32_t)seed; a.seed_hi = (uint32_t)(seed >> 32);
  a.elem_offset = elem_offset;
  a.fmt = format; a.rnd = rounding; a.flags = flags;
  a.d_amax = nullptr; a.flt_max = 0.0f; a.d_scale_out = nullptr; a.d_seed = seed_source();
  if (format == FPEMU_FMT_BF16 || format == FPEMU_FMT_FP16) { a.scale = 1.0f; a.d_scale = nullptr; }
  const bool scaled = (a.d_scale != nullptr) || (a.scale != 1.0f);
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  const bool aligned = ((((uintptr_t)in | (uintptr_t)out) & 31u) == 0);
  const bool rng_ok = (rounding != FPEMU_RND_STOCHASTIC) || ((elem_offset & 15u) == 0);
  F64Kernel k = nullptr;
  if (aligned && rng_ok && !(flags & FPEMU_FLAG_DAZ)) {
    switch (format) {
      case FPEMU_FMT_E5M2: k = f64_rnd<FPEMU_FMT_E5M2>(rounding, scaled); break;
      case FPEMU_FMT_E4M3: k = f64_rnd<FPEMU_FMT_E4M3>(rounding, scaled); break;
      case FPEMU_FMT_E3M4: k = f64_rnd<FPEMU_FMT_E3M4>(rounding, scaled); break;
      case FPEMU_FMT_BF16: k = f64_rnd<FPEMU_FMT_BF16>(rounding, false); break;
      case FPEMU_FMT_FP16: k = f64_rnd<FPEMU_FMT_FP16>(rounding, false); break;
      case FPEMU_FMT_FP4:
        if (rounding == FPEMU_RND_RNE)
          k = scaled ? cast_f64_kernel<FPEMU_FMT_FP4, FPEMU_RND_RNE, true> : cast_f64_kernel<FPEMU_FMT_FP4, FPEMU_RND_RNE, false>;
        break;
      default: break;
    }
  }
  if (k != nullptr) {
    int64_t want = (n / 8 + kThreads - 1) / kThreads;
    if (want < 1) want = 1;
    if (want > 0x7FFFFFFFLL) return FPEMU_ERR_TOO_LARGE;
    k<<<(unsigned)want, kThreads, 0, stream>>>(a);
  } else {
    int64_t want = (n + 255) / 256;
    const int64_t cap = (int64_t)device_info().sm_count * 16;
    cast_f64_generic_kernel<<<(int)(want < cap ? want : cap), 256, 0, stream>>>(a);
  }
  return (int)cudaGetLastError();
}
