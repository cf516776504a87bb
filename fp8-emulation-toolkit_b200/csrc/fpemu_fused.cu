// fpemu_fused.cu — per-tensor amax-scaled cast of a small / mid-size tensor in ONE launch
// (C ABI: fpemu_cast_amax_scaled_fused).
//
// The regular one-call path (fpemu_cast_amax_scaled) is memset + amax kernel + cast kernel: three enqueues and
// 12 B/elem (fp32) of traffic, which is what a tensor far larger than L2 needs.  The tensors the module hooks cast
// thousands of times per training step (activations and weights of a few MB, mpemu/qutils.py:274-287 called from
// :478-616) are launch-bound there.  Here the whole grid is resident at once (cooperative launch), every thread
// keeps its share of the tensor in registers across a grid-wide barrier, and the tensor is read once:
//     load R vectors per thread -> CTA maximum -> slot[blockIdx] -> grid.sync() -> every CTA reduces the slots ->
//     scale = clamp(flt_max / amax) -> round from the registers -> store.
// 8 B/elem of traffic, one enqueue, no memset (each slot is written before it is read), no atomics.
// Capacity = resident threads x R vectors (about 4.8 M fp32 / 9.7 M 16-bit elements on a B200: half of the register
// file holds data); anything larger, unaligned or in a mode without a packed kernel takes the three-launch path.
#include <cooperative_groups.h>

#include <mutex>

#include "fpemu_reduce.cuh"

namespace cg = cooperative_groups;

namespace fpemu {

int launch_cast(const CastArgs& a, int dtype, cudaStream_t stream);
int validate_cast(int64_t n, int dtype, int format, int rounding, float scale, const void* in, const void* out);

constexpr int kFusedR = 4;                 // 32-byte vectors a thread keeps in registers across the barrier

template <int FMT, int RND, int DT>
__global__ void __launch_bounds__(kThreads) fused_amax_cast_kernel(const CastArgs a_, float* d_work) {
  const CastArgs a = with_device_seed(a_);
  constexpr int VE = DtInfo<DT>::kVecElems;
  constexpr int R = kFusedR;
  cg::grid_group grid = cg::this_grid();
  const int64_t nvec = a.n / VE;
  const char* in = static_cast<const char*>(a.in);
  char* out = static_cast<char*>(a.out);
  uint32_t* slots = reinterpret_cast<uint32_t*>(d_work + 2);
  // vector j of this thread: CTAs interleave at 256-vector granularity so that a partly filled grid spreads evenly
  const int64_t first = (int64_t)blockIdx.x * kThreads + threadIdx.x;
  const int64_t step = (int64_t)gridDim.x * kThreads;
  Vec32 v[R];
  float m = 0.0f;
#pragma unroll
  for (int j = 0; j < R; ++j) {
    const int64_t vi = first + j * step;
    if (vi < nvec) v[j] = ldg256(in + vi * 32);
    else {
#pragma unroll
      for (int i = 0; i < 8; ++i) v[j].w[i] = 0u;
    }
  }
#pragma unroll
  for (int j = 0; j < R; ++j) m = absmax_nan(m, vec_absmax<DT, true>(v[j]));
  // ragged tail (< one vector): CTA 0 keeps it in a zero-padded register vector as well
  const int tail = (int)(a.n - nvec * VE);
  Vec32 tv;
#pragma unroll
  for (int i = 0; i < 8; ++i) tv.w[i] = 0u;
  const bool owns_tail = tail > 0 && blockIdx.x == 0 && threadIdx.x == 0;
  if (owns_tail) {
    const int64_t e0 = nvec * VE;
    if (DT == FPEMU_DT_F32) {
      const uint32_t* p = reinterpret_cast<const uint32_t*>(a.in) + e0;
      for (int i = 0; i < tail; ++i) tv.w[i] = p[i];
    } else {
      const unsigned short* p = reinterpret_cast<const unsigned short*>(a.in) + e0;
      for (int i = 0; i < tail; ++i) tv.w[i >> 1] |= (uint32_t)p[i] << ((i & 1) * 16);
    }
    m = absmax_nan(m, vec_absmax<DT, true>(tv));
  }
  __shared__ uint32_t warp_bits[kThreads / 32];
  uint32_t bits = warp_max_u32(__float_as_uint(m));
  if ((threadIdx.x & 31) == 0) warp_bits[threadIdx.x >> 5] = bits;
  __syncthreads();
  if (threadIdx.x == 0) {
#pragma unroll
    for (int w = 1; w < kThreads / 32; ++w) bits = max(bits, warp_bits[w]);
    slots[blockIdx.x] = bits;
  }
  grid.sync();                               // includes the memory fence that makes every slot visible
  uint32_t all = 0u;
  for (int i = threadIdx.x; i < (int)gridDim.x; i += kThreads) all = max(all, __ldcg(slots + i));
  all = warp_max_u32(all);
  __syncthreads();                           // warp_bits is reused
  if ((threadIdx.x & 31) == 0) warp_bits[threadIdx.x >> 5] = all;
  __syncthreads();
  all = warp_bits[0];
#pragma unroll
  for (int w = 1; w < kThreads / 32; ++w) all = max(all, warp_bits[w]);
  const float amax = __uint_as_float(all);
  const float s = scale_from_amax(amax, a.flt_max, true, DT);   // mpemu/qutils.py:283-285
  if (blockIdx.x == 0 && threadIdx.x == 0) { d_work[0] = amax; d_work[1] = s; }
  ScaleCtx sc;
  sc.unscale32 = (a.flags & FPEMU_FLAG_UNSCALE_FP32) != 0;
  sc.scale = effective_scale(s, DT);
  sc.rscale = __frcp_rn(sc.scale);
#pragma unroll
  for (int j = 0; j < R; ++j) {
    const int64_t vi = first + j * step;
    if (vi < nvec) {
      process_vec<FMT, RND, false, DT, true>(v[j], sc, a, a.elem_offset + (uint64_t)vi * VE);
      stg256(out + vi * 32, v[j]);
    }
  }
  if (owns_tail) {
    const int64_t e0 = nvec * VE;
    process_vec<FMT, RND, false, DT, true>(tv, sc, a, a.elem_offset + (uint64_t)e0);
    if (DT == FPEMU_DT_F32) {
      uint32_t* p = reinterpret_cast<uint32_t*>(a.out) + e0;
      for (int i = 0; i < tail; ++i) p[i] = tv.w[i];
    } else {
      unsigned short* p = reinterpret_cast<unsigned short*>(a.out) + e0;
      for (int i = 0; i < tail; ++i) p[i] = (unsigned short)(tv.w[i >> 1] >> ((i & 1) * 16));
    }
  }
}

typedef void (*FusedKernel)(const CastArgs, float*);

template <int FMT, int RND>
static FusedKernel fused_dt(int dt) {
  switch (dt) {
    case FPEMU_DT_F32:  return fused_amax_cast_kernel<FMT, RND, FPEMU_DT_F32>;
    case FPEMU_DT_F16:  return fused_amax_cast_kernel<FMT, RND, FPEMU_DT_F16>;
    default:            return fused_amax_cast_kernel<FMT, RND, FPEMU_DT_BF16>;
  }
}
template <int FMT>
static FusedKernel fused_rnd(int rnd, int dt) {
  switch (rnd) {
    case FPEMU_RND_RNE:        return fused_dt<FMT, FPEMU_RND_RNE>(dt);
    case FPEMU_RND_RTZ:        return fused_dt<FMT, FPEMU_RND_RTZ>(dt);
    case FPEMU_RND_STOCHASTIC: return fused_dt<FMT, FPEMU_RND_STOCHASTIC>(dt);
    default:                   return nullptr;
  }
}
static FusedKernel fused_pick(int fmt, int rnd, int dt) {
  switch (fmt) {
    case FPEMU_FMT_E5M2: return fused_rnd<FPEMU_FMT_E5M2>(rnd, dt);
    case FPEMU_FMT_E4M3: return fused_rnd<FPEMU_FMT_E4M3>(rnd, dt);
    case FPEMU_FMT_E3M4: return fused_rnd<FPEMU_FMT_E3M4>(rnd, dt);
    default:             return nullptr;
  }
}

// Resident CTAs of one kernel on the current device (cached per kernel and device by the runtime call's cost: a few us,
// so cache here).
static int resident_ctas(FusedKernel k) {
  struct Entry { FusedKernel k; int dev; int ctas; };
  static Entry cache[128];
  static int used = 0;
  static std::mutex mu;
  int dev = 0;
  cudaGetDevice(&dev);
  std::lock_guard<std::mutex> lock(mu);
  for (int i = 0; i < used; ++i)
    if (cache[i].k == k && cache[i].dev == dev) return cache[i].ctas;
  int per_sm = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k, kThreads, 0) != cudaSuccess || per_sm <= 0) return 0;
  const int ctas = per_sm * device_info().sm_count;
  if (used < 128) cache[used++] = Entry{k, dev, ctas};
  return ctas;
}

}  // namespace fpemu

extern "C" int fpemu_fused_work_floats(void) { return 2 + 4096; }

extern "C" int fpemu_cast_amax_scaled_fused(const void* in, void* out, int64_t n, int dtype, int format, int rounding,
                                            int flags, float flt_max, float* d_work, int work_floats, uint64_t seed,
                                            uint64_t elem_offset, void* stream_) {
  using namespace fpemu;
  int st = validate_cast(n, dtype, format, rounding, 1.0f, in, out);
  if (st != FPEMU_OK) return st;
  if (d_work == nullptr || work_floats < 2) return FPEMU_ERR_ARG;
  if (format == FPEMU_FMT_BF16 || format == FPEMU_FMT_FP16) return FPEMU_ERR_MODE;
  const bool aligned = (((uintptr_t)in | (uintptr_t)out) & 31u) == 0;
  const bool rng_ok = (rounding != FPEMU_RND_STOCHASTIC) || ((elem_offset & 15u) == 0);
  FusedKernel k = (aligned && rng_ok && !(flags & FPEMU_FLAG_DAZ) && n > 0) ? fused_pick(format, rounding, dtype) : nullptr;
  if (k != nullptr) {
    const int ve = dtype == FPEMU_DT_F32 ? 8 : 16;
    const int64_t nvec = n / ve;
    // as many CTAs as are resident at once, up to one vector per thread: short tensors spread over all SMs, long
    // ones put up to kFusedR vectors on a thread
    const int resident = resident_ctas(k);
    int64_t ctas = (nvec + kThreads - 1) / kThreads;
    if (ctas > resident) ctas = resident;
    if (ctas < 1) ctas = 1;
    if (ctas * kThreads * kFusedR >= nvec && ctas + 2 <= work_floats && resident > 0) {
      CastArgs a;
      a.in = in; a.out = out; a.n = n;
      a.scale = 1.0f; a.d_scale = nullptr;
      a.seed_lo = (uint32_t)seed; a.seed_hi = (uint32_t)(seed >> 32);
      a.elem_offset = elem_offset;
      a.fmt = format; a.rnd = rounding; a.flags = flags;
      a.d_amax = nullptr; a.flt_max = flt_max; a.d_scale_out = nullptr; a.d_seed = seed_source();
      void* args[] = {(void*)&a, (void*)&d_work};
      cudaError_t err = cudaLaunchCooperativeKernel((const void*)k, dim3((unsigned)ctas), dim3(kThreads), args, 0,
                                                    static_cast<cudaStream_t>(stream_));
      if (err != cudaErrorCooperativeLaunchTooLarge && err != cudaErrorNotSupported) return (int)err;
      cudaGetLastError();                  // fewer SMs available than the device reports (MPS limits, green contexts):
    }                                      // nothing was enqueued, take the three-launch path below
  }
  return fpemu_cast_amax_scaled(in, out, n, dtype, format, rounding, flags, flt_max, d_work, seed, elem_offset, stream_);
}
