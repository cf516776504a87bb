// fpemu_multi.cu — multi-tensor launches: ONE amax launch and ONE cast launch for a whole list of tensors
// (a model's parameter list, a rank's row slabs).  The reference casts tensor by tensor from Python
// (mpemu/qutils.py:470-475, hybrid_emu.py:110-139); at 16-45 M elements per tensor a B200 spends as long ramping
// kernels up and down as streaming, and below ~1 M elements the launches themselves dominate.
//
// The caller describes the tensors once (fpemu_multi_table_build, host side), uploads the resulting table and
// passes its device address to the launches.  CTA b of a launch finds its tensor by comparing b against the
// per-tensor first-CTA indices in the table (one coalesced table read per CTA, one ballot), then runs the same
// per-vector code as the single-tensor kernels.
#include "fpemu_codec_core.cuh"
#include "fpemu_reduce.cuh"

namespace fpemu {

struct MtEntry {
  const void* in;
  void* out;
  int64_t n;
  uint64_t elem_offset;
  uint32_t cta_cast;      // first CTA of this tensor in the cast launch
  uint32_t cta_amax;      // first CTA of this tensor in the amax launch
};
struct MtHeader {
  int32_t count;
  int32_t dtype;          // element type in the low byte; bit 8 (kMtAnyOffset): some elem_offset is not a multiple of 16
  uint32_t total_cast;
  uint32_t total_amax;
};
static_assert(sizeof(MtEntry) == 40, "table layout");
static_assert(sizeof(MtHeader) == 16, "table layout");

constexpr int kMtAmaxUnroll = 4;
constexpr int kMtAnyOffset = 0x100;

__device__ __forceinline__ const MtEntry* mt_entries(const void* table) {
  return reinterpret_cast<const MtEntry*>(static_cast<const char*>(table) + sizeof(MtHeader));
}

// Index of the tensor that owns CTA `b`: the last entry whose first-CTA index is <= b.  All threads of the
// CTA take part; `AMAX` selects which of the two CTA numberings is used.
template <bool AMAX>
__device__ __forceinline__ int mt_find(const void* table, int count, uint32_t b) {
  __shared__ int s_idx;
  const MtEntry* e = mt_entries(table);
  for (int base = 0; base < count; base += blockDim.x) {
    const int t = base + threadIdx.x;
    if (t < count) {
      const uint32_t first = AMAX ? e[t].cta_amax : e[t].cta_cast;
      const uint32_t next = (t + 1 < count) ? (AMAX ? e[t + 1].cta_amax : e[t + 1].cta_cast) : 0xFFFFFFFFu;
      if (first <= b && b < next) s_idx = t;           // exactly one thread of one round matches
    }
  }
  __syncthreads();
  return s_idx;
}

// Scale of tensor t: the caller's d_scales[t], or — when the launch was given the vector of absolute maxima instead
// (proto.d_amax) — derived here exactly as fpemu_scale_from_amax_dtype does, so that an amax-scaled sweep over a tensor
// list needs no separate scale launch; the tensor's first CTA stores the scale it used into proto.d_scale_out[t].
//
// Exchange mode (row-sharded lists, fpemu_cast_multi_xchg): proto.in is non-null and points at this rank's arrival flags
// for the sweep's parity (uint32[world]), proto.n holds the world size, proto.elem_offset the sweep's epoch, and
// proto.d_amax is this rank's receive area float[world][count], into which every rank's amax_multi_push_kernel has
// written (or is about to write) its local maxima over NVLink.  The CTA waits until all flags carry the epoch, then
// takes the maximum over ranks of tensor t's entries — the all-reduce(MAX) of the sharded sweep, done by the kernels
// that produce and consume the vector instead of a collective launched between them.
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}

template <int DT>
__device__ __forceinline__ float mt_scale(const CastArgs& proto, const float* d_scales, int t, int count, bool first_cta) {
  if (proto.d_amax == nullptr) return __ldg(d_scales + t);
  if (proto.in != nullptr) {
    __shared__ uint32_t s_amax_bits;
    const int world = (int)proto.n;
    const uint32_t epoch = (uint32_t)proto.elem_offset;
    if (threadIdx.x < 32) {
      uint32_t bits = 0u;
      if ((int)threadIdx.x < world) {
        const uint32_t* flag = static_cast<const uint32_t*>(proto.in) + threadIdx.x;
        const long long t0 = clock64();
        bool ok = true;
        while (ld_acquire_sys(flag) != epoch) {
          if (clock64() - t0 > 20000000000LL) { ok = false; break; }   // ~10 s: a peer died; poison instead of hanging
          __nanosleep(64);
        }
        const volatile uint32_t* recv = reinterpret_cast<const volatile uint32_t*>(proto.d_amax);
        bits = ok ? recv[(size_t)threadIdx.x * (size_t)count + (size_t)t] : 0x7FFFFFFFu;
      }
      bits = warp_max_u32(bits);                               // NaN patterns sort above everything: they propagate
      if (threadIdx.x == 0) s_amax_bits = bits;
    }
    __syncthreads();
    const float s = scale_from_amax(__uint_as_float(s_amax_bits), proto.flt_max, true, DT);
    if (proto.d_scale_out != nullptr && first_cta && threadIdx.x == 0) proto.d_scale_out[t] = s;
    return s;
  }
  const float s = scale_from_amax(__ldg(proto.d_amax + t), proto.flt_max, true, DT);
  if (proto.d_scale_out != nullptr && first_cta && threadIdx.x == 0) proto.d_scale_out[t] = s;
  return s;
}

template <int FMT, int RND, int DT, bool SCALED, bool ANYOFF = false>
__global__ void __launch_bounds__(kThreads) cast_multi_kernel(const void* table, const int count, const CastArgs proto,
                                                             const float* d_scales) {
  constexpr int VE = DtInfo<DT>::kVecElems;
  constexpr int U = UnrollFor<DT>::value;
  const int t = mt_find<false>(table, count, blockIdx.x);
  const MtEntry e = mt_entries(table)[t];
  CastArgs a = proto;
  a.in = e.in; a.out = e.out; a.n = e.n; a.elem_offset = e.elem_offset;
  a.seed_lo = proto.seed_lo + (uint32_t)t;                       // tensor t uses seed + t
  a.seed_hi = proto.seed_hi + (a.seed_lo < proto.seed_lo ? 1u : 0u);
  a = with_device_seed(a);
  ScaleCtx sc;
  sc.scale = 1.0f; sc.rscale = 1.0f;
  sc.unscale32 = (a.flags & FPEMU_FLAG_UNSCALE_FP32) != 0;
  const int64_t local_cta = (int64_t)(blockIdx.x - e.cta_cast);
  const int64_t nvec = a.n / VE;
  const char* in = static_cast<const char*>(a.in);
  char* out = static_cast<char*>(a.out);
  const int64_t base = local_cta * (kThreads * U) + threadIdx.x;
  Vec32 v[U];
#pragma unroll
  for (int j = 0; j < U; ++j) {
    const int64_t vi = base + j * kThreads;
    if (vi < nvec) v[j] = ldg256(in + vi * 32);
  }
  // the scale AFTER the loads have been issued: in exchange mode it involves system-scope flag reads and a CTA barrier,
  // whose latency would otherwise sit in front of every short-lived CTA's memory traffic
  if (SCALED) {
    const float s = effective_scale(mt_scale<DT>(proto, d_scales, t, count, local_cta == 0), DT);
    sc.scale = s;
    sc.rscale = __frcp_rn(s);
  }
#pragma unroll
  for (int j = 0; j < U; ++j) {
    const int64_t vi = base + j * kThreads;
    if (vi < nvec) {
      process_vec<FMT, RND, false, DT, SCALED, ANYOFF>(v[j], sc, a, a.elem_offset + (uint64_t)vi * VE);
      stg256(out + vi * 32, v[j]);
    }
  }
  // ragged tail of this tensor: its last CTA, thread 0, through a zero-padded register vector
  const int tail = (int)(a.n - nvec * VE);
  const int64_t ctas_of_t = (nvec + kThreads * U - 1) / (kThreads * U);
  if (tail > 0 && local_cta == (ctas_of_t > 0 ? ctas_of_t - 1 : 0) && threadIdx.x == 0) {
    Vec32 w;
#pragma unroll
    for (int i = 0; i < 8; ++i) w.w[i] = 0u;
    const int64_t e0 = nvec * VE;
    if (DT == FPEMU_DT_F32) {
      const uint32_t* p = reinterpret_cast<const uint32_t*>(a.in) + e0;
      for (int i = 0; i < tail; ++i) w.w[i] = p[i];
    } else {
      const unsigned short* p = reinterpret_cast<const unsigned short*>(a.in) + e0;
      for (int i = 0; i < tail; ++i) w.w[i >> 1] |= (uint32_t)p[i] << ((i & 1) * 16);
    }
    process_vec<FMT, RND, false, DT, SCALED, ANYOFF>(w, sc, a, a.elem_offset + (uint64_t)e0);
    if (DT == FPEMU_DT_F32) {
      uint32_t* p = reinterpret_cast<uint32_t*>(a.out) + e0;
      for (int i = 0; i < tail; ++i) p[i] = w.w[i];
    } else {
      unsigned short* p = reinterpret_cast<unsigned short*>(a.out) + e0;
      for (int i = 0; i < tail; ++i) p[i] = (unsigned short)(w.w[i >> 1] >> ((i & 1) * 16));
    }
  }
}

// Same CTA numbering as cast_multi_kernel (so one table serves both), codes instead of values: the table's `out`
// pointers are the byte buffers.
template <int FMT, int RND, int DT, bool SCALED>
__global__ void __launch_bounds__(kThreads) encode_multi_kernel(const void* table, const int count, const CastArgs proto,
                                                               const float* d_scales) {
  constexpr int VE = DtInfo<DT>::kVecElems;
  constexpr int U = UnrollFor<DT>::value;
  const int t = mt_find<false>(table, count, blockIdx.x);
  const MtEntry e = mt_entries(table)[t];
  CastArgs a = proto;
  a.in = e.in; a.out = e.out; a.n = e.n; a.elem_offset = e.elem_offset;
  a.seed_lo = proto.seed_lo + (uint32_t)t;                       // tensor t uses seed + t
  a.seed_hi = proto.seed_hi + (a.seed_lo < proto.seed_lo ? 1u : 0u);
  a = with_device_seed(a);
  ScaleCtx sc;
  sc.scale = 1.0f; sc.rscale = 1.0f; sc.unscale32 = false;
  const int64_t local_cta = (int64_t)(blockIdx.x - e.cta_cast);
  if (SCALED) {
    const float s = effective_scale(mt_scale<DT>(proto, d_scales, t, count, local_cta == 0), DT);
    sc.scale = s;
    sc.rscale = __frcp_rn(s);
  }
  const int64_t ctas_of_t = (a.n / VE + kThreads * U - 1) / (kThreads * U);
  encode_cta<FMT, RND, false, DT, SCALED, U>(a, sc, local_cta, local_cta == (ctas_of_t > 0 ? ctas_of_t - 1 : 0));
}

template <int DT>
__device__ __forceinline__ void amax_multi_body(const void* table, const int count, float* d_amax) {
  constexpr int VE = DtInfo<DT>::kVecElems;
  const int t = mt_find<true>(table, count, blockIdx.x);
  const MtEntry e = mt_entries(table)[t];
  const char* in = static_cast<const char*>(e.in);
  const int64_t nvec = e.n / VE;
  const int64_t local_cta = (int64_t)(blockIdx.x - e.cta_amax);
  const int64_t base = local_cta * (kThreads * kMtAmaxUnroll) + threadIdx.x;
  float m = 0.0f;
  Vec32 v[kMtAmaxUnroll];
#pragma unroll
  for (int j = 0; j < kMtAmaxUnroll; ++j) {
    const int64_t vi = base + j * kThreads;
    if (vi < nvec) v[j] = ldg256(in + vi * 32);
    else {
#pragma unroll
      for (int i = 0; i < 8; ++i) v[j].w[i] = 0u;
    }
  }
#pragma unroll
  for (int j = 0; j < kMtAmaxUnroll; ++j) m = absmax_nan(m, vec_absmax<DT, true>(v[j]));
  if (local_cta == 0 && threadIdx.x < (int)(e.n - nvec * VE))
    m = absmax_nan(m, load_elem(e.in, nvec * VE + threadIdx.x, DT));
  uint32_t bits = warp_max_u32(__float_as_uint(m));
  __shared__ uint32_t warp_bits[kThreads / 32];
  if ((threadIdx.x & 31) == 0) warp_bits[threadIdx.x >> 5] = bits;
  __syncthreads();
  if (threadIdx.x < 32) {
    bits = threadIdx.x < kThreads / 32 ? warp_bits[threadIdx.x] : 0u;
    bits = warp_max_u32(bits);
    unsigned int* dst = reinterpret_cast<unsigned int*>(d_amax + t);
    if (threadIdx.x == 0 && bits > *reinterpret_cast<volatile unsigned int*>(dst)) atomicMax(dst, bits);
  }
}

template <int DT>
__global__ void __launch_bounds__(kThreads) amax_multi_kernel(const void* table, const int count, float* d_amax) {
  amax_multi_body<DT>(table, count, d_amax);
}

// The exchange step of a row-sharded sweep, fused into the reduction: the same per-tensor maxima as amax_multi_kernel,
// and the LAST CTA of the launch to finish (a device counter decides) pushes the finished local vector into every
// rank's receive area over NVLink (plain stores to peer-mapped memory), fences at system scope and raises this rank's
// arrival flag there.  It also re-zeroes the local vector and the counter, so the next sweep needs no memset.
struct MtPeers {
  float* recv[16];          // rank r's receive area for this parity, float[world][count], as mapped into this process
  uint32_t* flags[16];      // rank r's arrival flags for this parity, uint32[world]
  int world, rank;
  uint32_t epoch;
};

__device__ __forceinline__ void mt_push(const int count, float* d_amax, unsigned int* d_counter, const MtPeers& peers) {
  __threadfence();
  volatile uint32_t* local = reinterpret_cast<volatile uint32_t*>(d_amax);
  for (int t = threadIdx.x; t < count; t += kThreads) {
    const uint32_t bits = local[t];
    local[t] = 0u;
    for (int r = 0; r < peers.world; ++r)
      reinterpret_cast<volatile uint32_t*>(peers.recv[r])[(size_t)peers.rank * count + t] = bits;
  }
  __threadfence_system();
  __syncthreads();
  if ((int)threadIdx.x < peers.world) st_release_sys(peers.flags[threadIdx.x] + peers.rank, peers.epoch);
  if (threadIdx.x == 0) *d_counter = 0u;
}

template <int DT>
__global__ void __launch_bounds__(kThreads) amax_multi_push_kernel(const void* table, const int count, float* d_amax,
                                                                  unsigned int* d_counter, const MtPeers peers) {
  amax_multi_body<DT>(table, count, d_amax);
  __shared__ bool s_last;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_last = (atomicAdd(d_counter, 1u) == gridDim.x - 1);
  __syncthreads();
  if (!s_last) return;
  mt_push(count, d_amax, d_counter, peers);
}

// A rank whose slabs are all empty still has to arrive.
__global__ void __launch_bounds__(kThreads) push_only_kernel(const int count, float* d_amax, unsigned int* d_counter,
                                                            const MtPeers peers) {
  mt_push(count, d_amax, d_counter, peers);
}

typedef void (*MultiCastKernel)(const void*, const int, const CastArgs, const float*);

template <int FMT, int RND, int DT>
static MultiCastKernel mpick_scaled(bool scaled, bool anyoff) {
  constexpr bool kSR = (RND == FPEMU_RND_STOCHASTIC);
  if (FMT == FPEMU_FMT_BF16 || FMT == FPEMU_FMT_FP16)
    return (kSR && anyoff) ? cast_multi_kernel<FMT, RND, DT, false, kSR> : cast_multi_kernel<FMT, RND, DT, false>;
  if (kSR && anyoff) return scaled ? cast_multi_kernel<FMT, RND, DT, true, kSR> : cast_multi_kernel<FMT, RND, DT, false, kSR>;
  return scaled ? cast_multi_kernel<FMT, RND, DT, true> : cast_multi_kernel<FMT, RND, DT, false>;
}
template <int FMT, int RND>
static MultiCastKernel mpick_dt(int dt, bool scaled, bool anyoff) {
  switch (dt) {
    case FPEMU_DT_F32:  return mpick_scaled<FMT, RND, FPEMU_DT_F32>(scaled, anyoff);
    case FPEMU_DT_F16:  return mpick_scaled<FMT, RND, FPEMU_DT_F16>(scaled, anyoff);
    default:            return mpick_scaled<FMT, RND, FPEMU_DT_BF16>(scaled, anyoff);
  }
}
template <int FMT>
static MultiCastKernel mpick_rnd(int rnd, int dt, bool scaled, bool anyoff) {
  switch (rnd) {
    case FPEMU_RND_RNE:        return mpick_dt<FMT, FPEMU_RND_RNE>(dt, scaled, anyoff);
    case FPEMU_RND_RTZ:        return (FMT == FPEMU_FMT_FP16) ? nullptr : mpick_dt<FMT, FPEMU_RND_RTZ>(dt, scaled, anyoff);
    case FPEMU_RND_STOCHASTIC: return mpick_dt<FMT, FPEMU_RND_STOCHASTIC>(dt, scaled, anyoff);
    default:                   return nullptr;
  }
}

template <int FMT, int RND>
static MultiCastKernel epick_dt(int dt, bool scaled) {
  switch (dt) {
    case FPEMU_DT_F32:  return scaled ? encode_multi_kernel<FMT, RND, FPEMU_DT_F32, true> : encode_multi_kernel<FMT, RND, FPEMU_DT_F32, false>;
    case FPEMU_DT_F16:  return scaled ? encode_multi_kernel<FMT, RND, FPEMU_DT_F16, true> : encode_multi_kernel<FMT, RND, FPEMU_DT_F16, false>;
    default:            return scaled ? encode_multi_kernel<FMT, RND, FPEMU_DT_BF16, true> : encode_multi_kernel<FMT, RND, FPEMU_DT_BF16, false>;
  }
}
template <int FMT>
static MultiCastKernel epick_rnd(int rnd, int dt, bool scaled) {
  switch (rnd) {
    case FPEMU_RND_RNE:        return epick_dt<FMT, FPEMU_RND_RNE>(dt, scaled);
    case FPEMU_RND_RTZ:        return epick_dt<FMT, FPEMU_RND_RTZ>(dt, scaled);
    case FPEMU_RND_STOCHASTIC: return epick_dt<FMT, FPEMU_RND_STOCHASTIC>(dt, scaled);
    default:                   return nullptr;
  }
}

}  // namespace fpemu

using namespace fpemu;

extern "C" int64_t fpemu_multi_table_bytes(int count) {
  if (count < 0) return FPEMU_ERR_ARG;
  return (int64_t)sizeof(MtHeader) + (int64_t)count * (int64_t)sizeof(MtEntry);
}

extern "C" int fpemu_multi_table_build(const fpemu_tensor_desc* descs, int count, int dtype, void* h_table) {
  if (count <= 0 || descs == nullptr || h_table == nullptr) return FPEMU_ERR_ARG;
  if (dtype < 0 || dtype >= FPEMU_DT_COUNT) return FPEMU_ERR_DTYPE;
  const int ve = dtype == FPEMU_DT_F32 ? 8 : 16;
  const int u = dtype == FPEMU_DT_F32 ? UnrollFor<FPEMU_DT_F32>::value : UnrollFor<FPEMU_DT_F16>::value;
  MtHeader* h = static_cast<MtHeader*>(h_table);
  MtEntry* e = reinterpret_cast<MtEntry*>(static_cast<char*>(h_table) + sizeof(MtHeader));
  uint64_t cast_ctas = 0, amax_ctas = 0;
  bool any_offset = false;
  for (int t = 0; t < count; ++t) {
    const fpemu_tensor_desc& d = descs[t];
    if (d.n < 0 || (d.n > 0 && (d.in == nullptr || d.out == nullptr))) return FPEMU_ERR_ARG;
    // the multi-tensor kernels only have the vector path: 32-byte aligned buffers (any element offset)
    if ((((uintptr_t)d.in | (uintptr_t)d.out) & 31u) != 0) return FPEMU_ERR_ALIGN;
    e[t].in = d.in; e[t].out = d.out; e[t].n = d.n; e[t].elem_offset = d.elem_offset;
    any_offset = any_offset || (d.elem_offset & 15u) != 0;
    e[t].cta_cast = (uint32_t)cast_ctas;
    e[t].cta_amax = (uint32_t)amax_ctas;
    const int64_t nvec = d.n / ve;
    int64_t c = (nvec + (int64_t)kThreads * u - 1) / ((int64_t)kThreads * u);
    int64_t m = (nvec + (int64_t)kThreads * kMtAmaxUnroll - 1) / ((int64_t)kThreads * kMtAmaxUnroll);
    if (d.n > 0 && c == 0) c = 1;                                  // a tensor that is only a ragged tail
    if (d.n > 0 && m == 0) m = 1;
    cast_ctas += (uint64_t)c;
    amax_ctas += (uint64_t)m;
    if (cast_ctas > 0x7FFFFFFFull || amax_ctas > 0x7FFFFFFFull) return FPEMU_ERR_TOO_LARGE;
  }
  h->count = count; h->dtype = dtype | (any_offset ? kMtAnyOffset : 0); h->total_cast = (uint32_t)cast_ctas; h->total_amax = (uint32_t)amax_ctas;
  return FPEMU_OK;
}

extern "C" int fpemu_amax_multi(const void* d_table, const void* h_table, int count, float* d_amax, void* stream_) {
  if (d_table == nullptr || h_table == nullptr || d_amax == nullptr || count <= 0) return FPEMU_ERR_ARG;
  const MtHeader* h = static_cast<const MtHeader*>(h_table);
  if (h->count != count) return FPEMU_ERR_ARG;
  if (h->total_amax == 0) return FPEMU_OK;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  switch (h->dtype & 0xFF) {
    case FPEMU_DT_F32:  amax_multi_kernel<FPEMU_DT_F32><<<h->total_amax, kThreads, 0, stream>>>(d_table, count, d_amax); break;
    case FPEMU_DT_F16:  amax_multi_kernel<FPEMU_DT_F16><<<h->total_amax, kThreads, 0, stream>>>(d_table, count, d_amax); break;
    default:            amax_multi_kernel<FPEMU_DT_BF16><<<h->total_amax, kThreads, 0, stream>>>(d_table, count, d_amax); break;
  }
  return (int)cudaGetLastError();
}

namespace {

// Shared body of the four list launches: values or codes, scales given or derived from a vector of maxima.
int launch_multi(bool codes, const void* d_table, const void* h_table, int count, int format, int rounding, int flags,
                 const float* d_scales, const float* d_amax, float flt_max, float* d_scales_out, uint64_t seed,
                 void* stream_, const uint32_t* xchg_flags = nullptr, int world = 0, uint32_t epoch = 0) {
  if (d_table == nullptr || h_table == nullptr || count <= 0) return FPEMU_ERR_ARG;
  const MtHeader* h = static_cast<const MtHeader*>(h_table);
  if (h->count != count) return FPEMU_ERR_ARG;
  if (format < 0 || format >= FPEMU_FMT_COUNT || rounding < 0 || rounding >= FPEMU_RND_COUNT) return FPEMU_ERR_MODE;
  if (flags & FPEMU_FLAG_DAZ) return FPEMU_ERR_MODE;             // DAZ variants: use the per-tensor entry point
  const bool takes_scale = format != FPEMU_FMT_BF16 && format != FPEMU_FMT_FP16;
  if (d_amax != nullptr && !takes_scale) return FPEMU_ERR_MODE;
  const bool scaled = (d_scales != nullptr || d_amax != nullptr) && takes_scale;
  MultiCastKernel k = nullptr;
  const int dt = h->dtype & 0xFF;
  const bool anyoff = (h->dtype & kMtAnyOffset) != 0 && rounding == FPEMU_RND_STOCHASTIC;
  if (codes && anyoff) return FPEMU_ERR_ALIGN;                   // the code kernels draw per aligned vector only
  if (codes) {
    switch (format) {
      case FPEMU_FMT_E5M2: k = epick_rnd<FPEMU_FMT_E5M2>(rounding, dt, scaled); break;
      case FPEMU_FMT_E4M3: k = epick_rnd<FPEMU_FMT_E4M3>(rounding, dt, scaled); break;
      case FPEMU_FMT_E3M4: k = epick_rnd<FPEMU_FMT_E3M4>(rounding, dt, scaled); break;
      default: break;                                            // only the three 8-bit formats have codes
    }
  } else {
    switch (format) {
      case FPEMU_FMT_E5M2: k = mpick_rnd<FPEMU_FMT_E5M2>(rounding, dt, scaled, anyoff); break;
      case FPEMU_FMT_E4M3: k = mpick_rnd<FPEMU_FMT_E4M3>(rounding, dt, scaled, anyoff); break;
      case FPEMU_FMT_E3M4: k = mpick_rnd<FPEMU_FMT_E3M4>(rounding, dt, scaled, anyoff); break;
      case FPEMU_FMT_BF16: k = mpick_rnd<FPEMU_FMT_BF16>(rounding, dt, false, anyoff); break;
      case FPEMU_FMT_FP16: k = mpick_rnd<FPEMU_FMT_FP16>(rounding, dt, false, anyoff); break;
      default: break;
    }
  }
  if (k == nullptr) return FPEMU_ERR_MODE;                       // rare modes: use the per-tensor entry point
  if (h->total_cast == 0) return FPEMU_OK;
  CastArgs proto;
  proto.in = xchg_flags; proto.out = nullptr; proto.n = world;       // exchange mode: see mt_scale()
  proto.scale = 1.0f; proto.d_scale = nullptr;
  proto.seed_lo = (uint32_t)seed; proto.seed_hi = (uint32_t)(seed >> 32);
  proto.elem_offset = epoch;
  proto.fmt = format; proto.rnd = rounding; proto.flags = flags;
  proto.d_amax = d_amax; proto.flt_max = flt_max; proto.d_scale_out = d_scales_out; proto.d_seed = seed_source();
  k<<<h->total_cast, kThreads, 0, static_cast<cudaStream_t>(stream_)>>>(d_table, count, proto, d_scales);
  return (int)cudaGetLastError();
}

}  // namespace

extern "C" int fpemu_cast_multi(const void* d_table, const void* h_table, int count, int format, int rounding, int flags,
                                const float* d_scales, uint64_t seed, void* stream_) {
  return launch_multi(false, d_table, h_table, count, format, rounding, flags, d_scales, nullptr, 0.0f, nullptr, seed, stream_);
}

extern "C" int fpemu_encode_multi(const void* d_table, const void* h_table, int count, int format, int rounding, int flags,
                                  const float* d_scales, uint64_t seed, void* stream_) {
  return launch_multi(true, d_table, h_table, count, format, rounding, flags, d_scales, nullptr, 0.0f, nullptr, seed, stream_);
}

extern "C" int fpemu_cast_multi_amax(const void* d_table, const void* h_table, int count, int format, int rounding,
                                     int flags, const float* d_amax, float flt_max, float* d_scales_out, uint64_t seed,
                                     void* stream_) {
  if (d_amax == nullptr) return FPEMU_ERR_ARG;
  return launch_multi(false, d_table, h_table, count, format, rounding, flags, nullptr, d_amax, flt_max, d_scales_out, seed,
                      stream_);
}

extern "C" int fpemu_encode_multi_amax(const void* d_table, const void* h_table, int count, int format, int rounding,
                                       int flags, const float* d_amax, float flt_max, float* d_scales_out, uint64_t seed,
                                       void* stream_) {
  if (d_amax == nullptr) return FPEMU_ERR_ARG;
  return launch_multi(true, d_table, h_table, count, format, rounding, flags, nullptr, d_amax, flt_max, d_scales_out, seed,
                      stream_);
}

// ---- exchange of the maxima over peer memory (row-sharded lists on the GPUs of one NVLink domain) ----------------------
// Exchange buffer of one rank, in floats: recv[parity 0..1][source rank][count], then flags[parity][source rank].
extern "C" int64_t fpemu_exchange_floats(int count, int world) {
  if (count <= 0 || world <= 0 || world > 16) return FPEMU_ERR_ARG;
  return 2LL * world * count + 2LL * world;
}

extern "C" int fpemu_amax_multi_push(const void* d_table, const void* h_table, int count, float* d_amax,
                                     unsigned int* d_counter, void* const* peer_bufs, int world, int rank, uint32_t epoch,
                                     void* stream_) {
  if (d_table == nullptr || h_table == nullptr || d_amax == nullptr || d_counter == nullptr || peer_bufs == nullptr ||
      count <= 0 || world <= 0 || world > 16 || rank < 0 || rank >= world || epoch == 0)
    return FPEMU_ERR_ARG;
  const MtHeader* h = static_cast<const MtHeader*>(h_table);
  if (h->count != count) return FPEMU_ERR_ARG;
  MtPeers peers;
  const int parity = (int)(epoch & 1u);
  for (int r = 0; r < 16; ++r) { peers.recv[r] = nullptr; peers.flags[r] = nullptr; }
  for (int r = 0; r < world; ++r) {
    if (peer_bufs[r] == nullptr) return FPEMU_ERR_ARG;
    float* base = static_cast<float*>(peer_bufs[r]);
    peers.recv[r] = base + (size_t)parity * world * count;
    peers.flags[r] = reinterpret_cast<uint32_t*>(base + 2 * (size_t)world * count) + (size_t)parity * world;
  }
  peers.world = world; peers.rank = rank; peers.epoch = epoch;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (h->total_amax == 0) {
    push_only_kernel<<<1, kThreads, 0, stream>>>(count, d_amax, d_counter, peers);
    return (int)cudaGetLastError();
  }
  switch (h->dtype & 0xFF) {
    case FPEMU_DT_F32:  amax_multi_push_kernel<FPEMU_DT_F32><<<h->total_amax, kThreads, 0, stream>>>(d_table, count, d_amax, d_counter, peers); break;
    case FPEMU_DT_F16:  amax_multi_push_kernel<FPEMU_DT_F16><<<h->total_amax, kThreads, 0, stream>>>(d_table, count, d_amax, d_counter, peers); break;
    default:            amax_multi_push_kernel<FPEMU_DT_BF16><<<h->total_amax, kThreads, 0, stream>>>(d_table, count, d_amax, d_counter, peers); break;
  }
  return (int)cudaGetLastError();
}

static int launch_xchg(bool codes, const void* d_table, const void* h_table, int count, int format, int rounding, int flags,
                       const void* my_buf, int world, uint32_t epoch, float flt_max, float* d_scales_out, uint64_t seed,
                       void* stream_) {
  if (my_buf == nullptr || world <= 0 || world > 16 || epoch == 0 || count <= 0) return FPEMU_ERR_ARG;
  const int parity = (int)(epoch & 1u);
  const float* base = static_cast<const float*>(my_buf);
  const float* recv = base + (size_t)parity * world * count;
  const uint32_t* flg = reinterpret_cast<const uint32_t*>(base + 2 * (size_t)world * count) + (size_t)parity * world;
  return launch_multi(codes, d_table, h_table, count, format, rounding, flags, nullptr, recv, flt_max, d_scales_out, seed,
                      stream_, flg, world, epoch);
}

extern "C" int fpemu_cast_multi_xchg(const void* d_table, const void* h_table, int count, int format, int rounding,
                                     int flags, const void* my_buf, int world, uint32_t epoch, float flt_max,
                                     float* d_scales_out, uint64_t seed, void* stream_) {
  return launch_xchg(false, d_table, h_table, count, format, rounding, flags, my_buf, world, epoch, flt_max, d_scales_out,
                     seed, stream_);
}

extern "C" int fpemu_encode_multi_xchg(const void* d_table, const void* h_table, int count, int format, int rounding,
                                       int flags, const void* my_buf, int world, uint32_t epoch, float flt_max,
                                       float* d_scales_out, uint64_t seed, void* stream_) {
  return launch_xchg(true, d_table, h_table, count, format, rounding, flags, my_buf, world, epoch, flt_max, d_scales_out,
                     seed, stream_);
}
