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
  int32_t dtype;          // element type in the low byte; bit 8 (kMtAnyOffset): some elem_offset is not a multiple of 16;
                          // bit 9 (kMtHasEmpty): some n is 0
  uint32_t total_cast;
  uint32_t total_amax;
};
static_assert(sizeof(MtEntry) == 40, "table layout");
static_assert(sizeof(MtHeader) == 16, "table layout");

constexpr int kMtAmaxUnroll = 4;
constexpr int kMtAnyOffset = 0x100;
constexpr int kMtHasEmpty = 0x200;        // some tensor has no elements

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
// Exchange mode (row-sharded lists, fpemu_cast_multi_xchg): proto.in is non-null, proto.n holds the world size,
// proto.elem_offset the sweep's epoch, and proto.d_amax is this rank's receive area for the sweep's parity: one 64-bit
// word {maximum bits, epoch} per (source rank, tensor), which every rank's push_kernel writes over NVLink with ONE
// 8-byte store per word.  The epoch travels with the value, so a word whose upper half equals the sweep's epoch is
// complete by itself: no separate flag, no fence, one load per rank on the reader's side.  Every warp polls its
// tensor's `world` words (one lane each, L2 hits) and takes the maximum over ranks — the all-reduce(MAX) of the
// sharded sweep, done by the kernels that produce and consume the vector instead of a collective launched between.
//
// The area holds kXchgCopies identical copies, uint64[copy][world][count], and a warp reads the copy its CTA and warp
// number select.  Polls must bypass L1, and at any moment all resident CTAs work on the same one or two tensors: with
// a single copy every poll of the launch (8 per CTA, ~10^6 per ms) lands on the same 32-byte sector of one L2 slice,
// which then runs ~40 % busy with polls alone and holds back the streaming traffic that is interleaved over all
// slices (measured: the cast launch 4 % slower).  Copies spread the polls over the slices.
constexpr int kXchgCopies = 32;

__device__ __forceinline__ uint64_t ld_relaxed_sys_u64(const uint64_t* p) {
  uint64_t v;
  asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed_sys_u64(uint64_t* p, uint64_t v) {
  asm volatile("st.relaxed.sys.global.u64 [%0], %1;" :: "l"(p), "l"(v) : "memory");
}

// The poll in two halves, so that a kernel can put its own loads between them: the first read of the word is issued
// BEFORE the CTA's vector loads and only looked at after them.  (Issued behind them it also comes back behind them —
// the requests of one SM queue in order on their way to L2 — and every CTA then waits one more round trip than it has
// to: measured 3 % of the cast launch.)  All 32 lanes of a warp call these together.
struct MtPoll {
  uint64_t* row;
  uint64_t w;
};

__device__ __forceinline__ MtPoll mt_poll_begin(const CastArgs& proto, int t, int count) {
  MtPoll p;
  p.row = nullptr; p.w = 0ull;
  const int lane = threadIdx.x & 31;
  if (lane < (int)proto.n) {
    const uint32_t copy = (blockIdx.x * (kThreads / 32) + (threadIdx.x >> 5)) % kXchgCopies;
    p.row = reinterpret_cast<uint64_t*>(const_cast<float*>(proto.d_amax)) +
            ((size_t)copy * (size_t)proto.n + (size_t)lane) * (size_t)count;
    p.w = ld_relaxed_sys_u64(p.row + t);
  }
  return p;
}

__device__ __forceinline__ float mt_poll_finish(const CastArgs& proto, MtPoll p, int t, int count) {
  const uint32_t epoch = (uint32_t)proto.elem_offset;
  uint32_t bits = 0u;
  if (p.row != nullptr) {
    uint64_t w = p.w;
    if ((uint32_t)(w >> 32) != epoch) {
      const long long t0 = clock64();
      do {
        if (clock64() - t0 > 20000000000LL) {
          // ~10 s: the peer died.  Poison its row of this copy with NaN for this epoch, so that the CTAs still to
          // come do not each wait out the limit: the launch ends after one limit, with NaN results, instead of hanging
          w = ((uint64_t)epoch << 32) | 0x7FFFFFFFull;
          for (int i = 0; i < count; ++i) st_relaxed_sys_u64(p.row + i, w);
          break;
        }
        __nanosleep(64);
        w = ld_relaxed_sys_u64(p.row + t);
      } while ((uint32_t)(w >> 32) != epoch);
    }
    bits = (uint32_t)w;
  }
  __syncwarp();
  return __uint_as_float(warp_max_u32(bits));                  // NaN patterns sort above everything: they propagate
}

__device__ __forceinline__ float mt_exchanged_amax(const CastArgs& proto, int t, int count) {
  return mt_poll_finish(proto, mt_poll_begin(proto, t, count), t, count);
}

// Scale of tensor t from its maximum `amax`; the tensor's first CTA stores it into proto.d_scale_out[t].
template <int DT>
__device__ __forceinline__ float mt_scale_of(const CastArgs& proto, float amax, int t, bool first_cta) {
  const float s = scale_from_amax(amax, proto.flt_max, true, DT);
  if (proto.d_scale_out != nullptr && first_cta && threadIdx.x == 0) proto.d_scale_out[t] = s;
  return s;
}

template <int DT>
__device__ __forceinline__ float mt_scale(const CastArgs& proto, const float* d_scales, int t, int count, bool first_cta) {
  if (proto.d_amax == nullptr) return __ldg(d_scales + t);
  return mt_scale_of<DT>(proto, __ldg(proto.d_amax + t), t, first_cta);
}

// A tensor without elements on this rank (an empty row slab) has no CTA in the cast launch, yet its scale is part of
// the sweep's result: one CTA stores the scales of all such tensors.
template <int DT>
__global__ void __launch_bounds__(kThreads) empty_scales_kernel(const void* table, const int count, const CastArgs proto) {
  const MtEntry* e = mt_entries(table);
  for (int t = threadIdx.x >> 5; t < count; t += kThreads / 32) {         // one warp per tensor
    if (e[t].n != 0) continue;
    const float amax = proto.in != nullptr ? mt_exchanged_amax(proto, t, count) : __ldg(proto.d_amax + t);
    if ((threadIdx.x & 31) == 0) proto.d_scale_out[t] = scale_from_amax(amax, proto.flt_max, true, DT);
  }
}

template <int FMT, int RND, int DT, bool SCALED, bool ANYOFF = false, bool XCHG = false>
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
  const int64_t first = local_cta * (kThreads * U);              // this CTA's first vector of the tensor
  const int64_t rest = nvec - first;
  const int left = (int)(rest < (int64_t)(kThreads * U) ? rest : (int64_t)(kThreads * U)) - (int)threadIdx.x;
  const char* in = static_cast<const char*>(a.in) + (first + threadIdx.x) * 32;
  char* out = static_cast<char*>(a.out) + (first + threadIdx.x) * 32;
  MtPoll poll;
  if (XCHG) poll = mt_poll_begin(proto, t, count);               // exchange mode is its own kernel instance: the poll's
                                                                 // registers would cost the plain kernel a resident CTA
  Vec32 v[U];
#pragma unroll
  for (int j = 0; j < U; ++j)
    if (j * kThreads < left) v[j] = ldg256(in + (size_t)j * kThreads * 32);
  // the scale AFTER the loads have been issued: in exchange mode it may have to wait for words that peers write, and
  // that latency would otherwise sit in front of every short-lived CTA's memory traffic
  if (SCALED) {
    if (XCHG) {
      const float s = effective_scale(mt_scale_of<DT>(proto, mt_poll_finish(proto, poll, t, count), t, local_cta == 0), DT);
      sc.scale = s;
      sc.rscale = __frcp_rn(s);
    } else {
      // one warp derives the scale and its reciprocal (two correctly rounded divisions: ~35 instructions that the
      // other seven warps of an issue-bound kernel need not repeat); the barrier sits behind the loads' latency
      __shared__ float s_scale[2];
      if (threadIdx.x < 32) {
        const float s = effective_scale(mt_scale<DT>(proto, d_scales, t, count, local_cta == 0), DT);
        if (threadIdx.x == 0) { s_scale[0] = s; s_scale[1] = __frcp_rn(s); }
      }
      __syncthreads();
      sc.scale = s_scale[0];
      sc.rscale = s_scale[1];
    }
  }
#pragma unroll
  for (int j = 0; j < U; ++j) {
    if (j * kThreads < left) {
      process_vec<FMT, RND, false, DT, SCALED, ANYOFF>(v[j], sc, a,
                                                       a.elem_offset + (uint64_t)(first + threadIdx.x + j * kThreads) * VE);
      stg256(out + (size_t)j * kThreads * 32, v[j]);
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
template <int FMT, int RND, int DT, bool SCALED, bool XCHG = false>
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
    const float s = effective_scale(XCHG ? mt_scale_of<DT>(proto, mt_exchanged_amax(proto, t, count), t, local_cta == 0)
                                         : mt_scale<DT>(proto, d_scales, t, count, local_cta == 0), DT);
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

// The exchange step of a row-sharded sweep: once the per-tensor maxima are complete (stream order: a small launch
// right behind amax_multi_kernel — a completion counter inside the reduction costs one same-address atomic per CTA,
// measured slower than the extra launch), every maximum is stored together with the sweep's epoch, as one 8-byte word,
// into every rank's receive area over NVLink (plain stores to peer-mapped memory); CTA c writes copy c.  The maxima
// are accumulated in one of two local vectors, by the epoch's parity: CTA 0 re-zeroes the OTHER one (which the
// previous sweep's push has finished reading), so no sweep needs a memset.
struct MtPeers {
  uint64_t* recv[16];       // rank r's receive area for this parity, uint64[copy][world][count], as mapped into this process
  int world, rank;
  uint32_t epoch;
};

__global__ void __launch_bounds__(kThreads) push_kernel(const int count, const float* mine, float* other, const MtPeers peers) {
  const size_t base = ((size_t)blockIdx.x * peers.world + peers.rank) * (size_t)count;
  for (int i = threadIdx.x; i < count * peers.world; i += kThreads) {
    const int r = i / count, t = i - r * count;
    const uint64_t w = ((uint64_t)peers.epoch << 32) | (uint64_t)__float_as_uint(__ldcg(mine + t));
    st_relaxed_sys_u64(peers.recv[r] + base + t, w);
  }
  if (blockIdx.x == 0)
    for (int t = threadIdx.x; t < count; t += kThreads) other[t] = 0.0f;
}

typedef void (*MultiCastKernel)(const void*, const int, const CastArgs, const float*);

template <int FMT, int RND, int DT>
static MultiCastKernel mpick_scaled(bool scaled, bool anyoff, bool xchg) {
  constexpr bool kSR = (RND == FPEMU_RND_STOCHASTIC);
  if (FMT == FPEMU_FMT_BF16 || FMT == FPEMU_FMT_FP16)
    return (kSR && anyoff) ? cast_multi_kernel<FMT, RND, DT, false, kSR> : cast_multi_kernel<FMT, RND, DT, false>;
  if (xchg)                                                      // always scaled
    return (kSR && anyoff) ? cast_multi_kernel<FMT, RND, DT, true, kSR, true> : cast_multi_kernel<FMT, RND, DT, true, false, true>;
  if (kSR && anyoff) return scaled ? cast_multi_kernel<FMT, RND, DT, true, kSR> : cast_multi_kernel<FMT, RND, DT, false, kSR>;
  return scaled ? cast_multi_kernel<FMT, RND, DT, true> : cast_multi_kernel<FMT, RND, DT, false>;
}
template <int FMT, int RND>
static MultiCastKernel mpick_dt(int dt, bool scaled, bool anyoff, bool xchg) {
  switch (dt) {
    case FPEMU_DT_F32:  return mpick_scaled<FMT, RND, FPEMU_DT_F32>(scaled, anyoff, xchg);
    case FPEMU_DT_F16:  return mpick_scaled<FMT, RND, FPEMU_DT_F16>(scaled, anyoff, xchg);
    default:            return mpick_scaled<FMT, RND, FPEMU_DT_BF16>(scaled, anyoff, xchg);
  }
}
template <int FMT>
static MultiCastKernel mpick_rnd(int rnd, int dt, bool scaled, bool anyoff, bool xchg) {
  switch (rnd) {
    case FPEMU_RND_RNE:        return mpick_dt<FMT, FPEMU_RND_RNE>(dt, scaled, anyoff, xchg);
    case FPEMU_RND_RTZ:        return (FMT == FPEMU_FMT_FP16) ? nullptr : mpick_dt<FMT, FPEMU_RND_RTZ>(dt, scaled, anyoff, xchg);
    case FPEMU_RND_STOCHASTIC: return mpick_dt<FMT, FPEMU_RND_STOCHASTIC>(dt, scaled, anyoff, xchg);
    default:                   return nullptr;
  }
}

template <int FMT, int RND, int DT>
static MultiCastKernel epick_scaled(bool scaled, bool xchg) {
  if (xchg) return encode_multi_kernel<FMT, RND, DT, true, true>;
  return scaled ? encode_multi_kernel<FMT, RND, DT, true> : encode_multi_kernel<FMT, RND, DT, false>;
}
template <int FMT, int RND>
static MultiCastKernel epick_dt(int dt, bool scaled, bool xchg) {
  switch (dt) {
    case FPEMU_DT_F32:  return epick_scaled<FMT, RND, FPEMU_DT_F32>(scaled, xchg);
    case FPEMU_DT_F16:  return epick_scaled<FMT, RND, FPEMU_DT_F16>(scaled, xchg);
    default:            return epick_scaled<FMT, RND, FPEMU_DT_BF16>(scaled, xchg);
  }
}
template <int FMT>
static MultiCastKernel epick_rnd(int rnd, int dt, bool scaled, bool xchg) {
  switch (rnd) {
    case FPEMU_RND_RNE:        return epick_dt<FMT, FPEMU_RND_RNE>(dt, scaled, xchg);
    case FPEMU_RND_RTZ:        return epick_dt<FMT, FPEMU_RND_RTZ>(dt, scaled, xchg);
    case FPEMU_RND_STOCHASTIC: return epick_dt<FMT, FPEMU_RND_STOCHASTIC>(dt, scaled, xchg);
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
  bool any_offset = false, has_empty = false;
  for (int t = 0; t < count; ++t) {
    const fpemu_tensor_desc& d = descs[t];
    if (d.n < 0 || (d.n > 0 && (d.in == nullptr || d.out == nullptr))) return FPEMU_ERR_ARG;
    // the multi-tensor kernels only have the vector path: 32-byte aligned buffers (any element offset)
    if ((((uintptr_t)d.in | (uintptr_t)d.out) & 31u) != 0) return FPEMU_ERR_ALIGN;
    e[t].in = d.in; e[t].out = d.out; e[t].n = d.n; e[t].elem_offset = d.elem_offset;
    any_offset = any_offset || (d.elem_offset & 15u) != 0;
    has_empty = has_empty || d.n == 0;
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
  h->count = count; h->dtype = dtype | (any_offset ? kMtAnyOffset : 0) | (has_empty ? kMtHasEmpty : 0); h->total_cast = (uint32_t)cast_ctas; h->total_amax = (uint32_t)amax_ctas;
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
                 void* stream_, bool xchg = false, int world = 0, uint32_t epoch = 0) {
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
      case FPEMU_FMT_E5M2: k = epick_rnd<FPEMU_FMT_E5M2>(rounding, dt, scaled, xchg); break;
      case FPEMU_FMT_E4M3: k = epick_rnd<FPEMU_FMT_E4M3>(rounding, dt, scaled, xchg); break;
      case FPEMU_FMT_E3M4: k = epick_rnd<FPEMU_FMT_E3M4>(rounding, dt, scaled, xchg); break;
      default: break;                                            // only the three 8-bit formats have codes
    }
  } else {
    switch (format) {
      case FPEMU_FMT_E5M2: k = mpick_rnd<FPEMU_FMT_E5M2>(rounding, dt, scaled, anyoff, xchg); break;
      case FPEMU_FMT_E4M3: k = mpick_rnd<FPEMU_FMT_E4M3>(rounding, dt, scaled, anyoff, xchg); break;
      case FPEMU_FMT_E3M4: k = mpick_rnd<FPEMU_FMT_E3M4>(rounding, dt, scaled, anyoff, xchg); break;
      case FPEMU_FMT_BF16: k = mpick_rnd<FPEMU_FMT_BF16>(rounding, dt, false, anyoff, false); break;
      case FPEMU_FMT_FP16: k = mpick_rnd<FPEMU_FMT_FP16>(rounding, dt, false, anyoff, false); break;
      default: break;
    }
  }
  if (k == nullptr) return FPEMU_ERR_MODE;                       // rare modes: use the per-tensor entry point
  CastArgs proto;
  proto.in = xchg ? d_amax : nullptr; proto.out = nullptr; proto.n = world;       // exchange mode: see mt_scale()
  proto.scale = 1.0f; proto.d_scale = nullptr;
  proto.seed_lo = (uint32_t)seed; proto.seed_hi = (uint32_t)(seed >> 32);
  proto.elem_offset = epoch;
  proto.fmt = format; proto.rnd = rounding; proto.flags = flags;
  proto.d_amax = d_amax; proto.flt_max = flt_max; proto.d_scale_out = d_scales_out; proto.d_seed = seed_source();
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (h->total_cast > 0) k<<<h->total_cast, kThreads, 0, stream>>>(d_table, count, proto, d_scales);
  if (d_amax != nullptr && d_scales_out != nullptr && (h->dtype & kMtHasEmpty) != 0) {
    switch (dt) {
      case FPEMU_DT_F32:  empty_scales_kernel<FPEMU_DT_F32><<<1, kThreads, 0, stream>>>(d_table, count, proto); break;
      case FPEMU_DT_F16:  empty_scales_kernel<FPEMU_DT_F16><<<1, kThreads, 0, stream>>>(d_table, count, proto); break;
      default:            empty_scales_kernel<FPEMU_DT_BF16><<<1, kThreads, 0, stream>>>(d_table, count, proto); break;
    }
  }
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
// Exchange buffer of one rank, counted in floats: uint64 recv[parity 0..1][copy][source rank][count].
extern "C" int64_t fpemu_exchange_floats(int count, int world) {
  if (count <= 0 || world <= 0 || world > 16) return FPEMU_ERR_ARG;
  return 4LL * kXchgCopies * world * count;
}

extern "C" int fpemu_amax_multi_push(const void* d_table, const void* h_table, int count, float* d_amax2,
                                     void* const* peer_bufs, int world, int rank, uint32_t epoch, void* stream_) {
  if (d_table == nullptr || h_table == nullptr || d_amax2 == nullptr || peer_bufs == nullptr || count <= 0 || world <= 0 ||
      world > 16 || rank < 0 || rank >= world || epoch == 0)
    return FPEMU_ERR_ARG;
  const MtHeader* h = static_cast<const MtHeader*>(h_table);
  if (h->count != count) return FPEMU_ERR_ARG;
  MtPeers peers;
  const size_t parity = (size_t)(epoch & 1u);
  for (int r = 0; r < 16; ++r) peers.recv[r] = nullptr;
  for (int r = 0; r < world; ++r) {
    if (peer_bufs[r] == nullptr || ((uintptr_t)peer_bufs[r] & 7u) != 0) return FPEMU_ERR_ARG;
    peers.recv[r] = static_cast<uint64_t*>(peer_bufs[r]) + parity * kXchgCopies * world * count;
  }
  peers.world = world; peers.rank = rank; peers.epoch = epoch;
  float* mine = d_amax2 + parity * count;
  float* other = d_amax2 + (parity ^ 1u) * count;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (h->total_amax > 0) {                                 // a rank whose slabs are all empty still has to arrive
    switch (h->dtype & 0xFF) {
      case FPEMU_DT_F32:  amax_multi_kernel<FPEMU_DT_F32><<<h->total_amax, kThreads, 0, stream>>>(d_table, count, mine); break;
      case FPEMU_DT_F16:  amax_multi_kernel<FPEMU_DT_F16><<<h->total_amax, kThreads, 0, stream>>>(d_table, count, mine); break;
      default:            amax_multi_kernel<FPEMU_DT_BF16><<<h->total_amax, kThreads, 0, stream>>>(d_table, count, mine); break;
    }
    const cudaError_t err = cudaGetLastError();
    if (err != cudaSuccess) return (int)err;
  }
  push_kernel<<<kXchgCopies, kThreads, 0, stream>>>(count, mine, other, peers);
  return (int)cudaGetLastError();
}

static int launch_xchg(bool codes, const void* d_table, const void* h_table, int count, int format, int rounding, int flags,
                       const void* my_buf, int world, uint32_t epoch, float flt_max, float* d_scales_out, uint64_t seed,
                       void* stream_) {
  if (my_buf == nullptr || ((uintptr_t)my_buf & 7u) != 0 || world <= 0 || world > 16 || epoch == 0 || count <= 0)
    return FPEMU_ERR_ARG;
  const uint64_t* recv = static_cast<const uint64_t*>(my_buf) + (size_t)(epoch & 1u) * kXchgCopies * world * count;
  return launch_multi(codes, d_table, h_table, count, format, rounding, flags, nullptr, reinterpret_cast<const float*>(recv),
                      flt_max, d_scales_out, seed, stream_, true, world, epoch);
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
