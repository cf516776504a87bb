// fpemu_group.cu — register-resident "group" casts: a group is one row of a row-wise (per-channel) cast or
// one quantization block of a blocked cast.  G threads (8 ... 256, a power of two chosen per launch so that a
// thread holds 2-4 vectors) load their group ONCE into registers (up to R 32-byte vectors per thread), reduce |x|
// with xor-shuffles (+ shared memory when a group spans several warps), derive the scale and cast straight from
// the registers: 8 B/elem (fp32) of DRAM traffic, no second pass.  Capacity: 256 * 8 vectors per group = 16 384
// fp32 / 32 768 16-bit elements; larger groups fall back to the two-pass kernels in fpemu_scaled.cu.
#include "fpemu_group_core.cuh"

namespace fpemu {

enum { kRuleRowClamp = 0, kRuleRowNoClamp = 1, kRuleBlockPow2 = 2 };

template <int FMT, int RND, int DT, int R>
__global__ void __launch_bounds__(256) group_reg_kernel(const CastArgs a_, const int64_t groups, const int vpg, const int lg,
                                                       const int rule, const float flt_max, float* d_scale_out) {
  const CastArgs a = with_device_seed(a_);
  constexpr int VE = DtInfo<DT>::kVecElems;
  const int G = 1 << lg;
  const int tg = threadIdx.x & (G - 1);
  const int64_t grp = (int64_t)blockIdx.x * (256 >> lg) + (threadIdx.x >> lg);
  const bool live = grp < groups;
  const int64_t v0 = grp * vpg;                                  // first vector of this group
  const char* in = static_cast<const char*>(a.in);
  char* out = static_cast<char*>(a.out);
  Vec32 v[R];
  float m = 0.0f;
#pragma unroll
  for (int j = 0; j < R; ++j) {
    const int vi = j * G + tg;
    if (live && vi < vpg) v[j] = ldg256(in + (v0 + vi) * 32);
    else {
#pragma unroll
      for (int i = 0; i < 8; ++i) v[j].w[i] = 0u;
    }
  }
  // rows follow torch.max(abs(row)) (a NaN makes the row's scale NaN, qutils.py:303), blocks follow absmax_block
  // (NaN ignored, fpemu_kernels.cu:150-170); `rule` is uniform across the launch
  if (rule == kRuleBlockPow2) {
#pragma unroll
    for (int j = 0; j < R; ++j) m = fmaxf(m, vec_absmax<DT, false>(v[j]));
  } else {
#pragma unroll
    for (int j = 0; j < R; ++j) m = absmax_nan(m, vec_absmax<DT, true>(v[j]));
  }
  const float amax = group_absmax(m, lg);
  float s;
  if (rule == kRuleBlockPow2) {
    s = block_scale(amax, FMT, DT);
  } else {
    const float s_host = scale_from_amax(amax, flt_max, rule == kRuleRowClamp, DT);
    if (d_scale_out != nullptr && live && tg == 0) d_scale_out[grp] = s_host;
    s = effective_scale(s_host, DT);
  }
  ScaleCtx sc;
  sc.scale = s;
  sc.rscale = __frcp_rn(s);
  sc.unscale32 = (a.flags & FPEMU_FLAG_UNSCALE_FP32) != 0;
#pragma unroll
  for (int j = 0; j < R; ++j) {
    const int vi = j * G + tg;
    if (live && vi < vpg) {
      process_vec<FMT, RND, false, DT, true>(v[j], sc, a, a.elem_offset + (uint64_t)(v0 + vi) * VE);
      stg256(out + (v0 + vi) * 32, v[j]);
    }
  }
}

// Rows whose vector count fits no power-of-two group shape well (576 bf16 columns = 36 vectors: the tightest shape,
// 16 threads x 3, leaves a quarter of its slots — and of the issue slots of an instruction-bound cast — empty).
// Here a CTA takes `rows_per_cta` WHOLE consecutive rows, which are one contiguous run of vectors, and deals the
// vectors to its threads without regard to the row boundaries: thread t holds vectors t, t + T, ... (R of them), so
// that only the last few slots of the CTA are idle (36 vectors: 14 rows = 504 of 512 slots).  The row maxima are
// collected in shared memory — per warp and slot one REDUX per distinct row (at most two when a row has 32 vectors or
// more) and one shared atomicMax by its first lane — the first `rows` threads turn them into scales, and every thread
// picks up the scale of each of its vectors' rows.
constexpr int kTileMaxRows = 64;

template <int FMT, int RND, int DT, int R>
__global__ void __launch_bounds__(256) rows_tile_kernel(const CastArgs a_, const int64_t rows, const int vpg,
                                                       const int rows_per_cta, const uint32_t inv, const int rule,
                                                       const float flt_max, float* d_scale_out) {
  const CastArgs a = with_device_seed(a_);
  constexpr int VE = DtInfo<DT>::kVecElems;
  __shared__ uint32_t row_bits[kTileMaxRows];
  __shared__ float2 row_scale[kTileMaxRows];                   // (scale, 1 / scale) per row
  const int T = blockDim.x;
  const int64_t row0 = (int64_t)blockIdx.x * rows_per_cta;
  const int nrows = (int)((rows - row0) < (int64_t)rows_per_cta ? (rows - row0) : (int64_t)rows_per_cta);
  const int nv = nrows * vpg;
  const int64_t v0 = row0 * vpg;
  const char* in = static_cast<const char*>(a.in) + v0 * 32;
  char* out = static_cast<char*>(a.out) + v0 * 32;
  if ((int)threadIdx.x < kTileMaxRows) row_bits[threadIdx.x] = 0u;
  Vec32 v[R];
  int row_of[R];
#pragma unroll
  for (int j = 0; j < R; ++j) {
    const int vi = j * T + threadIdx.x;
    row_of[j] = (int)(((uint32_t)vi * inv) >> 24);                         // vi / vpg (host-checked magic)
    if (vi < nv) v[j] = ldg256(in + vi * 32);
    else {
#pragma unroll
      for (int i = 0; i < 8; ++i) v[j].w[i] = 0u;
    }
  }
  __syncthreads();
  const int lane = threadIdx.x & 31;
#pragma unroll
  for (int j = 0; j < R; ++j) {
    const bool active = j * T + (int)threadIdx.x < nv;
    const int r = row_of[j];
    const uint32_t bits = __float_as_uint(vec_absmax<DT, true>(v[j]));     // a NaN sorts above every number
    uint32_t todo = __ballot_sync(0xFFFFFFFFu, active);
    while (todo != 0u) {
      const int leader = __ffs(todo) - 1;
      const int rl = __shfl_sync(0xFFFFFFFFu, r, leader);
      const bool mine = active && r == rl;
      const uint32_t m = __reduce_max_sync(0xFFFFFFFFu, mine ? bits : 0u);
      if (lane == leader && m > row_bits[rl]) atomicMax(&row_bits[rl], m);
      todo &= ~__ballot_sync(0xFFFFFFFFu, mine);
    }
  }
  __syncthreads();
  if ((int)threadIdx.x < nrows) {
    const float s_host = scale_from_amax(__uint_as_float(row_bits[threadIdx.x]), flt_max, rule == kRuleRowClamp, DT);
    if (d_scale_out != nullptr) d_scale_out[row0 + threadIdx.x] = s_host;
    const float s = effective_scale(s_host, DT);
    row_scale[threadIdx.x] = make_float2(s, __frcp_rn(s));
  }
  __syncthreads();
  ScaleCtx sc;
  sc.unscale32 = (a.flags & FPEMU_FLAG_UNSCALE_FP32) != 0;
#pragma unroll
  for (int j = 0; j < R; ++j) {
    const int vi = j * T + threadIdx.x;
    if (vi < nv) {
      const float2 s2 = row_scale[row_of[j]];
      sc.scale = s2.x;
      sc.rscale = s2.y;
      process_vec<FMT, RND, false, DT, true>(v[j], sc, a, a.elem_offset + (uint64_t)(v0 + vi) * VE);
      stg256(out + vi * 32, v[j]);
    }
  }
}

typedef void (*TileKernel)(const CastArgs, const int64_t, const int, const int, const uint32_t, const int, const float, float*);

template <int FMT, int RND, int DT>
static TileKernel tile_g(int r) {
  if (r == 2) return rows_tile_kernel<FMT, RND, DT, 2>;
  if (r == 3) return rows_tile_kernel<FMT, RND, DT, 3>;
  return rows_tile_kernel<FMT, RND, DT, 4>;
}
template <int FMT, int RND>
static TileKernel tile_dt(int dt, int r) {
  switch (dt) {
    case FPEMU_DT_F32:  return tile_g<FMT, RND, FPEMU_DT_F32>(r);
    case FPEMU_DT_F16:  return tile_g<FMT, RND, FPEMU_DT_F16>(r);
    default:            return tile_g<FMT, RND, FPEMU_DT_BF16>(r);
  }
}
template <int FMT>
static TileKernel tile_rnd(int rnd, int dt, int r) {
  switch (rnd) {
    case FPEMU_RND_RNE:        return tile_dt<FMT, FPEMU_RND_RNE>(dt, r);
    case FPEMU_RND_RTZ:        return tile_dt<FMT, FPEMU_RND_RTZ>(dt, r);
    case FPEMU_RND_STOCHASTIC: return tile_dt<FMT, FPEMU_RND_STOCHASTIC>(dt, r);
    default:                   return nullptr;
  }
}

// Threads per CTA (a multiple of 32), vectors per thread (2-4) and whole rows per CTA: among the shapes within 2 % of
// the fewest idle slots, the smallest CTA (then the fewest vectors per thread) — at equal slot use small CTAs measured
// faster (4608 bf16 columns: 96 x 3 = one row per CTA 6.15 TB/s, 192 x 3 = two rows 5.87 TB/s), the three CTA-wide
// barriers wait on fewer warps.  Returns the fraction of slots in use (0: rows of this length do not fit).
static double tile_shape(int64_t vpg, int* threads, int* r, int* rows_per_cta) {
  if (vpg < 16 || vpg > 256 * 4) return 0.0;
  double best = 0.0;
  for (int pass = 0; pass < 2; ++pass) {
    int64_t best_slots = -1;
    for (int rr = 2; rr <= 4; ++rr) {
      for (int t = 64; t <= 256; t += 32) {
        const int64_t slots = (int64_t)t * rr;
        if (slots < vpg) continue;
        int64_t rpc = slots / vpg;
        if (rpc > kTileMaxRows) rpc = kTileMaxRows;
        const double eff = (double)(rpc * vpg) / (double)slots;
        if (pass == 0) {
          if (eff > best) best = eff;
        } else if (eff >= best - 0.02 && (best_slots < 0 || slots < best_slots)) {
          best_slots = slots; *threads = t; *r = rr; *rows_per_cta = (int)rpc;
        }
      }
    }
  }
  return best - 0.02 > 0.0 ? (double)((int64_t)*rows_per_cta * vpg) / (double)((int64_t)*threads * *r) : 0.0;
}

typedef void (*GroupKernel)(const CastArgs, const int64_t, const int, const int, const int, const float, float*);

template <int FMT, int RND, int DT>
static GroupKernel pick_g(int r) {
  if (r == 2) return group_reg_kernel<FMT, RND, DT, 2>;
  if (r == 3) return group_reg_kernel<FMT, RND, DT, 3>;
  if (r == 4) return group_reg_kernel<FMT, RND, DT, 4>;
  if (r == 6) return group_reg_kernel<FMT, RND, DT, 6>;
  return group_reg_kernel<FMT, RND, DT, 8>;
}
template <int FMT, int RND>
static GroupKernel pick_dt(int dt, int r) {
  switch (dt) {
    case FPEMU_DT_F32:  return pick_g<FMT, RND, FPEMU_DT_F32>(r);
    case FPEMU_DT_F16:  return pick_g<FMT, RND, FPEMU_DT_F16>(r);
    default:            return pick_g<FMT, RND, FPEMU_DT_BF16>(r);
  }
}
template <int FMT>
static GroupKernel pick_rnd(int rnd, int dt, int r) {
  switch (rnd) {
    case FPEMU_RND_RNE:        return pick_dt<FMT, FPEMU_RND_RNE>(dt, r);
    case FPEMU_RND_RTZ:        return pick_dt<FMT, FPEMU_RND_RTZ>(dt, r);
    case FPEMU_RND_STOCHASTIC: return pick_dt<FMT, FPEMU_RND_STOCHASTIC>(dt, r);
    default:                   return nullptr;
  }
}

// Launches the register-resident kernel when the shape allows it; returns false when the caller must use
// another path (group too large, unsupported mode).  `rule`: 0 row-wise with clamp, 1 without, 2 block pow-2.
bool launch_group_reg(const CastArgs& a, int dtype, int64_t groups, int64_t vectors_per_group, int rule, float flt_max,
                      float* d_scale_out, cudaStream_t stream, int* status) {
  int lg, r;
  // rows only: the tile kernel when the best power-of-two shape would leave more than a tenth of its slots empty and
  // the tile shape does better
  if (rule != kRuleBlockPow2) {
    int tt = 0, tr = 0, rpc = 0;
    const double tile_eff = tile_shape(vectors_per_group, &tt, &tr, &rpc);
    double pow2_eff = 0.0;
    if (group_shape_tight(vectors_per_group, &lg, &r)) pow2_eff = (double)vectors_per_group / (double)((int64_t)r << lg);
    if (tile_eff > 0.0 && pow2_eff < 0.9 && tile_eff > pow2_eff + 0.04) {
      TileKernel tk = nullptr;
      switch (a.fmt) {
        case FPEMU_FMT_E5M2: tk = tile_rnd<FPEMU_FMT_E5M2>(a.rnd, dtype, tr); break;
        case FPEMU_FMT_E4M3: tk = tile_rnd<FPEMU_FMT_E4M3>(a.rnd, dtype, tr); break;
        case FPEMU_FMT_E3M4: tk = tile_rnd<FPEMU_FMT_E3M4>(a.rnd, dtype, tr); break;
        default: break;
      }
      const int64_t grid = (groups + rpc - 1) / rpc;
      if (tk != nullptr && grid <= 0x7FFFFFFFLL) {
        const uint32_t inv = (uint32_t)((1u << 24) / (uint32_t)vectors_per_group) + 1u;   // exact for vi < 1024, vpg >= 16
        tk<<<(unsigned)grid, tt, 0, stream>>>(a, groups, (int)vectors_per_group, rpc, inv, rule, flt_max, d_scale_out);
        *status = (int)cudaGetLastError();
        return true;
      }
    }
  }
  if (!group_shape_tight(vectors_per_group, &lg, &r)) return false;
  GroupKernel k = nullptr;
  switch (a.fmt) {
    case FPEMU_FMT_E5M2: k = pick_rnd<FPEMU_FMT_E5M2>(a.rnd, dtype, r); break;
    case FPEMU_FMT_E4M3: k = pick_rnd<FPEMU_FMT_E4M3>(a.rnd, dtype, r); break;
    case FPEMU_FMT_E3M4: k = pick_rnd<FPEMU_FMT_E3M4>(a.rnd, dtype, r); break;
    default: break;
  }
  if (k == nullptr) return false;
  const int64_t per_cta = 256 >> lg;
  const int64_t grid = (groups + per_cta - 1) / per_cta;
  if (grid > 0x7FFFFFFFLL) return false;
  k<<<(unsigned)grid, 256, 0, stream>>>(a, groups, (int)vectors_per_group, lg, rule, flt_max, d_scale_out);
  *status = (int)cudaGetLastError();
  return true;
}

}  // namespace fpemu
