// fpemu_group.cu — register-resident "group" casts: a group is one row of a row-wise (per-channel) cast or
// one quantization block of a blocked cast.  A warp (G = 32) or a whole CTA (G = 256) loads its group ONCE into
// registers (up to R 32-byte vectors per thread), reduces |x| with warp shuffles (+ shared memory across the
// 8 warps of a CTA), derives the scale and casts straight from the registers: 8 B/elem (fp32) of DRAM
// traffic, no second pass.  Capacity: G * R vectors per group = 1024 (warp) / 12 288 (CTA) fp32 elements,
// twice that for 16-bit tensors; larger groups fall back to the two-pass kernels in fpemu_scaled.cu.
#include "fpemu_reduce.cuh"

namespace fpemu {

enum { kRuleRowClamp = 0, kRuleRowNoClamp = 1, kRuleBlockPow2 = 2 };

template <int FMT, int RND, int DT, int G, int R>
__global__ void __launch_bounds__(256) group_reg_kernel(const CastArgs a, const int64_t groups, const int vpg,
                                                       const int rule, const float flt_max, float* d_scale_out) {
  constexpr int VE = DtInfo<DT>::kVecElems;
  constexpr int kGroupsPerCta = 256 / G;
  const int tg = threadIdx.x % G;
  const int64_t grp = (int64_t)blockIdx.x * kGroupsPerCta + threadIdx.x / G;
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
#pragma unroll
  for (int j = 0; j < R; ++j) m = fmaxf(m, vec_absmax<DT>(v[j]));
  uint32_t bits = warp_max_u32(__float_as_uint(m));
  if (G > 32) {
    __shared__ uint32_t warp_bits[8];
    if ((threadIdx.x & 31) == 0) warp_bits[threadIdx.x >> 5] = bits;
    __syncthreads();
    bits = warp_bits[0];
#pragma unroll
    for (int w = 1; w < 8; ++w) bits = max(bits, warp_bits[w]);
  }
  const float amax = __uint_as_float(bits);
  float s;
  if (rule == kRuleBlockPow2) {
    s = block_scale(amax, FMT, DT);
  } else {
    const float s_host = scale_from_amax(amax, flt_max, rule == kRuleRowClamp);
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

typedef void (*GroupKernel)(const CastArgs, const int64_t, const int, const int, const float, float*);

constexpr int kWarpR = 4;    // 32 * 4 = 128 vectors per group
constexpr int kCtaSmallR = 2;  // 256 * 2 = 512 vectors per group (fewer registers, more CTAs per SM)
constexpr int kCtaR = 6;     // 256 * 6 = 1536 vectors per group

template <int FMT, int RND, int DT>
static GroupKernel pick_g(int cta) {
  if (cta == 0) return group_reg_kernel<FMT, RND, DT, 32, kWarpR>;
  if (cta == 1) return group_reg_kernel<FMT, RND, DT, 256, kCtaSmallR>;
  return group_reg_kernel<FMT, RND, DT, 256, kCtaR>;
}
template <int FMT, int RND>
static GroupKernel pick_dt(int dt, int cta) {
  switch (dt) {
    case FPEMU_DT_F32:  return pick_g<FMT, RND, FPEMU_DT_F32>(cta);
    case FPEMU_DT_F16:  return pick_g<FMT, RND, FPEMU_DT_F16>(cta);
    default:            return pick_g<FMT, RND, FPEMU_DT_BF16>(cta);
  }
}
template <int FMT>
static GroupKernel pick_rnd(int rnd, int dt, int cta) {
  switch (rnd) {
    case FPEMU_RND_RNE:        return pick_dt<FMT, FPEMU_RND_RNE>(dt, cta);
    case FPEMU_RND_RTZ:        return pick_dt<FMT, FPEMU_RND_RTZ>(dt, cta);
    case FPEMU_RND_STOCHASTIC: return pick_dt<FMT, FPEMU_RND_STOCHASTIC>(dt, cta);
    default:                   return nullptr;
  }
}

// Launches the register-resident kernel when the shape allows it; returns false when the caller must use
// another path (group too large, unsupported mode).  `rule`: 0 row-wise with clamp, 1 without, 2 block pow-2.
bool launch_group_reg(const CastArgs& a, int dtype, int64_t groups, int64_t vectors_per_group, int rule, float flt_max,
                      float* d_scale_out, cudaStream_t stream, int* status) {
  if (vectors_per_group < 1 || vectors_per_group > 256 * kCtaR) return false;
  const int cta = vectors_per_group <= 32 * kWarpR ? 0 : (vectors_per_group <= 256 * kCtaSmallR ? 1 : 2);
  GroupKernel k = nullptr;
  switch (a.fmt) {
    case FPEMU_FMT_E5M2: k = pick_rnd<FPEMU_FMT_E5M2>(a.rnd, dtype, cta); break;
    case FPEMU_FMT_E4M3: k = pick_rnd<FPEMU_FMT_E4M3>(a.rnd, dtype, cta); break;
    case FPEMU_FMT_E3M4: k = pick_rnd<FPEMU_FMT_E3M4>(a.rnd, dtype, cta); break;
    default: break;
  }
  if (k == nullptr) return false;
  const int64_t per_cta = cta ? 1 : 8;   // a 256-thread CTA holds 8 warp-sized groups or one CTA-sized group
  const int64_t grid = (groups + per_cta - 1) / per_cta;
  if (grid > 0x7FFFFFFFLL) return false;
  k<<<(unsigned)grid, 256, 0, stream>>>(a, groups, (int)vectors_per_group, rule, flt_max, d_scale_out);
  *status = (int)cudaGetLastError();
  return true;
}

}  // namespace fpemu
