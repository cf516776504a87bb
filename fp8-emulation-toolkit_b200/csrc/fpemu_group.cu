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
