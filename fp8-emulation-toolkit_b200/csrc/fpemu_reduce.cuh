// fpemu_reduce.cuh — absolute-maximum and scale-derivation device helpers shared by the scaled kernels.
#pragma once

#include "fpemu_cast_core.cuh"

namespace fpemu {

// ------------------------------------------------------------------------------------------ amax
// max |x| of one 32-byte vector as a non-negative float.
//   NANPROP = true : a NaN lane makes the result NaN, like torch.max(abs(t)) — the per-tensor / per-channel scale
//                    derivations of mpemu/qutils.py:283, :303, :327 (the scale, and with it every output, is then NaN);
//   NANPROP = false: NaN lanes are ignored (fmaxf semantics) — the reference's absmax_block of the blocked mode,
//                    fpemu_kernels.cu:150-170.
template <int DT, bool NANPROP>
__device__ __forceinline__ float vec_absmax(const Vec32& v) {
  if (DT == FPEMU_DT_F32) {
    float m = 0.0f;
#pragma unroll
    for (int i = 0; i < 8; ++i) m = NANPROP ? absmax_nan(m, __uint_as_float(v.w[i])) : fmaxf(m, fabsf(__uint_as_float(v.w[i])));
    return m;
  } else if (DT == FPEMU_DT_F16) {
    __half2 m = as_h2(0u);
#pragma unroll
    for (int i = 0; i < 8; ++i) m = NANPROP ? __hmax2_nan(m, __habs2(as_h2(v.w[i]))) : __hmax2(m, __habs2(as_h2(v.w[i])));   // |x| as an operand modifier
    return NANPROP ? absmax_nan(__low2float(m), __high2float(m)) : fmaxf(__low2float(m), __high2float(m));
  } else {
    uint32_t z = 0u;
    __nv_bfloat162 m = *reinterpret_cast<__nv_bfloat162*>(&z);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      uint32_t w = v.w[i];
      const __nv_bfloat162 a = __habs2(*reinterpret_cast<__nv_bfloat162*>(&w));
      m = NANPROP ? __hmax2_nan(m, a) : __hmax2(m, a);
    }
    return NANPROP ? absmax_nan(__low2float(m), __high2float(m)) : fmaxf(__low2float(m), __high2float(m));
  }
}

// Power-of-two block scale from the block maximum (fpemu_kernels.cu:409-413, 577-581, 909-911, 998).
__device__ __forceinline__ float block_scale(float maxval, int fmt, int dt) {
  const float p2 = __uint_as_float(__float_as_uint(maxval) & 0x7F800000u);
  switch (fmt) {
    case FPEMU_FMT_E5M2:
      if (dt == FPEMU_DT_F16) {                              // `scalar_t scale` is a half in this instantiation
        float s = __half2float(__float2half_rn(2.0f * p2));
        return __half2float(__float2half_rn(s * (1.0f / 16384.0f)));
      }
      return (2.0f * p2) * (1.0f / 16384.0f);
    case FPEMU_FMT_E4M3:
    case FPEMU_FMT_E4M3_IEEE: return (2.0f * p2) * 0.125f;
    case FPEMU_FMT_E3M4:      return 2.0f * p2;
    default:                  return __fdiv_rn(1.0f, maxval);    // FP4: 1 / max
  }
}

}  // namespace fpemu
