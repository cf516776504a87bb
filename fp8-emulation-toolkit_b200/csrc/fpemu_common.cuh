// fpemu_common.cuh — vector I/O, launch helpers and argument checks shared by the kernels.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "fpemu_math.cuh"

namespace fpemu {

// One 256-bit global transaction per thread (LDG.E.256 / STG.E.256 on sm_100a): 8 fp32 or 16 fp16/bf16.
struct alignas(32) Vec32 { uint32_t w[8]; };

// Streaming accesses: every byte is touched exactly once, so do not allocate in L1.  Plain (coherent)
// ld.global rather than .nc because the in-place variant reads the buffer it writes.
__device__ __forceinline__ Vec32 ldg256(const void* p) {
  Vec32 r;
  asm volatile("ld.global.L1::no_allocate.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(r.w[0]), "=r"(r.w[1]), "=r"(r.w[2]), "=r"(r.w[3]), "=r"(r.w[4]), "=r"(r.w[5]), "=r"(r.w[6]), "=r"(r.w[7])
               : "l"(p));
  return r;
}
__device__ __forceinline__ void stg256(void* p, const Vec32& r) {
  asm volatile("st.global.L1::no_allocate.v8.u32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
               :: "l"(p), "r"(r.w[0]), "r"(r.w[1]), "r"(r.w[2]), "r"(r.w[3]), "r"(r.w[4]), "r"(r.w[5]), "r"(r.w[6]), "r"(r.w[7])
               : "memory");
}

template <int DT> struct DtInfo;
template <> struct DtInfo<FPEMU_DT_F32>  { static constexpr int kBytes = 4; static constexpr int kVecElems = 8; };
template <> struct DtInfo<FPEMU_DT_F16>  { static constexpr int kBytes = 2; static constexpr int kVecElems = 16; };
template <> struct DtInfo<FPEMU_DT_BF16> { static constexpr int kBytes = 2; static constexpr int kVecElems = 16; };

__host__ __device__ inline int dtype_bytes(int dt) { return dt == FPEMU_DT_F32 ? 4 : 2; }

// Scalar element access for the generic / tail paths.
__device__ __forceinline__ float load_elem(const void* p, int64_t i, int dt) {
  if (dt == FPEMU_DT_F32) return reinterpret_cast<const float*>(p)[i];
  const unsigned short b = reinterpret_cast<const unsigned short*>(p)[i];
  if (dt == FPEMU_DT_F16) return __half2float(__ushort_as_half(b));
  return __uint_as_float((uint32_t)b << 16);
}
__device__ __forceinline__ void store_elem(void* p, int64_t i, int dt, float v) {
  if (dt == FPEMU_DT_F32) { reinterpret_cast<float*>(p)[i] = v; return; }
  unsigned short b;
  if (dt == FPEMU_DT_F16) b = __half_as_ushort(__float2half_rn(v));
  else b = __bfloat16_as_ushort(__float2bfloat16_rn(v));
  reinterpret_cast<unsigned short*>(p)[i] = b;
}

// The scale a kernel instantiated for an fp16 tensor sees is rounded to half first
// (`const scalar_t in_scale`, reference fpemu_kernels.cu:379, 469, 547, 877).
__device__ __forceinline__ float effective_scale(float s, int dt) {
  return dt == FPEMU_DT_F16 ? __half2float(__float2half_rn(s)) : s;
}

// warp / block max of |x| values held as bit patterns: for non-negative floats the integer order of the encodings is
// the numeric order, and every NaN encoding sorts above +Inf, so a NaN that reached the pattern stays the maximum
__device__ __forceinline__ uint32_t warp_max_u32(uint32_t v) {
  return __reduce_max_sync(0xFFFFFFFFu, v);
}

// max(m, |x|) that returns NaN when either operand is NaN (what torch.max(abs(t)) does, mpemu/qutils.py:283) — one
// FMNMX.NAN with the |.| source modifier
__device__ __forceinline__ float absmax_nan(float m, float x) {
  float r;
  asm("max.NaN.f32 %0, %1, %2;" : "=f"(r) : "f"(m), "f"(fabsf(x)));
  return r;
}

// The amax-derived scale of quantize_tensor (mpemu/qutils.py:283-285, :303-305, :328):
//     scale = flt_max / vmax;  scale = 6.55e+04 if scale > 3.275e+04 else scale
// `vmax` is a 0-dim tensor there, so `flt_max / vmax` is Tensor.__rtruediv__ = `vmax.reciprocal() * flt_max`
// (torch/_tensor.py `__rdiv__`), evaluated in the tensor's dtype: two roundings, not one division.
//   float32 tensors: fl32(fl32(1 / vmax) * flt_max);
//   float16 tensors: the reciprocal is rounded to half, the product again, and the comparison with 3.275e4 runs in
//                    half (3.275e4 -> 32752); the clamp value 6.55e4 reaches the kernel as a float and is rounded to
//                    half there (65504) by effective_scale();
//   bfloat16 tensors (not dispatched by the reference): float32 arithmetic on the widened values.
__device__ __forceinline__ float scale_from_amax(float amax, float flt_max, bool clamp, int dt) {
  if (dt == FPEMU_DT_F16) {
    const float r = __half2float(__float2half_rn(__frcp_rn(amax)));
    const float s = __half2float(__float2half_rn(__fmul_rn(r, flt_max)));
    return (clamp && s > 32752.0f) ? 6.55e4f : s;
  }
  const float s = __fmul_rn(__frcp_rn(amax), flt_max);
  return (clamp && s > 3.275e4f) ? 6.55e4f : s;
}

struct DeviceInfo { int sm_count; };
const DeviceInfo& device_info();   // cached per current device

}  // namespace fpemu
