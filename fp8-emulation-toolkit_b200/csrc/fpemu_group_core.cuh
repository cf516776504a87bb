// fpemu_group_core.cuh — pieces shared by the register-resident group kernels (fpemu_group.cu, fpemu_rows_codec.cu).
#pragma once

#include "fpemu_reduce.cuh"

namespace fpemu {

// Maximum of a non-negative float over the 2^lg threads of a group (threads of one group are contiguous; every thread
// of the CTA calls this, groups narrower than a warp use xor-shuffles only, wider ones add one pass through shared
// memory).  lg is uniform across the CTA.
__device__ __forceinline__ float group_absmax(float m, const int lg) {
  uint32_t bits = __float_as_uint(m);
  if (lg >= 5) {
    bits = __reduce_max_sync(0xFFFFFFFFu, bits);                  // one REDUX for a whole warp
  } else {
    for (int o = (1 << lg) >> 1; o > 0; o >>= 1) bits = max(bits, __shfl_xor_sync(0xFFFFFFFFu, bits, o));
  }
  if (lg > 5) {
    __shared__ uint32_t warp_bits[8];
    if ((threadIdx.x & 31) == 0) warp_bits[threadIdx.x >> 5] = bits;
    __syncthreads();
    const int wpg = 1 << (lg - 5);                                // warps per group
    const int w0 = (threadIdx.x >> lg) * wpg;
    bits = warp_bits[w0];
    for (int i = 1; i < wpg; ++i) bits = max(bits, warp_bits[w0 + i]);
  }
  return __uint_as_float(bits);
}

// Group width (log2 of the thread count, 8 ... 256 threads) and vectors per thread (4, 6 or 8) for a group of
// `vpg` vectors: the narrowest group that still fits with 4 vectors per thread, so that every thread has 2-4 vectors
// in flight; 6 or 8 vectors per thread on 256 threads for the longest groups.  false: does not fit in registers.
inline bool group_shape(int64_t vpg, int* lg, int* r) {
  if (vpg < 1 || vpg > 256 * 8) return false;
  if (vpg > 256 * 6) { *lg = 8; *r = 8; return true; }
  if (vpg > 256 * 4) { *lg = 8; *r = 6; return true; }
  int l = 3;
  while ((4LL << l) < vpg) ++l;
  *lg = l; *r = 4;
  return true;
}

// The same choice with 2 and 3 vectors per thread allowed as well (fpemu_group.cu has those instantiations): the shape
// with the fewest idle vector slots, ties to the narrower group.  Rows whose length is not a power of two
// (576, 4608, 11 008 ...) otherwise leave up to half of the slots of a 4-per-thread shape empty, and the registers those
// slots hold are what limits the resident warps.
inline bool group_shape_tight(int64_t vpg, int* lg, int* r) {
  if (vpg < 1 || vpg > 256 * 8) return false;
  static const int kR[] = {2, 3, 4, 6, 8};
  int64_t best_slots = -1;
  for (int l = 3; l <= 8; ++l) {
    for (int ri = 0; ri < 5; ++ri) {
      const int64_t slots = (int64_t)kR[ri] << l;
      if (slots < vpg) continue;
      if (kR[ri] > 4 && l < 8) continue;                         // 6 / 8 per thread only when 256 threads are not enough
      if (best_slots < 0 || slots < best_slots) { best_slots = slots; *lg = l; *r = kR[ri]; }
      break;                                                     // larger R at this width only adds idle slots
    }
  }
  return best_slots >= 0;
}

}  // namespace fpemu
