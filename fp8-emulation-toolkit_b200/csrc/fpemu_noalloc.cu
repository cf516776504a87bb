// fpemu_noalloc.cu — fpemu_cast_amax_scaled_noalloc: the per-tensor amax-scaled cast for callers that
// cannot give this library any scratch memory.
//
// Who needs this.  The regular amax-scaled path (fpemu_amax -> fpemu_scale_from_amax -> fpemu_cast) wants two
// device floats from the caller for the running maximum and the derived scale.  Some callers have no way to
// provide them: a plugin loaded into a server whose memory pool was sized once at start-up and then handed
// over completely to the host framework; a C or Fortran code that received its two device buffers from
// someone else and owns no allocator at all; a real-time loop in which cudaMalloc is off limits because it
// synchronises the device and may fail at a moment nobody can handle.  For them the scratch lives inside the
// library image: a file-scope `static __device__` array that every GPU loading the library holds from load
// until process exit (65 792 bytes per GPU, see kScratchWords).
//
// How it works.  Three launches on the caller's stream, nothing else:
//   1. partial_amax_kernel: CTA b reduces its 32 KiB slice of the input and stores the maximum in slot b;
//   2. finish_scale_kernel: one CTA reduces the slots that step 1 wrote and stores amax and
//      scale = clamp(flt_max / amax) (mpemu/qutils.py:283-285) behind them;
//   3. the ordinary fpemu_cast, reading the scale from the scratch (its device-scalar argument).
// Because there is one slot per CTA the input size is bounded: kMaxParts slices of 32 KiB = 512 MiB
// (2^27 float32 or 2^28 float16 / bfloat16 elements).  Larger tensors are rejected with
// FPEMU_ERR_TOO_LARGE; cast them in pieces with the caller-scratch entry points instead.
//
// Sharing.  All calls on one GPU use the same array and it keeps whatever the previous call left.  Every
// slot that steps 2 and 3 read has been written by step 1 / 2 of the SAME call first, so nothing has to be
// cleared and stale slots of an earlier, larger call are never looked at.  Two calls that overlap on the GPU
// (different streams, or different host threads on different streams) race on the array and produce garbage:
// serialise them on one stream or with events.
//
// Set-up.  Host code cannot use the array's name as a pointer (in host code the name is only a handle for
// the runtime), so the first call on each device asks the runtime for the real device address with
// cudaGetSymbolAddress while that device is current, checks the result and caches it.  Make that first call
// before any stream capture begins.
#include <mutex>

#include "fpemu_common.cuh"

namespace {

using namespace fpemu;

constexpr int kPartThreads = 256;
constexpr int kPartUnroll = 4;                                  // 1024 vectors = 32 KiB of input per CTA
constexpr int kSliceBytes = kPartThreads * kPartUnroll * 32;
constexpr int kMaxParts = 16384;                                // x 32 KiB = 512 MiB of input per call
constexpr int kSlotAmax = kMaxParts;                            // word index of the reduced maximum
constexpr int kSlotScale = kMaxParts + 1;                       // word index of the scale (a float)
constexpr int kScratchWords = kMaxParts + 64;                   // 65 792 bytes

static __device__ __align__(256) uint32_t g_noalloc_scratch[kScratchWords];

// |x| maximum of one CTA's 32 KiB slice -> parts[blockIdx.x].  A NaN in the slice makes the slot NaN, as in fpemu_amax.
__global__ void __launch_bounds__(kPartThreads) partial_amax_kernel(const void* in_, int64_t n, int dt, int aligned,
                                                                  uint32_t* parts) {
  const int esz = dtype_bytes(dt);
  const int64_t e_per_cta = kSliceBytes / esz;
  const int64_t e0 = (int64_t)blockIdx.x * e_per_cta;
  int64_t e1 = e0 + e_per_cta;
  if (e1 > n) e1 = n;
  float m = 0.0f;
  if (aligned) {
    const int ve = 32 / esz;
    const char* in = static_cast<const char*>(in_);
    const int64_t v0 = e0 / ve, v1 = e1 / ve;                   // whole vectors of this slice
    for (int64_t vi = v0 + threadIdx.x; vi < v1; vi += kPartThreads) {
      const Vec32 v = ldg256(in + vi * 32);
      if (dt == FPEMU_DT_F32) {
#pragma unroll
        for (int i = 0; i < 8; ++i) m = absmax_nan(m, __uint_as_float(v.w[i]));
      } else if (dt == FPEMU_DT_F16) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float2 f = unpack_f16x2(v.w[i]);
          m = absmax_nan(absmax_nan(m, f.x), f.y);
        }
      } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float2 f = unpack_bf16x2(v.w[i]);
          m = absmax_nan(absmax_nan(m, f.x), f.y);
        }
      }
    }
    for (int64_t i = v1 * ve + threadIdx.x; i < e1; i += kPartThreads) m = absmax_nan(m, load_elem(in_, i, dt));
  } else {
    for (int64_t i = e0 + threadIdx.x; i < e1; i += kPartThreads) m = absmax_nan(m, load_elem(in_, i, dt));
  }
  uint32_t bits = warp_max_u32(__float_as_uint(m));             // non-negative floats order like their bits
  __shared__ uint32_t warp_bits[kPartThreads / 32];
  if ((threadIdx.x & 31) == 0) warp_bits[threadIdx.x >> 5] = bits;
  __syncthreads();
  if (threadIdx.x < 32) {
    bits = threadIdx.x < kPartThreads / 32 ? warp_bits[threadIdx.x] : 0u;
    bits = warp_max_u32(bits);
    if (threadIdx.x == 0) parts[blockIdx.x] = bits;             // always written: no slot needs clearing
  }
}

// One CTA: maximum of parts[0 .. nparts) -> scratch[kSlotAmax]; clamp(flt_max / amax) -> scratch[kSlotScale].
__global__ void __launch_bounds__(1024) finish_scale_kernel(uint32_t* scratch, int nparts, float flt_max, int dt) {
  uint32_t bits = 0u;
  for (int i = threadIdx.x; i < nparts; i += 1024) bits = max(bits, scratch[i]);
  bits = warp_max_u32(bits);
  __shared__ uint32_t warp_bits[32];
  if ((threadIdx.x & 31) == 0) warp_bits[threadIdx.x >> 5] = bits;
  __syncthreads();
  if (threadIdx.x < 32) {
    bits = warp_max_u32(warp_bits[threadIdx.x]);
    if (threadIdx.x == 0) {
      const float amax = __uint_as_float(bits);
      const float s = scale_from_amax(amax, flt_max, true, dt);  // mpemu/qutils.py:283-285
      scratch[kSlotAmax] = bits;
      scratch[kSlotScale] = __float_as_uint(s);
    }
  }
}

// Device address of g_noalloc_scratch on the CURRENT device, looked up once per device.
int scratch_address(uint32_t** out) {
  static uint32_t* cached[64];
  static std::mutex mu;
  int dev = 0;
  cudaError_t err = cudaGetDevice(&dev);
  if (err != cudaSuccess) return (int)err;
  if (dev < 0 || dev >= 64) return FPEMU_ERR_ARG;
  std::lock_guard<std::mutex> lock(mu);
  if (cached[dev] == nullptr) {
    uint32_t* p = nullptr;
    err = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_noalloc_scratch);
    if (err != cudaSuccess) return (int)err;
    if (p == nullptr) return (int)cudaErrorInvalidSymbol;
    cached[dev] = p;
  }
  *out = cached[dev];
  return FPEMU_OK;
}

}  // namespace

extern "C" int fpemu_cast_amax_scaled_noalloc(const void* in, void* out, int64_t n, int dtype, int format,
                                              int rounding, int flags, float flt_max, uint64_t seed,
                                              uint64_t elem_offset, void* stream_) {
  if (n < 0 || (n > 0 && (in == nullptr || out == nullptr))) return FPEMU_ERR_ARG;
  if (dtype < 0 || dtype >= FPEMU_DT_COUNT) return FPEMU_ERR_DTYPE;
  if (format < 0 || format >= FPEMU_FMT_COUNT || rounding < 0 || rounding >= FPEMU_RND_COUNT) return FPEMU_ERR_MODE;
  if (format == FPEMU_FMT_BF16 || format == FPEMU_FMT_FP16) return FPEMU_ERR_MODE;   // these formats take no scale
  const int esz = dtype_bytes(dtype);
  if (((uintptr_t)in % (uintptr_t)esz) || ((uintptr_t)out % (uintptr_t)esz)) return FPEMU_ERR_ALIGN;
  const int64_t e_per_cta = kSliceBytes / esz;
  if (n > (int64_t)kMaxParts * e_per_cta) return FPEMU_ERR_TOO_LARGE;                 // more than 512 MiB of input
  if (n == 0) return FPEMU_OK;
  uint32_t* scratch = nullptr;
  int st = scratch_address(&scratch);
  if (st != FPEMU_OK) return st;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  const int nparts = (int)((n + e_per_cta - 1) / e_per_cta);
  const int aligned = (((uintptr_t)in) & 31u) == 0 ? 1 : 0;
  partial_amax_kernel<<<nparts, kPartThreads, 0, stream>>>(in, n, dtype, aligned, scratch);
  cudaError_t err = cudaGetLastError();
  if (err != cudaSuccess) return (int)err;
  finish_scale_kernel<<<1, 1024, 0, stream>>>(scratch, nparts, flt_max, dtype);
  err = cudaGetLastError();
  if (err != cudaSuccess) return (int)err;
  return fpemu_cast(in, out, n, dtype, format, rounding, flags, 1.0f,
                    reinterpret_cast<const float*>(scratch + kSlotScale), seed, elem_offset, stream_);
}
