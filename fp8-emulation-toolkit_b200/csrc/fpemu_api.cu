// fpemu_api.cu — host-side helpers of the C ABI: mode-string parser, error text, device info cache.
#include <string.h>

#include <mutex>

#include "fpemu_cast_core.cuh"

namespace fpemu {

const DeviceInfo& device_info() {
  static DeviceInfo info[64];
  static bool have[64];
  static std::mutex mu;
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) dev = 0;
  std::lock_guard<std::mutex> lock(mu);
  if (!have[dev]) {
    int sms = 0;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0) sms = 148;
    info[dev].sm_count = sms;
    have[dev] = true;
  }
  return info[dev];
}

namespace {
const unsigned long long* g_seed_source[64];
std::mutex g_seed_mu;
__global__ void advance_seed_kernel(unsigned long long* p) { *p += 0x9E3779B97F4A7C15ull; }
}  // namespace

const unsigned long long* seed_source() {
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) return nullptr;
  std::lock_guard<std::mutex> lock(g_seed_mu);
  return g_seed_source[dev];
}

}  // namespace fpemu

extern "C" int fpemu_set_seed_source(const uint64_t* d_seed) {
  int dev = 0;
  cudaError_t err = cudaGetDevice(&dev);
  if (err != cudaSuccess) return (int)err;
  if (dev < 0 || dev >= 64) return FPEMU_ERR_ARG;
  std::lock_guard<std::mutex> lock(fpemu::g_seed_mu);
  fpemu::g_seed_source[dev] = reinterpret_cast<const unsigned long long*>(d_seed);
  return FPEMU_OK;
}

extern "C" int fpemu_advance_seed_source(void* stream) {
  const unsigned long long* p = fpemu::seed_source();
  if (p == nullptr) return FPEMU_ERR_ARG;
  fpemu::advance_seed_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(const_cast<unsigned long long*>(p));
  return (int)cudaGetLastError();
}

extern "C" int fpemu_version(void) { return 100; }  // 0.1.0

extern "C" const char* fpemu_strerror(int status) {
  switch (status) {
    case FPEMU_OK:            return "success";
    case FPEMU_ERR_MODE:      return "fpemu: unknown or unsupported mode (format / rounding combination)";
    case FPEMU_ERR_DTYPE:     return "fpemu: unsupported element type (float32, float16, bfloat16 only)";
    case FPEMU_ERR_SCALE:     return "fpemu: invalid scale factor, must not be larger than 65504";
    case FPEMU_ERR_ARG:       return "fpemu: invalid argument (null pointer, negative size or bad block size)";
    case FPEMU_ERR_ALIGN:     return "fpemu: buffer not aligned to its element size";
    case FPEMU_ERR_TOO_LARGE: return "fpemu: problem exceeds a fixed capacity of this entry point";
    default: break;
  }
  if (status > 0) return cudaGetErrorString(static_cast<cudaError_t>(status));
  return "fpemu: unknown status";
}

namespace {

struct Prefix { const char* text; int fmt; int flags; };
const Prefix kPrefixes[] = {
    {"E5M2_DAZ_", FPEMU_FMT_E5M2, FPEMU_FLAG_DAZ}, {"E4M3_IEEE_", FPEMU_FMT_E4M3_IEEE, 0},
    {"FLOAT16_DAZ_", FPEMU_FMT_FP16, FPEMU_FLAG_DAZ}, {"E5M2_", FPEMU_FMT_E5M2, 0},
    {"E4M3_", FPEMU_FMT_E4M3, 0}, {"E3M4_", FPEMU_FMT_E3M4, 0},
    {"BFLOAT16_", FPEMU_FMT_BF16, 0}, {"FLOAT16_", FPEMU_FMT_FP16, 0},
};
struct Suffix { const char* text; int rnd; };
const Suffix kSuffixes[] = {
    {"RTZ", FPEMU_RND_RTZ}, {"RNE", FPEMU_RND_RNE}, {"STOCHASTIC", FPEMU_RND_STOCHASTIC}, {"RNAZ", FPEMU_RND_RNAZ},
    {"RNTZ", FPEMU_RND_RNTZ}, {"RPINF", FPEMU_RND_RPINF}, {"RNINF", FPEMU_RND_RNINF},
};

// Which suffixes each family accepts: the reference's dispatcher strings (fpemu_kernels.cu:1058-1364)
// plus <FMT>_RTZ for E4M3 / E3M4 / BFLOAT16 / E4M3_IEEE (truncate: the kernels' "no mask set" path).
bool accepted(int fmt, int flags, int rnd) {
  const bool daz = (flags & FPEMU_FLAG_DAZ) != 0;
  switch (fmt) {
    case FPEMU_FMT_E5M2:
      if (daz) return rnd == FPEMU_RND_RNE || rnd == FPEMU_RND_STOCHASTIC || rnd == FPEMU_RND_RNAZ || rnd == FPEMU_RND_RNTZ;
      return true;
    case FPEMU_FMT_FP16:
      if (daz) return rnd == FPEMU_RND_RNE;
      return rnd == FPEMU_RND_RNE || rnd == FPEMU_RND_STOCHASTIC;
    default:
      return rnd == FPEMU_RND_RNE || rnd == FPEMU_RND_STOCHASTIC || rnd == FPEMU_RND_RTZ;
  }
}

}  // namespace

extern "C" int fpemu_parse_mode(const char* mode, int* format, int* rounding, int* flags) {
  if (mode == nullptr || format == nullptr || rounding == nullptr || flags == nullptr) return FPEMU_ERR_ARG;
  if (strcmp(mode, "FP4_NEAREST") == 0) {
    *format = FPEMU_FMT_FP4; *rounding = FPEMU_RND_RNE; *flags = 0;
    return FPEMU_OK;
  }
  for (const Prefix& p : kPrefixes) {
    const size_t len = strlen(p.text);
    if (strncmp(mode, p.text, len) != 0) continue;
    for (const Suffix& s : kSuffixes) {
      if (strcmp(mode + len, s.text) == 0 && accepted(p.fmt, p.flags, s.rnd)) {
        *format = p.fmt; *rounding = s.rnd; *flags = p.flags;
        return FPEMU_OK;
      }
    }
    return FPEMU_ERR_MODE;
  }
  return FPEMU_ERR_MODE;
}
