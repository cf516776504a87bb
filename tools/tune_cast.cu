// tune_cast.cu — launch-shape sweep for the cast kernel (development tool, not part of the library).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -o tools/tune_cast tools/tune_cast.cu
//   ./tools/tune_cast [log2_elems]
#include <cstdio>
#include <cstdlib>
#include <functional>
#include <algorithm>
#include <string>
#include <vector>

#include "../fp8-emulation-toolkit_b200/csrc/fpemu_cast_core.cuh"

using namespace fpemu;
namespace fpemu { const DeviceInfo& device_info() { static DeviceInfo d{148}; return d; } }

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

template <int FMT, int RND, int DT, bool SCALED, int THREADS, int UNROLL>
__global__ void __launch_bounds__(THREADS) k_cast(const CastArgs a) {
  constexpr int VE = DtInfo<DT>::kVecElems;
  ScaleCtx sc; sc.scale = 1.f; sc.rscale = 1.f; sc.unscale32 = false;
  if (SCALED) { sc.scale = a.scale; sc.rscale = __frcp_rn(a.scale); }
  const int64_t nvec = a.n / VE;
  const char* in = (const char*)a.in; char* out = (char*)a.out;
  const int64_t stride = (int64_t)gridDim.x * (THREADS * UNROLL);
  for (int64_t base = (int64_t)blockIdx.x * (THREADS * UNROLL) + threadIdx.x; base < nvec; base += stride) {
    Vec32 v[UNROLL];
#pragma unroll
    for (int j = 0; j < UNROLL; ++j) { int64_t vi = base + j * THREADS; if (vi < nvec) v[j] = ldg256(in + vi * 32); }
#pragma unroll
    for (int j = 0; j < UNROLL; ++j) {
      int64_t vi = base + j * THREADS;
      if (vi < nvec) { process_vec<FMT, RND, false, DT, SCALED>(v[j], sc, a, a.elem_offset + (uint64_t)vi * VE); stg256(out + vi * 32, v[j]); }
    }
  }
}

// pure copies for the ceiling
template <int THREADS, int UNROLL, int KIND>
__global__ void __launch_bounds__(THREADS) k_copy(const char* in, char* out, int64_t nvec) {
  const int64_t stride = (int64_t)gridDim.x * (THREADS * UNROLL);
  for (int64_t base = (int64_t)blockIdx.x * (THREADS * UNROLL) + threadIdx.x; base < nvec; base += stride) {
    Vec32 v[UNROLL];
#pragma unroll
    for (int j = 0; j < UNROLL; ++j) {
      int64_t vi = base + j * THREADS;
      if (vi < nvec) {
        if (KIND == 0) v[j] = ldg256(in + vi * 32);
        else if (KIND == 1 || KIND == 3) v[j] = *reinterpret_cast<const Vec32*>(in + vi * 32);
        else if (KIND == 4) asm volatile("ld.global.cs.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(v[j].w[0]), "=r"(v[j].w[1]), "=r"(v[j].w[2]), "=r"(v[j].w[3]), "=r"(v[j].w[4]), "=r"(v[j].w[5]), "=r"(v[j].w[6]), "=r"(v[j].w[7]) : "l"(in + vi * 32));
        else if (KIND == 5) asm volatile("ld.global.L1::no_allocate.L2::evict_first.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(v[j].w[0]), "=r"(v[j].w[1]), "=r"(v[j].w[2]), "=r"(v[j].w[3]), "=r"(v[j].w[4]), "=r"(v[j].w[5]), "=r"(v[j].w[6]), "=r"(v[j].w[7]) : "l"(in + vi * 32));
        else asm volatile("ld.global.nc.L1::no_allocate.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(v[j].w[0]), "=r"(v[j].w[1]), "=r"(v[j].w[2]), "=r"(v[j].w[3]), "=r"(v[j].w[4]), "=r"(v[j].w[5]), "=r"(v[j].w[6]), "=r"(v[j].w[7]) : "l"(in + vi * 32));
      }
    }
#pragma unroll
    for (int j = 0; j < UNROLL; ++j) {
      int64_t vi = base + j * THREADS;
      if (vi < nvec) {
        if (KIND == 1) *reinterpret_cast<Vec32*>(out + vi * 32) = v[j];
        else if (KIND == 3 || KIND == 4) asm volatile("st.global.cs.v8.u32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
               :: "l"(out + vi * 32), "r"(v[j].w[0]), "r"(v[j].w[1]), "r"(v[j].w[2]), "r"(v[j].w[3]), "r"(v[j].w[4]), "r"(v[j].w[5]), "r"(v[j].w[6]), "r"(v[j].w[7]) : "memory");
        else if (KIND == 5) asm volatile("st.global.L1::no_allocate.L2::evict_first.v8.u32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
               :: "l"(out + vi * 32), "r"(v[j].w[0]), "r"(v[j].w[1]), "r"(v[j].w[2]), "r"(v[j].w[3]), "r"(v[j].w[4]), "r"(v[j].w[5]), "r"(v[j].w[6]), "r"(v[j].w[7]) : "memory");
        else stg256(out + vi * 32, v[j]);
      }
    }
  }
}


// ---- TMA bulk staging experiment (north_star: "TMA bulk staging only if ncu shows it helps") ---------------------------
// One tile of TILE bytes per CTA iteration: one elected thread issues a 1-D bulk load (cp.async.bulk, global -> shared,
// completion on an mbarrier), all threads wait, optionally round their 32-byte vectors in place in shared memory
// (E4M3_RNE, the same process_vec as the library kernel), then one thread issues a 1-D bulk store (shared -> global,
// bulk async-group).  STAGES tiles are in flight per CTA (ring of buffers); several CTAs share an SM.
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t phase) {
  asm volatile(
      "{\n.reg .pred p;\nWAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE;\nbra WAIT_LOOP;\nDONE:\n}" :: "r"(smem_u32(bar)), "r"(phase) : "memory");
}
__device__ __forceinline__ void bulk_load(void* smem, const void* gmem, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               :: "r"(smem_u32(smem)), "l"(gmem), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void bulk_store(void* gmem, const void* smem, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" :: "l"(gmem), "r"(smem_u32(smem)), "r"(bytes) : "memory");
}

template <int TILE, int STAGES, bool CAST>
__global__ void __launch_bounds__(256) k_bulk(const CastArgs a, int64_t ntiles, int tiles_per_cta) {
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ uint64_t bars[STAGES];
  const int64_t t0 = (int64_t)blockIdx.x * tiles_per_cta;
  int64_t t1 = t0 + tiles_per_cta;
  if (t1 > ntiles) t1 = ntiles;
  const int my = (int)(t1 - t0);
  const char* in = (const char*)a.in; char* out = (char*)a.out;
  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) mbar_init(&bars[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES && s < my; ++s) {
      mbar_expect_tx(&bars[s], TILE);
      bulk_load(smem + s * TILE, in + (t0 + s) * TILE, TILE, &bars[s]);
    }
  }
  ScaleCtx sc; sc.scale = 1.f; sc.rscale = 1.f; sc.unscale32 = false;
  for (int i = 0; i < my; ++i) {
    const int s = i % STAGES;
    mbar_wait(&bars[s], (i / STAGES) & 1);
    unsigned char* buf = smem + s * TILE;
    if (CAST) {
#pragma unroll
      for (int j = 0; j < TILE / (256 * 32); ++j) {
        Vec32* p = reinterpret_cast<Vec32*>(buf) + j * 256 + threadIdx.x;
        Vec32 v = *p;
        process_vec<FPEMU_FMT_E4M3, FPEMU_RND_RNE, false, FPEMU_DT_F32, false>(v, sc, a, 0);
        *p = v;
      }
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      bulk_store(out + (t0 + i) * TILE, buf, TILE);
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      if (i + STAGES < my) {
        asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");     // the store has read the buffer: reuse it
        mbar_expect_tx(&bars[s], TILE);
        bulk_load(buf, in + (t0 + i + STAGES) * TILE, TILE, &bars[s]);
      }
    }
  }
  if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}

static float time_ms(int reps, const std::function<void()>& f) {
  cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
  for (int i = 0; i < 3; ++i) f();
  CK(cudaEventRecord(a));
  for (int i = 0; i < reps; ++i) f();
  CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b));
  float ms; CK(cudaEventElapsedTime(&ms, a, b));
  CK(cudaGetLastError());
  cudaEventDestroy(a); cudaEventDestroy(b);
  return ms / reps;
}

template <int FMT, int RND, int DT, bool SCALED, int THREADS, int UNROLL>
void run_cast(const char* name, const void* in, void* out, int64_t n, const std::vector<int>& cps) {
  CastArgs a; a.in = in; a.out = out; a.n = n; a.scale = SCALED ? 3.7f : 1.f; a.d_scale = nullptr; a.seed_lo = 1; a.seed_hi = 2;
  a.elem_offset = 0; a.fmt = FMT; a.rnd = RND; a.flags = 0; a.d_amax = nullptr; a.flt_max = 0.f; a.d_scale_out = nullptr; a.d_seed = nullptr;
  const int ve = DtInfo<DT>::kVecElems; const double bytes = 2.0 * n * DtInfo<DT>::kBytes;
  for (int c : cps) {
    int64_t want = (n / ve + THREADS * UNROLL - 1) / (THREADS * UNROLL);
    int grid = c > 0 ? (int)std::min<int64_t>(want, 148LL * c) : (int)want;
    float ms = time_ms(10, [&] { k_cast<FMT, RND, DT, SCALED, THREADS, UNROLL><<<grid, THREADS>>>(a); });
    printf("%-28s T=%4d U=%d ctas/sm=%3d grid=%8d  %8.4f ms  %7.1f GB/s\n", name, THREADS, UNROLL, c, grid, ms, bytes / ms / 1e6);
  }
}

template <int TILE, int STAGES, bool CAST>
void run_bulk(const char* name, const void* in, void* out, int64_t bytes_one, const std::vector<int>& tpcs) {
  CastArgs a; a.in = in; a.out = out; a.n = bytes_one / 4; a.scale = 1.f; a.d_scale = nullptr; a.seed_lo = 1; a.seed_hi = 2;
  a.elem_offset = 0; a.fmt = FPEMU_FMT_E4M3; a.rnd = FPEMU_RND_RNE; a.flags = 0; a.d_amax = nullptr; a.flt_max = 0.f;
  a.d_scale_out = nullptr; a.d_seed = nullptr;
  const int64_t ntiles = bytes_one / TILE;
  const int smem = TILE * STAGES;
  CK(cudaFuncSetAttribute(k_bulk<TILE, STAGES, CAST>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  for (int tpc : tpcs) {
    const int64_t grid = (ntiles + tpc - 1) / tpc;
    float ms = time_ms(10, [&] { k_bulk<TILE, STAGES, CAST><<<(unsigned)grid, 256, smem>>>(a, ntiles, tpc); });
    printf("%-28s tile=%3d KiB stages=%d tiles/cta=%4d grid=%8lld  %8.4f ms  %7.1f GB/s\n", name, TILE / 1024, STAGES, tpc,
           (long long)grid, ms, 2.0 * bytes_one / ms / 1e6);
  }
}

template <int THREADS, int UNROLL, int KIND>
void run_copy(const char* name, const void* in, void* out, int64_t bytes_one, const std::vector<int>& cps) {
  int64_t nvec = bytes_one / 32;
  for (int c : cps) {
    int64_t want = (nvec + THREADS * UNROLL - 1) / (THREADS * UNROLL);
    int grid = c > 0 ? (int)std::min<int64_t>(want, 148LL * c) : (int)want;
    float ms = time_ms(10, [&] { k_copy<THREADS, UNROLL, KIND><<<grid, THREADS>>>((const char*)in, (char*)out, nvec); });
    printf("%-28s T=%4d U=%d ctas/sm=%3d grid=%8d  %8.4f ms  %7.1f GB/s\n", name, THREADS, UNROLL, c, grid, ms, 2.0 * bytes_one / ms / 1e6);
  }
}

#define CASTS(FMT, RND, DT, SC, NAME) \
  run_cast<FMT, RND, DT, SC, 256, 1>(NAME, in, out, n_for(DT), cps); \
  run_cast<FMT, RND, DT, SC, 256, 2>(NAME, in, out, n_for(DT), cps); \
  run_cast<FMT, RND, DT, SC, 256, 4>(NAME, in, out, n_for(DT), cps); \
  run_cast<FMT, RND, DT, SC, 512, 2>(NAME, in, out, n_for(DT), cps); \
  run_cast<FMT, RND, DT, SC, 128, 4>(NAME, in, out, n_for(DT), cps);

int main(int argc, char** argv) {
  int lg = argc > 1 ? atoi(argv[1]) : 30;
  const int64_t n32 = 1LL << lg;
  const int64_t bytes = n32 * 4;
  void *in, *out;
  CK(cudaMalloc(&in, bytes)); CK(cudaMalloc(&out, bytes));
  // fill with a pattern of finite floats around 1.0 (N(0,1)-like exponents)
  {
    std::vector<uint32_t> h(1 << 20);
    uint32_t s = 12345;
    for (auto& w : h) { s = s * 1664525u + 1013904223u; uint32_t e = 120 + ((s >> 9) % 10); w = (s & 0x807FFFFFu) | (e << 23); }
    for (int64_t off = 0; off < bytes; off += (int64_t)h.size() * 4) CK(cudaMemcpy((char*)in + off, h.data(), std::min<int64_t>(h.size() * 4, bytes - off), cudaMemcpyHostToDevice));
  }
  auto n_for = [&](int dt) { return dt == FPEMU_DT_F32 ? n32 : 2 * n32; };   // same bytes
  std::vector<int> cps = {2, 4, 8, 16, 32, 0};
  printf("# %lld MiB per buffer\n", (long long)(bytes >> 20));
  float ms = time_ms(10, [&] { CK(cudaMemcpyAsync(out, in, bytes, cudaMemcpyDeviceToDevice)); });
  printf("%-28s %8.4f ms  %7.1f GB/s\n", "cudaMemcpyAsync D2D", ms, 2.0 * bytes / ms / 1e6);
  run_copy<256, 1, 0>("copy ld.na.256", in, out, bytes, cps);
  run_copy<256, 2, 0>("copy ld.na.256", in, out, bytes, cps);
  run_copy<256, 4, 0>("copy ld.na.256", in, out, bytes, cps);
  run_copy<512, 2, 0>("copy ld.na.256", in, out, bytes, cps);
  run_copy<256, 2, 1>("copy plain 256", in, out, bytes, cps);
  run_copy<256, 2, 2>("copy ld.nc.na.256", in, out, bytes, cps);
  if (argc > 2 && std::string(argv[2]) == "bulk") {   // TMA bulk staging vs the plain 256-bit kernels
    std::vector<int> one = {0};
    for (int rep = 0; rep < 2; ++rep) {
      run_copy<256, 2, 0>("copy ld.na.256 / st.na.256", in, out, bytes, one);
      run_cast<FPEMU_FMT_E4M3, FPEMU_RND_RNE, FPEMU_DT_F32, false, 256, 2>("E4M3_RNE f32 (library shape)", in, out, n32, one);
      run_bulk<16384, 1, false>("bulk copy", in, out, bytes, {1, 4});
      run_bulk<32768, 1, false>("bulk copy", in, out, bytes, {1, 4});
      run_bulk<16384, 2, false>("bulk copy", in, out, bytes, {2, 8, 64});
      run_bulk<32768, 2, false>("bulk copy", in, out, bytes, {2, 8, 64});
      run_bulk<16384, 4, false>("bulk copy", in, out, bytes, {8, 64, 885});
      run_bulk<32768, 3, false>("bulk copy", in, out, bytes, {8, 64, 443});
      run_bulk<16384, 1, true>("bulk E4M3_RNE f32", in, out, bytes, {1, 4});
      run_bulk<32768, 1, true>("bulk E4M3_RNE f32", in, out, bytes, {1, 4});
      run_bulk<16384, 2, true>("bulk E4M3_RNE f32", in, out, bytes, {2, 8, 64});
      run_bulk<32768, 2, true>("bulk E4M3_RNE f32", in, out, bytes, {2, 8, 64});
      run_bulk<16384, 4, true>("bulk E4M3_RNE f32", in, out, bytes, {8, 64, 885});
    }
    return 0;
  }
  if (argc > 2) {   // cache-policy experiments only
    std::vector<int> one = {0};
    for (int rep = 0; rep < 3; ++rep) {
      run_copy<256, 2, 0>("ld.na / st.na", in, out, bytes, one);
      run_copy<256, 2, 1>("ld / st (default)", in, out, bytes, one);
      run_copy<256, 2, 3>("ld / st.cs", in, out, bytes, one);
      run_copy<256, 2, 4>("ld.cs / st.cs", in, out, bytes, one);
      run_copy<256, 2, 5>("ld.na.evict_first / st.na.evict_first", in, out, bytes, one);
      run_copy<512, 2, 0>("ld.na / st.na T512", in, out, bytes, one);
      run_copy<256, 4, 0>("ld.na / st.na U4", in, out, bytes, one);
      run_copy<128, 2, 0>("ld.na / st.na T128", in, out, bytes, one);
    }
    return 0;
  }
  CASTS(FPEMU_FMT_E5M2, FPEMU_RND_RTZ, FPEMU_DT_F32, false, "E5M2_RTZ f32")
  CASTS(FPEMU_FMT_E5M2, FPEMU_RND_RNE, FPEMU_DT_F32, false, "E5M2_RNE f32")
  CASTS(FPEMU_FMT_E4M3, FPEMU_RND_RNE, FPEMU_DT_F32, false, "E4M3_RNE f32")
  CASTS(FPEMU_FMT_E4M3, FPEMU_RND_STOCHASTIC, FPEMU_DT_F32, false, "E4M3_SR f32")
  CASTS(FPEMU_FMT_E4M3, FPEMU_RND_RNE, FPEMU_DT_F32, true, "E4M3_RNE f32 scaled")
  CASTS(FPEMU_FMT_E3M4, FPEMU_RND_STOCHASTIC, FPEMU_DT_F32, false, "E3M4_SR f32")
  CASTS(FPEMU_FMT_BF16, FPEMU_RND_RNE, FPEMU_DT_F32, false, "BF16_RNE f32")
  CASTS(FPEMU_FMT_E4M3, FPEMU_RND_RNE, FPEMU_DT_BF16, false, "E4M3_RNE bf16")
  CASTS(FPEMU_FMT_E4M3, FPEMU_RND_RNE, FPEMU_DT_BF16, true, "E4M3_RNE bf16 scaled")
  CASTS(FPEMU_FMT_E4M3, FPEMU_RND_STOCHASTIC, FPEMU_DT_BF16, false, "E4M3_SR bf16")
  CASTS(FPEMU_FMT_E4M3, FPEMU_RND_RNE, FPEMU_DT_F16, false, "E4M3_RNE f16")
  return 0;
}
