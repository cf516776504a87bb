// fpemu_host.cu — host buffer in, host buffer(s) out: the paths the end-to-end measurement times.
//
//   fpemu_cast_host        one cast of a host tensor, staged through two device buffers so the H2D copy, the cast kernel
//                          and the D2H copy of consecutive chunks overlap (PCIe is full duplex; the kernel is ~100x faster
//                          than either copy).
//   fpemu_cast_host_multi  M casts of ONE host tensor (a format x rounding sweep, plus per-tensor amax-scaled casts):
//                          the input crosses the bus once, chunk by chunk, into a device-resident copy; every chunk is
//                          cast M times into a two-slot staging ring and each result goes back to its own host buffer.
//                          Amax-scaled jobs reduce |x| chunk by chunk while the plain jobs run and are cast in a second
//                          pass over the resident copy (their scale needs the whole tensor).  Bus traffic: 1 x H2D + M x D2H
//                          instead of M x (H2D + D2H).
//
// Events are created once per host thread and device and reused by every call (cudaStreamWaitEvent captures the
// most recent record at the time it is called, so one event per dependency edge is enough).
#include "fpemu_cast_core.cuh"

namespace fpemu {
int validate_cast(int64_t n, int dtype, int format, int rounding, float scale, const void* in, const void* out);
int launch_cast(const CastArgs& a, int dtype, cudaStream_t stream);
}  // namespace fpemu

extern "C" int fpemu_cast(const void* in, void* out, int64_t n, int dtype, int format, int rounding, int flags, float scale,
                          const float* d_scale, uint64_t seed, uint64_t elem_offset, void* stream);
extern "C" int fpemu_amax(const void* in, int64_t n, int dtype, float* d_amax, void* stream);

namespace {

struct HostEvents {
  cudaEvent_t in, k, out[2];
  bool ready;
};

// Events of the calling thread for the current device (created on first use, kept for the life of the thread).
int host_events(HostEvents** ev) {
  thread_local HostEvents cache[64];
  int dev = 0;
  cudaError_t err = cudaGetDevice(&dev);
  if (err != cudaSuccess) return (int)err;
  if (dev < 0 || dev >= 64) return FPEMU_ERR_ARG;
  HostEvents& e = cache[dev];
  if (!e.ready) {
    cudaEvent_t* all[4] = {&e.in, &e.k, &e.out[0], &e.out[1]};
    for (int i = 0; i < 4; ++i) {
      err = cudaEventCreateWithFlags(all[i], cudaEventDisableTiming);
      if (err != cudaSuccess) {
        for (int j = 0; j < i; ++j) cudaEventDestroy(*all[j]);
        return (int)err;
      }
    }
    e.ready = true;
  }
  *ev = &e;
  return FPEMU_OK;
}

#define FPEMU_TRY(expr)                          \
  do {                                           \
    if (status == 0) status = (int)(expr);       \
  } while (0)

}  // namespace

extern "C" int fpemu_cast_host(const void* h_in, void* h_out, int64_t n, int dtype, int format, int rounding,
                               int flags, float scale, uint64_t seed, void* d_stage0, void* d_stage1,
                               int64_t chunk_elems, void* stream_h2d, void* stream_compute, void* stream_d2h) {
  if (n < 0 || chunk_elems <= 0 || d_stage0 == nullptr || d_stage1 == nullptr) return FPEMU_ERR_ARG;
  if (n > 0 && (h_in == nullptr || h_out == nullptr)) return FPEMU_ERR_ARG;
  if (dtype < 0 || dtype >= FPEMU_DT_COUNT) return FPEMU_ERR_DTYPE;
  if (chunk_elems % 16 != 0) return FPEMU_ERR_ARG;          // keeps every chunk on a Philox-group boundary
  cudaStream_t s_in = static_cast<cudaStream_t>(stream_h2d);
  cudaStream_t s_k = static_cast<cudaStream_t>(stream_compute);
  cudaStream_t s_out = static_cast<cudaStream_t>(stream_d2h);
  const size_t esz = (size_t)fpemu::dtype_bytes(dtype);
  void* stage[2] = {d_stage0, d_stage1};
  HostEvents* ev = nullptr;
  int status = host_events(&ev);
  if (status != 0) return status;
  int64_t c = 0;
  for (int64_t off = 0; off < n && status == 0; off += chunk_elems, ++c) {
    const int b = (int)(c & 1);
    const int64_t m = (n - off < chunk_elems) ? (n - off) : chunk_elems;
    const char* src = static_cast<const char*>(h_in) + (size_t)off * esz;
    char* dst = static_cast<char*>(h_out) + (size_t)off * esz;
    if (c >= 2) FPEMU_TRY(cudaStreamWaitEvent(s_in, ev->out[b], 0));          // stage buffer b is free again
    FPEMU_TRY(cudaMemcpyAsync(stage[b], src, (size_t)m * esz, cudaMemcpyHostToDevice, s_in));
    FPEMU_TRY(cudaEventRecord(ev->in, s_in));
    FPEMU_TRY(cudaStreamWaitEvent(s_k, ev->in, 0));
    FPEMU_TRY(fpemu_cast(stage[b], stage[b], m, dtype, format, rounding, flags, scale, nullptr, seed, (uint64_t)off, s_k));
    FPEMU_TRY(cudaEventRecord(ev->k, s_k));
    FPEMU_TRY(cudaStreamWaitEvent(s_out, ev->k, 0));
    FPEMU_TRY(cudaMemcpyAsync(dst, stage[b], (size_t)m * esz, cudaMemcpyDeviceToHost, s_out));
    FPEMU_TRY(cudaEventRecord(ev->out[b], s_out));
  }
  cudaError_t e1 = cudaStreamSynchronize(s_in);
  cudaError_t e2 = cudaStreamSynchronize(s_k);
  cudaError_t e3 = cudaStreamSynchronize(s_out);
  if (status != 0) return status;
  if (e1 != cudaSuccess) return (int)e1;
  if (e2 != cudaSuccess) return (int)e2;
  return (int)e3;
}

extern "C" int fpemu_cast_host_multi(const void* h_in, int64_t n, int dtype, const fpemu_host_job* jobs, int njobs,
                                     void* d_in, void* d_stage0, void* d_stage1, float* d_work, int64_t chunk_elems,
                                     void* stream_h2d, void* stream_compute, void* stream_d2h) {
  using namespace fpemu;
  if (n < 0 || njobs <= 0 || jobs == nullptr || chunk_elems <= 0 || chunk_elems % 16 != 0) return FPEMU_ERR_ARG;
  if (d_in == nullptr || d_stage0 == nullptr || d_stage1 == nullptr) return FPEMU_ERR_ARG;
  if (n > 0 && h_in == nullptr) return FPEMU_ERR_ARG;
  if (dtype < 0 || dtype >= FPEMU_DT_COUNT) return FPEMU_ERR_DTYPE;
  int n_amax = 0;
  for (int j = 0; j < njobs; ++j) {
    const fpemu_host_job& job = jobs[j];
    int st = validate_cast(n, dtype, job.format, job.rounding, job.amax_scaled ? 1.0f : job.scale, d_in, d_in);
    if (st != FPEMU_OK) return st;
    if (n > 0 && job.h_out == nullptr) return FPEMU_ERR_ARG;
    if (job.amax_scaled) {
      if (job.format == FPEMU_FMT_BF16 || job.format == FPEMU_FMT_FP16) return FPEMU_ERR_MODE;   // these take no scale
      ++n_amax;
    }
  }
  if (n_amax > 0 && d_work == nullptr) return FPEMU_ERR_ARG;
  if (n == 0) return FPEMU_OK;
  cudaStream_t s_in = static_cast<cudaStream_t>(stream_h2d);
  cudaStream_t s_k = static_cast<cudaStream_t>(stream_compute);
  cudaStream_t s_out = static_cast<cudaStream_t>(stream_d2h);
  const size_t esz = (size_t)dtype_bytes(dtype);
  void* stage[2] = {d_stage0, d_stage1};
  HostEvents* ev = nullptr;
  int status = host_events(&ev);
  if (status != 0) return status;
  int64_t slot_uses = 0;                                    // staging slot = slot_uses & 1

  // one cast of one chunk of the resident copy into the next staging slot, and its way back to the host
  auto emit = [&](const fpemu_host_job& job, int64_t off, int64_t m, const float* d_amax) {
    const int b = (int)(slot_uses & 1);
    if (slot_uses >= 2) FPEMU_TRY(cudaStreamWaitEvent(s_k, ev->out[b], 0));   // the slot's last D2H has finished
    ++slot_uses;
    if (status != 0) return;
    CastArgs a;
    a.in = static_cast<const char*>(d_in) + (size_t)off * esz;
    a.out = stage[b];
    a.n = m;
    a.scale = job.scale; a.d_scale = nullptr;
    a.seed_lo = (uint32_t)job.seed; a.seed_hi = (uint32_t)(job.seed >> 32);
    a.elem_offset = (uint64_t)off;
    a.fmt = job.format; a.rnd = job.rounding; a.flags = job.flags;
    a.d_amax = d_amax; a.flt_max = job.flt_max; a.d_scale_out = nullptr; a.d_seed = seed_source();
    if (job.format == FPEMU_FMT_BF16 || job.format == FPEMU_FMT_FP16) a.scale = 1.0f;
    FPEMU_TRY(launch_cast(a, dtype, s_k));
    FPEMU_TRY(cudaEventRecord(ev->k, s_k));
    FPEMU_TRY(cudaStreamWaitEvent(s_out, ev->k, 0));
    FPEMU_TRY(cudaMemcpyAsync(static_cast<char*>(job.h_out) + (size_t)off * esz, stage[b], (size_t)m * esz,
                              cudaMemcpyDeviceToHost, s_out));
    FPEMU_TRY(cudaEventRecord(ev->out[b], s_out));
  };

  if (n_amax > 0) FPEMU_TRY(cudaMemsetAsync(d_work, 0, sizeof(float), s_k));
  for (int64_t off = 0; off < n && status == 0; off += chunk_elems) {
    const int64_t m = (n - off < chunk_elems) ? (n - off) : chunk_elems;
    char* d_chunk = static_cast<char*>(d_in) + (size_t)off * esz;
    FPEMU_TRY(cudaMemcpyAsync(d_chunk, static_cast<const char*>(h_in) + (size_t)off * esz, (size_t)m * esz,
                              cudaMemcpyHostToDevice, s_in));
    FPEMU_TRY(cudaEventRecord(ev->in, s_in));
    FPEMU_TRY(cudaStreamWaitEvent(s_k, ev->in, 0));
    if (n_amax > 0) FPEMU_TRY(fpemu_amax(d_chunk, m, dtype, d_work, s_k));
    for (int j = 0; j < njobs && status == 0; ++j)
      if (!jobs[j].amax_scaled) emit(jobs[j], off, m, nullptr);
  }
  // second pass over the resident copy for the jobs whose scale needed the whole tensor
  for (int j = 0; j < njobs && status == 0; ++j) {
    if (!jobs[j].amax_scaled) continue;
    for (int64_t off = 0; off < n && status == 0; off += chunk_elems)
      emit(jobs[j], off, (n - off < chunk_elems) ? (n - off) : chunk_elems, d_work);
  }
  cudaError_t e1 = cudaStreamSynchronize(s_in);
  cudaError_t e2 = cudaStreamSynchronize(s_k);
  cudaError_t e3 = cudaStreamSynchronize(s_out);
  if (status != 0) return status;
  if (e1 != cudaSuccess) return (int)e1;
  if (e2 != cudaSuccess) return (int)e2;
  return (int)e3;
}
