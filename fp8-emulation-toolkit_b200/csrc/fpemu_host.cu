// fpemu_host.cu — fpemu_cast_host: host buffer in, host buffer out, staged through two device buffers so the
// H2D copy, the cast kernel and the D2H copy of consecutive chunks overlap (PCIe is full duplex; the kernel
// is ~100x faster than either copy).  This is the path the end-to-end measurement times.
#include "fpemu_common.cuh"

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
  cudaEvent_t ev_in[2], ev_k[2], ev_out[2];
  cudaError_t err = cudaSuccess;
  int created = 0;
  for (int i = 0; i < 2 && err == cudaSuccess; ++i) {
    err = cudaEventCreateWithFlags(&ev_in[i], cudaEventDisableTiming);
    if (err == cudaSuccess) err = cudaEventCreateWithFlags(&ev_k[i], cudaEventDisableTiming);
    if (err == cudaSuccess) err = cudaEventCreateWithFlags(&ev_out[i], cudaEventDisableTiming);
    if (err == cudaSuccess) created = i + 1;
  }
  int status = (int)err;
  int64_t c = 0;
  for (int64_t off = 0; off < n && status == 0; off += chunk_elems, ++c) {
    const int b = (int)(c & 1);
    const int64_t m = (n - off < chunk_elems) ? (n - off) : chunk_elems;
    const char* src = static_cast<const char*>(h_in) + (size_t)off * esz;
    char* dst = static_cast<char*>(h_out) + (size_t)off * esz;
    if (c >= 2) status = (int)cudaStreamWaitEvent(s_in, ev_out[b], 0);        // stage buffer b is free again
    if (status == 0) status = (int)cudaMemcpyAsync(stage[b], src, (size_t)m * esz, cudaMemcpyHostToDevice, s_in);
    if (status == 0) status = (int)cudaEventRecord(ev_in[b], s_in);
    if (status == 0) status = (int)cudaStreamWaitEvent(s_k, ev_in[b], 0);
    if (status == 0) status = fpemu_cast(stage[b], stage[b], m, dtype, format, rounding, flags, scale, nullptr, seed,
                                         (uint64_t)off, s_k);
    if (status == 0) status = (int)cudaEventRecord(ev_k[b], s_k);
    if (status == 0) status = (int)cudaStreamWaitEvent(s_out, ev_k[b], 0);
    if (status == 0) status = (int)cudaMemcpyAsync(dst, stage[b], (size_t)m * esz, cudaMemcpyDeviceToHost, s_out);
    if (status == 0) status = (int)cudaEventRecord(ev_out[b], s_out);
  }
  cudaError_t e1 = cudaStreamSynchronize(s_in);
  cudaError_t e2 = cudaStreamSynchronize(s_k);
  cudaError_t e3 = cudaStreamSynchronize(s_out);
  for (int i = 0; i < created; ++i) { cudaEventDestroy(ev_in[i]); cudaEventDestroy(ev_k[i]); cudaEventDestroy(ev_out[i]); }
  if (status != 0) return status;
  if (e1 != cudaSuccess) return (int)e1;
  if (e2 != cudaSuccess) return (int)e2;
  return (int)e3;
}
