/*
 * fpemu_b200.h — C ABI of libfpemu_b200.so, the B200 (sm_100a) replacement for the native side of
 * the FP8-Emulation-Toolkit `fpemu_cuda` operator.
 *
 * Every entry point takes plain pointers and sizes (no torch types), enqueues work on the stream the
 * caller passes (a cudaStream_t as void*; NULL = legacy default stream), never synchronises the device
 * and, unless stated otherwise, never allocates or frees device memory.
 *
 * Return convention (replaces the reference's `fprintf(stderr) + exit(1)`,
 * mpemu/pytquant/cuda/fpemu_kernels.cu:1037-1041, and AT_CUDA_CHECK throws, :1071):
 *      0  success
 *     <0  invalid argument (FPEMU_ERR_*)
 *     >0  a cudaError_t raised by the launch
 * fpemu_strerror() turns any of them into text.
 */
#ifndef FPEMU_B200_H
#define FPEMU_B200_H

#include <stdint.h>

#if defined(__GNUC__)
#define FPEMU_API __attribute__((visibility("default")))
#else
#define FPEMU_API
#endif

#ifdef __cplusplus
extern "C" {
#endif

/* Target number formats.  Reference kernels: fpemu_kernels.cu E5M2_Kernel :375-463, E5M2_DAZ_Kernel
 * :465-541 (format E5M2 + FPEMU_FLAG_DAZ), E4M3_Kernel :543-654, E3M4_Kernel :873-984,
 * BFLOAT16_Kernel :243-305, FLOAT16_Kernel :307-373, E4M3_IEEE_Kernel :656-767,
 * FP4_Nearest_Kernel :986-1024. */
enum {
  FPEMU_FMT_E5M2 = 0,
  FPEMU_FMT_E4M3 = 1,
  FPEMU_FMT_E3M4 = 2,
  FPEMU_FMT_BF16 = 3,
  FPEMU_FMT_FP16 = 4,
  FPEMU_FMT_E4M3_IEEE = 5,
  FPEMU_FMT_FP4 = 6,
  FPEMU_FMT_COUNT = 7
};

/* Rounding modes; numeric values equal the reference's ROUNDING_MODES (fpemu_kernels.cu:17-18). */
enum {
  FPEMU_RND_RTZ = 0,        /* truncate                                  */
  FPEMU_RND_RNE = 1,        /* nearest, ties to even                     */
  FPEMU_RND_STOCHASTIC = 2, /* P(round away) = dropped fraction          */
  FPEMU_RND_RNAZ = 3,       /* nearest, ties away from zero              */
  FPEMU_RND_RNTZ = 4,       /* nearest, ties toward zero                 */
  FPEMU_RND_RPINF = 5,      /* reference "round to +INF" (sic: >= half)  */
  FPEMU_RND_RNINF = 6,      /* reference "round to -INF" (sic: >= half)  */
  FPEMU_RND_COUNT = 7
};

/* Flags. */
enum {
  FPEMU_FLAG_DAZ = 1,          /* denormal flush variants: E5M2_DAZ_*, FLOAT16_DAZ_RNE               */
  FPEMU_FLAG_UNSCALE_FP32 = 2, /* multiply by 1/scale in fp32 like the reference CPU op
                                  (cpp/fpemu_impl.cpp:761,1131,1716).  Default (flag clear) rounds the
                                  un-scaled value through fp16 like the reference CUDA op
                                  (fpemu_kernels.cu:461,539,652,982).                                 */
  FPEMU_FLAG_NO_SCALE_CLAMP = 4, /* fpemu_cast_rowwise only: skip the `6.55e4 if scale > 3.275e4` clamp,
                                  as the fine-grained branch of quantize_tensor does (qutils.py:328-330)  */
  FPEMU_FLAG_EXPECT_NORMALS = 8 /* fpemu_encode / fpemu_encode_multi, E4M3 + RNE only: a performance hint that
                                  values below the format's smallest normal (after scaling) are rare, as they are
                                  when the scale was derived from the tensor's maximum.  Enables the hardware
                                  fp16 -> e4m3 conversion for vectors made of normals, zeros and overflows (where
                                  it is bit-identical to the reference's rounding); other vectors are redone
                                  exactly.  Never changes a result; costs throughput when the hint is wrong.   */
};

/* Element types of the in/out buffers (both buffers always share one type). */
enum { FPEMU_DT_F32 = 0, FPEMU_DT_F16 = 1, FPEMU_DT_BF16 = 2, FPEMU_DT_COUNT = 3 };

/* Error codes. */
enum {
  FPEMU_OK = 0,
  FPEMU_ERR_MODE = -1,     /* unknown mode string / format / rounding combination */
  FPEMU_ERR_DTYPE = -2,    /* unsupported element type                            */
  FPEMU_ERR_SCALE = -3,    /* scale > 65504 (reference: exit(1))                  */
  FPEMU_ERR_ARG = -4,      /* null pointer, negative size, bad block size, ...    */
  FPEMU_ERR_ALIGN = -5,    /* pointer not aligned to its element size             */
  FPEMU_ERR_TOO_LARGE = -6 /* problem exceeds a documented fixed capacity         */
};

/* Library version (major*10000 + minor*100 + patch). */
FPEMU_API int fpemu_version(void);

/* Text for a status returned by any entry point (static storage, never NULL). */
FPEMU_API const char* fpemu_strerror(int status);

/* Stochastic rounding under CUDA-graph replay.  The `seed` argument of every entry point is a host value: a captured
 * graph would replay the same random bits forever.  fpemu_set_seed_source(d_seed) installs, for the CURRENT device, the
 * address of a device-resident uint64 that every stochastic kernel adds to its seed when it starts (NULL removes it);
 * fpemu_advance_seed_source(stream) enqueues a one-thread kernel that moves that word on.  Capture the advance together
 * with the casts and every replay draws fresh bits; results stay a pure function of (seed + word, element index).
 * Process-wide per-device setting, read when a call is made (a captured graph keeps the address it was captured with:
 * keep that word alive and change its VALUE, not its address, afterwards); deterministic modes never look at it. */
FPEMU_API int fpemu_set_seed_source(const uint64_t* d_seed);
FPEMU_API int fpemu_advance_seed_source(void* stream);

/* Parse a reference mode string ("E4M3_RNE", "E5M2_DAZ_STOCHASTIC", "BFLOAT16_RNE", ...) into
 * (format, rounding, flags).  Replaces the `mode.compare(...)` chain of fpemu_cuda_forward,
 * fpemu_kernels.cu:1058-1364.  Besides the reference's strings it accepts <FMT>_RTZ for E4M3, E3M4,
 * BFLOAT16 (truncate = "same pipeline, rounding add 0").  Unknown strings return
 * FPEMU_ERR_MODE (the reference silently returns zeros, :1375-1380). */
FPEMU_API int fpemu_parse_mode(const char* mode, int* format, int* rounding, int* flags);

/* Elementwise quantize-dequantize cast of n elements: out[i] = QDQ(in[i]).  in == out is allowed
 * (in-place).  Replaces the kernel launches of fpemu_cuda_forward (fpemu_kernels.cu:1028-1381) for
 * block_norm == false.
 *   scale    host scalar applied before rounding; the result is multiplied by 1/scale.  Larger than 65504:
 *            FPEMU_ERR_SCALE (reference: exit(1), :1037-1041); a NaN passes that test there and here, and
 *            turns every output into NaN.
 *   d_scale  optional DEVICE pointer to one float; when non-NULL it overrides `scale` (no host sync;
 *            fed by fpemu_scale_from_amax).
 *   seed, elem_offset  stochastic rounding uses Philox4x32-10 keyed by seed with the counter derived
 *            from (elem_offset + i), so a tensor cast in slabs gives the same bits as in one piece.
 */
FPEMU_API int fpemu_cast(const void* in, void* out, int64_t n, int dtype, int format, int rounding, int flags,
               float scale, const float* d_scale, uint64_t seed, uint64_t elem_offset, void* stream);

/* The same cast for float64 tensors (in / out: n doubles, 8-byte aligned; in == out allowed).  The reference
 * dispatches its kernels for double too (AT_DISPATCH_FLOATING_TYPES_AND_HALF, fpemu_kernels.cu:1059 ...): the element is
 * narrowed to float — `(float)x`, and for the formats that take a scale `(float)(x * (double)scale)`, both operands
 * double (`scalar_t scale`, :405, :420) — rounded like a float tensor's element and widened back (:136-148).  Same
 * random field per element index as fpemu_cast.  Only this entry point takes doubles. */
FPEMU_API int fpemu_cast_f64(const void* in, void* out, int64_t n, int format, int rounding, int flags, float scale,
                             const float* d_scale, uint64_t seed, uint64_t elem_offset, void* stream);

/* Blocked cast ("block_norm"): every run of block_size consecutive elements gets its own power-of-two
 * scale derived from its absolute maximum, 2*2^floor(log2 max)/K with K = 16384 (E5M2), 8 (E4M3,
 * E4M3_IEEE), 1 (E3M4); FP4 uses 1/max.  Replaces absmax_block + the per-kernel scale derivation,
 * fpemu_kernels.cu:150-170, :407-414, :575-582, :905-912.  n must be a multiple of block_size
 * (the reference silently disables blocking otherwise, :1048-1051; this ABI returns FPEMU_ERR_ARG and
 * leaves that policy to the caller).  Formats whose reference kernels take no block_norm argument
 * (BFLOAT16, FLOAT16, E5M2_DAZ) return FPEMU_ERR_MODE. */
FPEMU_API int fpemu_cast_blocked(const void* in, void* out, int64_t n, int64_t block_size, int dtype, int format,
                       int rounding, int flags, uint64_t seed, uint64_t elem_offset, void* stream);

/* Single-pass absolute maximum of n elements, max-accumulated into *d_amax (device float the caller
 * initialised, normally to 0).  A NaN element makes the result NaN, exactly as
 * `torch.max(torch.abs(tensor.flatten()))` of mpemu/qutils.py:283 does (the scale derived from it and every
 * value cast with that scale are then NaN too — DESIGN.md §2 "NaN in the maximum"); the maximum is kept as a bit
 * pattern, where every NaN sorts above +Inf, so the accumulation stays one integer atomicMax. */
FPEMU_API int fpemu_amax(const void* in, int64_t n, int dtype, float* d_amax, void* stream);

/* d_scale[i] = the scale quantize_tensor derives from the maximum d_amax[i], for i < count:
 *     scale = flt_max / vmax;  scale = 6.55e4 if scale > 3.275e4 else scale        (mpemu/qutils.py:283-285)
 * `vmax` is a 0-dim tensor there, so torch evaluates the division as `vmax.reciprocal() * flt_max` in the tensor's
 * element type: fl32(fl32(1 / vmax) * flt_max) for float32 (and bfloat16, which the reference does not dispatch);
 * reciprocal, product and comparison each rounded to / done in half for float16.  `dtype` names that element type;
 * fpemu_scale_from_amax is the float32 form.  Device-side, no host sync.  Every entry point below that derives a
 * scale from a maximum (amax-scaled, row-wise, fused, no-alloc, multi) uses the same rule for its tensor's type. */
FPEMU_API int fpemu_scale_from_amax(const float* d_amax, int count, float flt_max, float* d_scale, void* stream);
FPEMU_API int fpemu_scale_from_amax_dtype(const float* d_amax, int count, float flt_max, int dtype, float* d_scale,
                                          void* stream);

/* Per-tensor `mean` scaling of quantize_tensor (mpemu/qutils.py:276-281), on the device, no host sync:
 *     mean = mean|x|;  mean = flt_min if not |mean| > mean_floor;  *d_scale = max(1, flt_min / mean)
 * (the division evaluated like torch's for a 0-dim tensor: reciprocal * flt_min in the tensor's element type).
 * Two launches, deterministic (fixed partial-sum assignment, double accumulation).  d_work: device scratch of
 * fpemu_mean_scale_work_bytes() bytes, 8-byte aligned. */
FPEMU_API int fpemu_mean_scale_work_bytes(void);
FPEMU_API int fpemu_scale_from_abs_mean(const void* in, int64_t n, int dtype, float flt_min, float mean_floor, void* d_work,
                                        float* d_scale, void* stream);

/* *d_scale = the blocked mode's power-of-two scale 2*2^floor(log2 amax)/K (fpemu_kernels.cu:409-413, 577-581, 909-911;
 * FP4: 1/amax, :998) for ONE block whose maximum is *d_amax — used when a single block spans a whole large tensor
 * (block_size == size, the reference CPU op's reading, cpp/fpemu_impl.cpp:778-779): fpemu_amax, this, fpemu_cast. */
FPEMU_API int fpemu_block_scale_from_amax(const float* d_amax, int format, int dtype, float* d_scale, void* stream);

/* Per-tensor amax-scaled cast in one call (the per-tensor `max` branch of quantize_tensor, mpemu/qutils.py:274-287):
 * clears d_work[0..1], reduces max|x| into d_work[0] and casts with the scale fpemu_scale_from_amax_dtype defines,
 * which the cast kernel derives itself and stores in d_work[1].  d_work: two device floats owned by the caller (scratch; calls that
 * overlap on the GPU need separate d_work).  Three enqueues on `stream`, no host sync. */
FPEMU_API int fpemu_cast_amax_scaled(const void* in, void* out, int64_t n, int dtype, int format, int rounding,
                                     int flags, float flt_max, float* d_work, uint64_t seed, uint64_t elem_offset,
                                     void* stream);

/* The same cast, one enqueue when the tensor is small enough to stay in registers: a cooperative launch whose
 * threads keep their share of the tensor across a grid-wide barrier (read once: 8 B/elem fp32 instead of 12, no
 * memset, no atomics).  Eligible: E5M2 / E4M3 / E3M4 with RNE / STOCHASTIC / RTZ, 32-byte aligned buffers,
 * elem_offset % 16 == 0 for stochastic rounding, and n up to (resident threads x 4 vectors) — about 4.8 M float32 or
 * 9.7 M 16-bit elements on a B200; everything else silently takes fpemu_cast_amax_scaled's three-launch path, so
 * the results never depend on which path ran.  d_work: work_floats device floats, at least 2; the single-launch
 * path needs 2 + its grid size (fpemu_fused_work_floats() is always enough).  d_work[0] = amax and d_work[1] = scale
 * on return, as above.  This is the call the module hooks make thousands of times per training step
 * (mpemu/qutils.py:274-287 from :478-616). */
FPEMU_API int fpemu_fused_work_floats(void);
FPEMU_API int fpemu_cast_amax_scaled_fused(const void* in, void* out, int64_t n, int dtype, int format, int rounding,
                                 int flags, float flt_max, float* d_work, int work_floats, uint64_t seed,
                                 uint64_t elem_offset, void* stream);

/* Row-wise (per-channel) amax-scaled cast of a rows x cols row-major matrix: each row gets
 * scale = clamp(flt_max / max|row|).  One launch replaces the Python per-channel loop of
 * mpemu/qutils.py:289-311.  d_row_scale (optional, rows floats) receives the scales used. */
FPEMU_API int fpemu_cast_rowwise(const void* in, void* out, int64_t rows, int64_t cols, int dtype, int format,
                       int rounding, int flags, float flt_max, float* d_row_scale, uint64_t seed,
                       uint64_t elem_offset, void* stream);

/* ---- 8-bit wire form --------------------------------------------------------------
 * The reference defines the E5M2 / E4M3 / E3M4 bit layouts (README.md:17-27) but only ever produces the
 * de-quantised value; these two entry points materialise the byte.
 *
 * fpemu_encode: the rounding of fpemu_cast (same arguments, same random fields for stochastic rounding), result
 * stored as one code per element: E5M2 S.EEEEE.MM bias 15 (Inf 0x7C, NaN 0x7D-0x7F), E4M3 S.EEEE.MMM bias 7
 * (max 0x7E = 448, 0x7F = NaN, never produced), E3M4 S.EEE.MMMM bias 3 (max 0x7E = 30), gradual underflow.
 * format must be one of those three.  Where the reference's rounded value lies between two points of the
 * format's subnormal grid (E4M3 below 2^-6, E3M4 below 2^-2: the reference keeps full mantissa precision there,
 * fpemu_kernels.cu:594-606, 924-936) the code is the nearest grid point, ties to the even code.
 * Fast path: `in` 32-byte aligned, `codes` 8-byte (fp32 input) / 16-byte (16-bit input) aligned, elem_offset a
 * multiple of 16 when rounding is stochastic; otherwise an element-wise kernel runs.
 *
 * fpemu_decode: codes -> values of element type `dtype`, multiplied by 1/scale exactly as fpemu_cast un-scales
 * (FPEMU_FLAG_UNSCALE_FP32 honoured).  fpemu_decode(fpemu_encode(x)) equals fpemu_cast(x) bit for bit for every
 * element whose rounded value is a member of the format (always, for E5M2). */
FPEMU_API int fpemu_encode(const void* in, void* codes, int64_t n, int dtype, int format, int rounding, int flags,
                 float scale, const float* d_scale, uint64_t seed, uint64_t elem_offset, void* stream);
FPEMU_API int fpemu_decode(const void* codes, void* out, int64_t n, int dtype, int format, int flags, float scale,
                 const float* d_scale, void* stream);

/* Per-channel wire form: the rounding of fpemu_cast_rowwise (each row scaled by clamp(flt_max / max|row|), mpemu/qutils.py:
 * 289-311; FPEMU_FLAG_NO_SCALE_CLAMP for the fine-grained branch :313-337) stored as one code per element plus one float
 * per row in d_row_scale (required), and the way back.  Rows up to 16 384 float32 / 32 768 16-bit elements are encoded in
 * a single pass from registers (`in` 32-byte aligned, `codes` 8- / 16-byte aligned, cols a multiple of 8 / 16, and of 16
 * with elem_offset % 16 == 0 for stochastic rounding); anything else runs one warp per row in two passes.
 * fpemu_decode_rowwise(fpemu_encode_rowwise(x)) equals fpemu_cast_rowwise(x) wherever the rounded value is a member of
 * the format (fpemu_encode above). */
FPEMU_API int fpemu_encode_rowwise(const void* in, void* codes, int64_t rows, int64_t cols, int dtype, int format,
                         int rounding, int flags, float flt_max, float* d_row_scale, uint64_t seed,
                         uint64_t elem_offset, void* stream);
FPEMU_API int fpemu_decode_rowwise(const void* codes, void* out, int64_t rows, int64_t cols, int dtype, int format,
                         int flags, const float* d_row_scale, void* stream);

/* Row-wise cast with caller-supplied scales: row r is cast with scale d_row_scale[r] (read on the device, no host sync).
 * For per-channel scales that are not derived from the row maximum — the reference's per-channel `mean` scaling takes
 * torch.mode of each channel and calls the op once per channel (mpemu/qutils.py:289-311 with :294-301); this is that
 * loop as one launch.  Formats: every one that takes a scale (E5M2, E4M3, E3M4, E4M3_IEEE, FP4; BFLOAT16 / FLOAT16
 * return FPEMU_ERR_MODE).  Vector path: buffers 32-byte aligned, cols a multiple of 8 (float32) / 16 (16-bit); FP4
 * always runs the element-wise kernel. */
FPEMU_API int fpemu_cast_rows_scaled(const void* in, void* out, int64_t rows, int64_t cols, int dtype, int format,
                           int rounding, int flags, const float* d_row_scale, uint64_t seed, uint64_t elem_offset,
                           void* stream);

/* Host-buffer convenience used by the end-to-end measurement: casts n elements that live in HOST
 * memory (pinned for full speed) into a HOST output, staging through two caller-provided device
 * buffers of chunk_elems elements each (d_stage[0], d_stage[1]) so the H2D copy, the kernel and the
 * D2H copy of consecutive chunks overlap on the three streams given.  Synchronises the three streams
 * before returning. */
FPEMU_API int fpemu_cast_host(const void* h_in, void* h_out, int64_t n, int dtype, int format, int rounding,
                    int flags, float scale, uint64_t seed, void* d_stage0, void* d_stage1,
                    int64_t chunk_elems, void* stream_h2d, void* stream_compute, void* stream_d2h);

/* Several casts of ONE host tensor (a format x rounding sweep over the same input, with or without per-tensor
 * amax-scaled members): the input crosses the bus once.  Chunks of chunk_elems elements are uploaded into d_in (a device
 * buffer of n elements that keeps the whole input resident), every chunk is cast once per job into the two-slot
 * staging ring d_stage0 / d_stage1 (chunk_elems elements each) and copied to that job's h_out; jobs with
 * amax_scaled != 0 use scale = the scale of fpemu_scale_from_amax_dtype for max|input| (reduced chunk by chunk while the
 * other jobs run, then cast in a second pass over d_in; d_work: one device float, required when such a job exists).
 * Same streams, overlap and final synchronisation as fpemu_cast_host.  Bus traffic 1 x H2D + njobs x D2H.
 * Replaces M round trips `x.cuda(); forward(mode_i); .cpu()` (and, for amax-scaled jobs, the per-tensor branch of
 * quantize_tensor, mpemu/qutils.py:274-287) of a CPU-resident caller. */
typedef struct fpemu_host_job {
  void* h_out;       /* host output, n elements                                              */
  int format, rounding, flags;
  float scale;       /* host scale (ignored when amax_scaled)                                */
  int amax_scaled;   /* != 0: scale derived from max|input| with flt_max                     */
  float flt_max;
  uint64_t seed;     /* stochastic rounding: Philox key; element i of the tensor uses index i */
} fpemu_host_job;
FPEMU_API int fpemu_cast_host_multi(const void* h_in, int64_t n, int dtype, const fpemu_host_job* jobs, int njobs,
                                    void* d_in, void* d_stage0, void* d_stage1, float* d_work, int64_t chunk_elems,
                                    void* stream_h2d, void* stream_compute, void* stream_d2h);

/* ---- multi-tensor launches: one amax launch and one cast launch for a whole list of tensors -----------------
 * Replaces the reference's tensor-by-tensor Python loops (mpemu/qutils.py:470-475 quantize_model_weights,
 * mpemu/hybrid_emu.py:110-139 master-parameter loop, and any per-parameter sweep of a weight set).
 * Usage: fill fpemu_tensor_desc[count]; fpemu_multi_table_build() writes a table of fpemu_multi_table_bytes(count)
 * bytes into HOST memory; copy that table to the device once (it stays valid while the buffers do) and pass both
 * copies to the launches.  Buffers must be 32-byte aligned and elem_offset a multiple of 16 (FPEMU_ERR_ALIGN
 * otherwise: fall back to the per-tensor entry points); all tensors share one element type. */
typedef struct fpemu_tensor_desc {
  const void* in;        /* device input                                         */
  void* out;             /* device output (may equal in)                         */
  int64_t n;             /* elements                                             */
  uint64_t elem_offset;  /* global index of element 0 (stochastic rounding)      */
} fpemu_tensor_desc;

FPEMU_API int64_t fpemu_multi_table_bytes(int count);
FPEMU_API int fpemu_multi_table_build(const fpemu_tensor_desc* descs, int count, int dtype, void* h_table);
/* d_amax[t] = max(d_amax[t], max|tensor t|) for every t, one launch. */
FPEMU_API int fpemu_amax_multi(const void* d_table, const void* h_table, int count, float* d_amax, void* stream);
/* Cast every tensor in one launch; tensor t uses scale d_scales[t] (NULL: scale 1) and seed + t. */
FPEMU_API int fpemu_cast_multi(const void* d_table, const void* h_table, int count, int format, int rounding,
                               int flags, const float* d_scales, uint64_t seed, void* stream);

/* fpemu_encode for a whole list in one launch: same table, same CTA numbering and per-tensor seeds as
 * fpemu_cast_multi, but every descriptor's `out` is a byte buffer of n codes (32-byte aligned like the rest).
 * format must be E5M2, E4M3 or E3M4. */
FPEMU_API int fpemu_encode_multi(const void* d_table, const void* h_table, int count, int format, int rounding, int flags,
                       const float* d_scales, uint64_t seed, void* stream);

/* Per-tensor amax-scaled cast that needs NO scratch memory from the caller and allocates none.
 * (No counterpart in the reference: its Python layer computes the maximum with torch and passes a float.)
 * Same result as fpemu_amax + fpemu_scale_from_amax + fpemu_cast with a device scalar, but the running
 * maximum and the scale live in a 65 792-byte `static __device__` array inside the library, which every GPU
 * that loads the library keeps until the process exits.  Meant for callers that cannot allocate device memory
 * at all: a plugin inside a server whose pool was sized at start-up and belongs to the host framework, code
 * that was handed its two buffers and owns no allocator, loops in which cudaMalloc is forbidden because it
 * synchronises the device and can fail.  The caller still passes `in` and `out` (in == out allowed).
 *   - Limit: at most 512 MiB of input per call (2^27 float32 or 2^28 float16/bfloat16 elements); more returns
 *     FPEMU_ERR_TOO_LARGE and launches nothing.  BFLOAT16 / FLOAT16 targets take no scale: FPEMU_ERR_MODE.
 *   - The first call on each device looks up the array's device address (cudaGetSymbolAddress); make that
 *     first call before any stream capture starts.
 *   - All calls on a GPU share the array.  Nothing needs clearing between calls (each call overwrites every
 *     slot it later reads), but calls that overlap on the GPU race: keep them on one stream or order them with
 *     events.
 *   - Calls no cudaMalloc / cudaFree and no allocating library routine. */
FPEMU_API int fpemu_cast_amax_scaled_noalloc(const void* in, void* out, int64_t n, int dtype, int format,
                                             int rounding, int flags, float flt_max, uint64_t seed,
                                             uint64_t elem_offset, void* stream);

/* The amax-scaled sweep of a tensor list in two launches: fpemu_amax_multi, then one of these, which take the vector of
 * maxima itself (d_amax[t], after the caller's all-reduce when the tensors are row-sharded) and derive tensor t's
 * scale inside the cast kernel exactly as fpemu_scale_from_amax_dtype does for the table's element type
 * (mpemu/qutils.py:283-285); tensor t's scale is stored in d_scales_out[t] (optional).  Replaces the same Python loops
 * as fpemu_cast_multi with the per-tensor `max` scaling of mpemu/qutils.py:274-287 applied to each tensor. */
FPEMU_API int fpemu_cast_multi_amax(const void* d_table, const void* h_table, int count, int format, int rounding,
                                    int flags, const float* d_amax, float flt_max, float* d_scales_out, uint64_t seed,
                                    void* stream);
FPEMU_API int fpemu_encode_multi_amax(const void* d_table, const void* h_table, int count, int format, int rounding,
                                      int flags, const float* d_amax, float flt_max, float* d_scales_out, uint64_t seed,
                                      void* stream);

/* ---- the exchange step of a row-sharded sweep, fused into the two kernels that surround it ----------------------------
 * when the tensors of a list are row-sharded over the GPUs of one NVLink domain, the only communication
 * is an all-reduce(MAX) of the float[count] vector of maxima between the amax launch and the cast launch.  These entry
 * points remove the collective launch: every rank owns an exchange buffer of fpemu_exchange_floats(count, world)
 * floats (8-byte aligned) that all ranks can address (peer-mapped / symmetric memory; zero it once, before the first
 * sweep, and make sure every rank has done so).  fpemu_amax_multi_push reduces like fpemu_amax_multi into one half of
 * d_amax2 (2 x count floats, zero before the first call; the halves alternate with the epoch and the call re-zeroes
 * the one the next sweep uses), and a small launch behind the reduction stores every maximum, paired with the sweep's
 * epoch in one 8-byte word, into EVERY rank's buffer over NVLink.  The epoch travels with the value, so there is no
 * separate arrival flag and no fence; the buffer holds several copies of the words so that the readers' polls spread
 * over the L2 slices.
 * fpemu_cast_multi_xchg / fpemu_encode_multi_xchg are fpemu_cast_multi_amax / fpemu_encode_multi_amax whose warps poll
 * their tensor's `world` words until they carry the sweep's epoch and take the maximum over ranks themselves.
 *   peer_bufs  host array of `world` pointers: rank r's exchange buffer as mapped into THIS process
 *   my_buf     this rank's own exchange buffer
 *   epoch      sweep number, the same on every rank, starting at 1 and incremented per sweep (its parity selects one
 *              of two halves of the buffer, so a rank may run one sweep ahead of its peers)
 * A rank that never arrives would stall its peers: after about 10 s the waiting kernel gives up and poisons that rank's
 * maxima with NaN for the epoch, so the launch ends (with NaN scales) instead of hanging the GPU.  No counterpart in
 * the reference (its only multi-GPU code is stock DDP in an example script); the results equal the all-reduce path bit
 * for bit.  With d_scales_out, the scales of tensors that have no elements on this rank are stored as well (all three
 * of fpemu_*_multi_amax / _xchg). */
FPEMU_API int64_t fpemu_exchange_floats(int count, int world);
FPEMU_API int fpemu_amax_multi_push(const void* d_table, const void* h_table, int count, float* d_amax2,
                                    void* const* peer_bufs, int world, int rank, uint32_t epoch, void* stream);
FPEMU_API int fpemu_cast_multi_xchg(const void* d_table, const void* h_table, int count, int format, int rounding,
                                    int flags, const void* my_buf, int world, uint32_t epoch, float flt_max,
                                    float* d_scales_out, uint64_t seed, void* stream);
FPEMU_API int fpemu_encode_multi_xchg(const void* d_table, const void* h_table, int count, int format, int rounding,
                                      int flags, const void* my_buf, int world, uint32_t epoch, float flt_max,
                                      float* d_scales_out, uint64_t seed, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* FPEMU_B200_H */
