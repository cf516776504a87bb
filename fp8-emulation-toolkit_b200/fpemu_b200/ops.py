"""Tensor-level operations over the C ABI.  PyTorch is plumbing here (device memory, streams); every
number is produced by the CUDA kernels of libfpemu_b200.so."""
import itertools

import torch

from . import _lib as L

_DT = {torch.float32: L.DT_F32, torch.float16: L.DT_F16, torch.bfloat16: L.DT_BF16}

# flt_max per format, mpemu/qutils.py:100-114
FLT_MAX = {L.FMT_E5M2: 57344.0, L.FMT_E4M3: 448.0, L.FMT_E4M3_IEEE: 240.0, L.FMT_E3M4: 30.0, L.FMT_FP4: 1.0}

_seed_base = 42           # the reference seeds torch with 42 at import (cuda/fpemu.py:19); we keep our own stream
_seed_counter = itertools.count()


def manual_seed(seed):
    """Reset the stochastic-rounding seed sequence (one fresh Philox key per stochastic call)."""
    global _seed_base, _seed_counter
    _seed_base = int(seed) & 0xFFFFFFFF
    _seed_counter = itertools.count()


def next_seed():
    return (_seed_base << 32) | (next(_seed_counter) & 0xFFFFFFFF)


_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)
_device_seeds = {}


def enable_device_seed(device=None, value=0):
    """Install a device-resident 64-bit word that every stochastic kernel on `device` adds to its seed
    (fpemu_set_seed_source).  With `advance_device_seed()` captured in the same CUDA graph as the casts, every replay
    of the graph draws fresh random bits; without a graph it is simply a seed offset that lives on the GPU.
    Returns the int64 tensor holding the word: a captured graph keeps its address, so set a new value with
    `word.fill_(v)` rather than calling this again."""
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    word = torch.tensor([int(value)], dtype=torch.int64, device=dev)
    with torch.cuda.device(dev):
        L.check(L.lib().fpemu_set_seed_source(word.data_ptr()), "fpemu_set_seed_source")
    _device_seeds[dev.index] = word                      # keeps the word alive
    return word


def disable_device_seed(device=None):
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    with torch.cuda.device(dev):
        L.check(L.lib().fpemu_set_seed_source(None), "fpemu_set_seed_source")
    _device_seeds.pop(dev.index, None)


def advance_device_seed(device=None):
    """Enqueue (on the current stream) the one-thread kernel that moves the device seed word on."""
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    if dev.index not in _device_seeds:
        raise RuntimeError("fpemu: no device seed installed on %s (enable_device_seed first)" % dev)
    with torch.cuda.device(dev):
        st = L.lib().fpemu_advance_seed_source(_raw_stream(dev.index) if _raw_stream is not None
                                               else torch.cuda.current_stream(dev).cuda_stream)
    L.check(st, "fpemu_advance_seed_source")


def _dt(t):
    try:
        return _DT[t.dtype]
    except KeyError:
        raise TypeError("fpemu: unsupported dtype %s (float32, float16, bfloat16)" % t.dtype)


def _check_tensor(t, name="input"):
    if not t.is_cuda:
        raise RuntimeError("%s must be a CUDA tensor" % name)           # CHECK_CUDA, cuda/fpemu_impl.cpp:23
    if not t.is_contiguous():
        raise RuntimeError("%s must be contiguous" % name)              # CHECK_CONTIGUOUS, :24


def _check_out(out, x, n):
    """A caller-supplied output: CUDA, contiguous, same device and dtype as the input, room for n elements."""
    _check_tensor(out, "out")
    if out.dtype != x.dtype or out.device != x.device or out.numel() < n:
        raise ValueError("fpemu: out must match the input's dtype and device and hold at least %d elements" % n)


def _stream(t):
    """cudaStream_t (as int) of torch's current stream on t's device."""
    if _raw_stream is not None:
        return _raw_stream(t.device.index if t.device.index is not None else torch.cuda.current_device())
    return torch.cuda.current_stream(t.device).cuda_stream


class _on_device(object):
    """Make t's device current for the duration of a launch (cheap no-op when it already is)."""
    def __init__(self, t):
        self.idx = t.device.index
        self.prev = None

    def __enter__(self):
        cur = torch.cuda.current_device()
        if self.idx is not None and cur != self.idx:
            self.prev = cur
            torch.cuda.set_device(self.idx)

    def __exit__(self, *exc):
        if self.prev is not None:
            torch.cuda.set_device(self.prev)


def empty_like_coaligned(x):
    """torch.empty_like(x), except that for a view starting off a 32-byte boundary the result starts off by the same
    amount (it is then a view into a slightly larger allocation): input and output misaligned alike is the case the
    vector kernels handle at full speed, an aligned output for a misaligned input is not."""
    mis = x.data_ptr() & 31
    esz = x.element_size()
    if mis == 0 or mis % esz or not x.is_contiguous():
        return torch.empty_like(x)
    k = mis // esz
    buf = torch.empty(x.numel() + 32 // esz, dtype=x.dtype, device=x.device)
    k = (k - (buf.data_ptr() & 31) // esz) % (32 // esz)
    return buf[k:k + x.numel()].view(x.shape)


def cast(x, mode, out=None, scale=1.0, d_scale=None, seed=None, elem_offset=0, n=None, unscale_fp32=False):
    """QDQ-cast `x` (CUDA, contiguous) into `out` (default: new tensor; pass out=x for in-place).
    `d_scale`: optional 1-element float32 CUDA tensor used instead of `scale` (no host sync)."""
    _check_tensor(x)
    fmt, rnd, flags = L.parse_mode(mode)
    if unscale_fp32:
        flags |= L.FLAG_UNSCALE_FP32
    if out is None:
        out = empty_like_coaligned(x)
    else:
        _check_tensor(out, "out")
        if out.dtype != x.dtype or out.numel() < (x.numel() if n is None else n):
            raise ValueError("fpemu: out must match input dtype and size")
    n = x.numel() if n is None else int(n)
    if n > x.numel():
        raise ValueError("fpemu: size larger than the tensor")
    if rnd == L.RND_STOCHASTIC and seed is None:
        seed = next_seed()
    dptr = None
    if d_scale is not None:
        if not (d_scale.is_cuda and d_scale.dtype == torch.float32 and d_scale.numel() >= 1):
            raise ValueError("fpemu: d_scale must be a float32 CUDA tensor")
        dptr = d_scale.data_ptr()
    with _on_device(x):
        if x.dtype == torch.float64:                  # the one entry point that takes doubles (fpemu_kernels.cu:1059)
            st = L.lib().fpemu_cast_f64(x.data_ptr(), out.data_ptr(), n, fmt, rnd, flags, float(scale), dptr,
                                        int(seed or 0), int(elem_offset), _stream(x))
        else:
            st = L.lib().fpemu_cast(x.data_ptr(), out.data_ptr(), n, _dt(x), fmt, rnd, flags, float(scale), dptr,
                                    int(seed or 0), int(elem_offset), _stream(x))
    L.check(st, "fpemu_cast(%s)" % mode)
    return out


def _d_scale_ptr(d_scale):
    if d_scale is None:
        return None
    if not (d_scale.is_cuda and d_scale.dtype == torch.float32 and d_scale.numel() >= 1):
        raise ValueError("fpemu: d_scale must be a float32 CUDA tensor")
    return d_scale.data_ptr()


def encode(x, mode, out=None, scale=1.0, d_scale=None, seed=None, elem_offset=0, n=None, expect_normals=False):
    """Round `x` like `cast(x, mode, ...)` and return the format's one-byte codes (uint8 tensor of x's shape):
    E5M2 / E4M3 / E3M4 bit layouts of the reference's format table (README.md:17-27).  `expect_normals`: performance
    hint for E4M3_RNE that scaled values below 2^-6 are rare (amax-derived scales); never changes a code."""
    _check_tensor(x)
    fmt, rnd, flags = L.parse_mode(mode)
    if expect_normals:
        flags |= L.FLAG_EXPECT_NORMALS
    n = x.numel() if n is None else int(n)
    if n > x.numel():
        raise ValueError("fpemu: size larger than the tensor")
    if out is None:
        out = torch.empty(x.shape, dtype=torch.uint8, device=x.device)
    elif not (out.is_cuda and out.dtype == torch.uint8 and out.is_contiguous() and out.numel() >= n):
        raise ValueError("fpemu: codes must be a contiguous uint8 CUDA tensor of at least n elements")
    if rnd == L.RND_STOCHASTIC and seed is None:
        seed = next_seed()
    with _on_device(x):
        st = L.lib().fpemu_encode(x.data_ptr(), out.data_ptr(), n, _dt(x), fmt, rnd, flags, float(scale),
                                  _d_scale_ptr(d_scale), int(seed or 0), int(elem_offset), _stream(x))
    L.check(st, "fpemu_encode(%s)" % mode)
    return out


def decode(codes, mode, dtype=torch.float32, out=None, scale=1.0, d_scale=None, n=None, unscale_fp32=False):
    """Codes written by `encode` -> values of `dtype`, divided by the scale exactly as `cast` un-scales.
    `mode` may be a full mode string or just the format name ("E4M3")."""
    if not (codes.is_cuda and codes.dtype == torch.uint8 and codes.is_contiguous()):
        raise ValueError("fpemu: codes must be a contiguous uint8 CUDA tensor")
    fmt = {"E5M2": L.FMT_E5M2, "E4M3": L.FMT_E4M3, "E3M4": L.FMT_E3M4}.get(mode)
    if fmt is None:
        fmt = L.parse_mode(mode)[0]
    flags = 0
    if unscale_fp32:
        flags |= L.FLAG_UNSCALE_FP32
    n = codes.numel() if n is None else int(n)
    if n > codes.numel():
        raise ValueError("fpemu: size larger than the tensor")
    if out is None:
        out = torch.empty(codes.shape, dtype=dtype, device=codes.device)
    else:
        _check_tensor(out, "out")
        if out.numel() < n:
            raise ValueError("fpemu: out too small")
    with _on_device(codes):
        st = L.lib().fpemu_decode(codes.data_ptr(), out.data_ptr(), n, _dt(out), fmt, flags, float(scale),
                                  _d_scale_ptr(d_scale), _stream(codes))
    L.check(st, "fpemu_decode(%s)" % mode)
    return out


def cast_blocked(x, mode, block_size, out=None, seed=None, elem_offset=0, n=None, unscale_fp32=False):
    _check_tensor(x)
    fmt, rnd, flags = L.parse_mode(mode)
    if unscale_fp32:
        flags |= L.FLAG_UNSCALE_FP32
    n = x.numel() if n is None else int(n)
    if n > x.numel():
        raise ValueError("fpemu: size larger than the tensor")
    if out is None:
        out = torch.empty_like(x)
    else:
        _check_out(out, x, n)
    if rnd == L.RND_STOCHASTIC and seed is None:
        seed = next_seed()
    if int(block_size) == n and n > (1 << 16) and fmt in (L.FMT_E5M2, L.FMT_E4M3, L.FMT_E3M4, L.FMT_E4M3_IEEE) \
            and not (flags & L.FLAG_DAZ):
        # one block = the whole (large) tensor (the reference CPU op's reading of block_size == size,
        # cpp/fpemu_impl.cpp:778-779): amax kernel, then the power-of-two block scale derived on the device from
        # the exponent field of the maximum (fpemu_kernels.cu:409-413, 577-581, 909-911), then the plain cast
        work = torch.zeros(2, dtype=torch.float32, device=x.device)
        amax(x, out=work[0:1], n=n)
        with _on_device(x):
            st = L.lib().fpemu_block_scale_from_amax(work.data_ptr(), fmt, _dt(x), work.data_ptr() + 4, _stream(x))
        L.check(st, "fpemu_block_scale_from_amax")
        return cast(x, mode, out=out, d_scale=work[1:2], seed=seed, elem_offset=elem_offset, n=n,
                    unscale_fp32=unscale_fp32)
    with _on_device(x):
        st = L.lib().fpemu_cast_blocked(x.data_ptr(), out.data_ptr(), n, int(block_size), _dt(x), fmt, rnd, flags,
                                        int(seed or 0), int(elem_offset), _stream(x))
    L.check(st, "fpemu_cast_blocked(%s, %d)" % (mode, block_size))
    return out


def amax(x, out=None, n=None):
    """max|x| as a 1-element float32 CUDA tensor, max-accumulated into `out` when given (no host sync)."""
    _check_tensor(x)
    if out is None:
        out = torch.zeros(1, dtype=torch.float32, device=x.device)
    n = x.numel() if n is None else int(n)
    with _on_device(x):
        st = L.lib().fpemu_amax(x.data_ptr(), n, _dt(x), out.data_ptr(), _stream(x))
    L.check(st, "fpemu_amax")
    return out


def scale_from_amax(d_amax, flt_max, out=None, dtype=torch.float32):
    """The scale quantize_tensor derives from a maximum (`flt_max / vmax` + clamp, qutils.py:283-285), on the device,
    elementwise over a float32 CUDA vector.  `dtype`: element type of the tensors the maxima came from — torch
    evaluates `flt_max / vmax` as `vmax.reciprocal() * flt_max` in that type (two roundings in half for float16)."""
    if out is None:
        out = torch.empty_like(d_amax)
    with _on_device(d_amax):
        st = L.lib().fpemu_scale_from_amax_dtype(d_amax.data_ptr(), d_amax.numel(), float(flt_max), _DT[dtype],
                                                 out.data_ptr(), _stream(d_amax))
    L.check(st, "fpemu_scale_from_amax")
    return out


def mean_scale(x, flt_min, mean_floor=1e-5, out=None):
    """Per-tensor `mean` scale of quantize_tensor (qutils.py:276-281) as a 1-element float32 CUDA tensor:
    max(1, flt_min / mean|x|) with the reference's floor on the mean; two launches, no host sync."""
    _check_tensor(x)
    if out is None:
        out = torch.empty(1, dtype=torch.float32, device=x.device)
    stream = _stream(x)
    work = _work_for(x, stream)                       # >= 8 KiB, 512-byte aligned: room for the partial sums
    with _on_device(x):
        st = L.lib().fpemu_scale_from_abs_mean(x.data_ptr(), x.numel(), _dt(x), float(flt_min), float(mean_floor),
                                               work.data_ptr() + 8, out.data_ptr(), stream)
    L.check(st, "fpemu_scale_from_abs_mean")
    return out


_work_cache = {}


_work_floats = None


def _work_size():
    global _work_floats
    if _work_floats is None:
        _work_floats = int(L.lib().fpemu_fused_work_floats())
    return _work_floats


def _work_for(x, stream):
    """Scratch floats (running maximum, scale, per-CTA maxima of the single-launch path) per (device, stream): calls
    on one stream execute in order, so they can share the scratch; a caller that wants to keep the scale gets a
    fresh buffer instead."""
    key = (x.device.index, stream)
    w = _work_cache.get(key)
    if w is None:
        w = torch.zeros(_work_size(), dtype=torch.float32, device=x.device)
        _work_cache[key] = w
    return w


def cast_amax_scaled(x, mode, out=None, flt_max=None, seed=None, elem_offset=0, unscale_fp32=False,
                     d_amax=None, return_scale=False, fused=True):
    """Per-tensor amax-scaled cast without a host sync (the `per-tensor` / `max` branch of quantize_tensor,
    qutils.py:274-287).  One library call: a single cooperative launch for tensors that fit in registers (about
    4.8 M fp32 elements), else memset + amax kernel + cast kernel, on the current stream; `fused=False` forces the
    three-launch path (same results).
    `d_amax` lets a caller supply an already reduced (e.g. all-reduced across ranks) amax; then only the scale
    and cast kernels run."""
    fmt, rnd, flags = L.parse_mode(mode)
    if flt_max is None:
        flt_max = FLT_MAX[fmt]
    if d_amax is not None:
        d_scale = scale_from_amax(d_amax, flt_max, dtype=x.dtype)
        y = cast(x, mode, out=out, d_scale=d_scale, seed=seed, elem_offset=elem_offset, unscale_fp32=unscale_fp32)
        return (y, d_scale) if return_scale else y
    _check_tensor(x)
    if unscale_fp32:
        flags |= L.FLAG_UNSCALE_FP32
    if out is None:
        out = torch.empty_like(x)
    else:
        _check_tensor(out, "out")
        if out.dtype != x.dtype or out.numel() < x.numel():
            raise ValueError("fpemu: out must match input dtype and size")
    if rnd == L.RND_STOCHASTIC and seed is None:
        seed = next_seed()
    stream = _stream(x)
    work = torch.empty(_work_size(), dtype=torch.float32, device=x.device) if return_scale else _work_for(x, stream)
    with _on_device(x):
        if fused:
            st = L.lib().fpemu_cast_amax_scaled_fused(x.data_ptr(), out.data_ptr(), x.numel(), _dt(x), fmt, rnd, flags,
                                                      float(flt_max), work.data_ptr(), work.numel(), int(seed or 0),
                                                      int(elem_offset), stream)
        else:
            st = L.lib().fpemu_cast_amax_scaled(x.data_ptr(), out.data_ptr(), x.numel(), _dt(x), fmt, rnd, flags,
                                                float(flt_max), work.data_ptr(), int(seed or 0), int(elem_offset), stream)
    L.check(st, "fpemu_cast_amax_scaled(%s)" % mode)
    return (out, work[1:2]) if return_scale else out


def cast_amax_scaled_noalloc(x, mode, out, flt_max=None, seed=None, elem_offset=0, unscale_fp32=False):
    """Per-tensor amax-scaled cast for callers that cannot allocate: `out` must be supplied (pass `x` for
    in-place) and the library keeps the running maximum and the scale in its own static device array, so this
    call allocates nothing on the device.  At most 512 MiB of input; calls sharing a GPU must not overlap
    (see include/fpemu_b200.h).  Same numbers as `cast_amax_scaled`."""
    _check_tensor(x)
    _check_tensor(out, "out")
    if out.dtype != x.dtype or out.numel() < x.numel():
        raise ValueError("fpemu: out must match input dtype and size")
    fmt, rnd, flags = L.parse_mode(mode)
    if unscale_fp32:
        flags |= L.FLAG_UNSCALE_FP32
    if flt_max is None:
        flt_max = FLT_MAX[fmt]
    if rnd == L.RND_STOCHASTIC and seed is None:
        seed = next_seed()
    with _on_device(x):
        st = L.lib().fpemu_cast_amax_scaled_noalloc(x.data_ptr(), out.data_ptr(), x.numel(), _dt(x), fmt, rnd, flags,
                                                    float(flt_max), int(seed or 0), int(elem_offset), _stream(x))
    L.check(st, "fpemu_cast_amax_scaled_noalloc(%s)" % mode)
    return out


def cast_rowwise(x2d, mode, out=None, flt_max=None, clamp=True, seed=None, elem_offset=0, unscale_fp32=False,
                 return_scales=False, rows=None, cols=None):
    """Per-channel (dim-0) amax-scaled cast in ONE launch (replaces the loop of qutils.py:289-311)."""
    _check_tensor(x2d)
    fmt, rnd, flags = L.parse_mode(mode)
    if unscale_fp32:
        flags |= L.FLAG_UNSCALE_FP32
    if not clamp:
        flags |= L.FLAG_NO_SCALE_CLAMP
    if flt_max is None:
        flt_max = FLT_MAX[fmt]
    if rows is None:
        rows = x2d.shape[0] if x2d.dim() > 0 else 1
        cols = x2d.numel() // max(rows, 1)
    if rows * cols > x2d.numel():
        raise ValueError("fpemu: rows * cols larger than the tensor")
    if out is None:
        out = torch.empty_like(x2d)
    else:
        _check_out(out, x2d, rows * cols)
    scales = torch.empty(rows, dtype=torch.float32, device=x2d.device) if return_scales else None
    if rnd == L.RND_STOCHASTIC and seed is None:
        seed = next_seed()
    with _on_device(x2d):
        st = L.lib().fpemu_cast_rowwise(x2d.data_ptr(), out.data_ptr(), int(rows), int(cols), _dt(x2d), fmt, rnd,
                                        flags, float(flt_max), scales.data_ptr() if scales is not None else None,
                                        int(seed or 0), int(elem_offset), _stream(x2d))
    L.check(st, "fpemu_cast_rowwise(%s)" % mode)
    return (out, scales) if return_scales else out


def cast_rows_scaled(x2d, mode, scales, out=None, seed=None, elem_offset=0, unscale_fp32=False, rows=None, cols=None):
    """Row r of `x2d` cast with scale `scales[r]` (float32 CUDA vector, read on the device): one launch for a whole
    per-channel loop whose scales are not the row maxima."""
    _check_tensor(x2d)
    fmt, rnd, flags = L.parse_mode(mode)
    if unscale_fp32:
        flags |= L.FLAG_UNSCALE_FP32
    if rows is None:
        rows = x2d.shape[0] if x2d.dim() > 0 else 1
        cols = x2d.numel() // max(rows, 1)
    if not (scales.is_cuda and scales.dtype == torch.float32 and scales.is_contiguous() and scales.numel() >= rows):
        raise ValueError("fpemu: scales must be a contiguous float32 CUDA tensor with one entry per row")
    if out is None:
        out = torch.empty_like(x2d)
    else:
        _check_tensor(out, "out")
        if out.dtype != x2d.dtype or out.numel() < rows * cols:
            raise ValueError("fpemu: out must match input dtype and size")
    if rnd == L.RND_STOCHASTIC and seed is None:
        seed = next_seed()
    with _on_device(x2d):
        st = L.lib().fpemu_cast_rows_scaled(x2d.data_ptr(), out.data_ptr(), int(rows), int(cols), _dt(x2d), fmt, rnd, flags,
                                            scales.data_ptr(), int(seed or 0), int(elem_offset), _stream(x2d))
    L.check(st, "fpemu_cast_rows_scaled(%s)" % mode)
    return out


def encode_rowwise(x2d, mode, codes=None, scales=None, flt_max=None, clamp=True, seed=None, elem_offset=0, rows=None,
                   cols=None):
    """Per-channel wire form: (codes uint8 of x2d's shape, scales float32[rows]) such that
    `decode_rowwise(codes, scales, ...)` reproduces `cast_rowwise(x2d, mode)`; one launch, each row read once."""
    _check_tensor(x2d)
    fmt, rnd, flags = L.parse_mode(mode)
    if not clamp:
        flags |= L.FLAG_NO_SCALE_CLAMP
    if flt_max is None:
        flt_max = FLT_MAX[fmt]
    if rows is None:
        rows = x2d.shape[0] if x2d.dim() > 0 else 1
        cols = x2d.numel() // max(rows, 1)
    if codes is None:
        codes = torch.empty(x2d.shape, dtype=torch.uint8, device=x2d.device)
    elif not (codes.is_cuda and codes.dtype == torch.uint8 and codes.is_contiguous() and codes.numel() >= rows * cols):
        raise ValueError("fpemu: codes must be a contiguous uint8 CUDA tensor of rows * cols elements")
    if scales is None:
        scales = torch.empty(rows, dtype=torch.float32, device=x2d.device)
    elif not (scales.is_cuda and scales.dtype == torch.float32 and scales.is_contiguous() and scales.numel() >= rows):
        raise ValueError("fpemu: scales must be a contiguous float32 CUDA tensor with one entry per row")
    if rnd == L.RND_STOCHASTIC and seed is None:
        seed = next_seed()
    with _on_device(x2d):
        st = L.lib().fpemu_encode_rowwise(x2d.data_ptr(), codes.data_ptr(), int(rows), int(cols), _dt(x2d), fmt, rnd, flags,
                                          float(flt_max), scales.data_ptr(), int(seed or 0), int(elem_offset), _stream(x2d))
    L.check(st, "fpemu_encode_rowwise(%s)" % mode)
    return codes, scales


def decode_rowwise(codes, scales, mode, dtype=torch.float32, out=None, unscale_fp32=False, rows=None, cols=None):
    """Codes + per-row scales -> values of `dtype` (what `cast_rowwise` returns)."""
    if not (codes.is_cuda and codes.dtype == torch.uint8 and codes.is_contiguous()):
        raise ValueError("fpemu: codes must be a contiguous uint8 CUDA tensor")
    fmt = {"E5M2": L.FMT_E5M2, "E4M3": L.FMT_E4M3, "E3M4": L.FMT_E3M4}.get(mode)
    if fmt is None:
        fmt = L.parse_mode(mode)[0]
    if rows is None:
        rows = codes.shape[0] if codes.dim() > 0 else 1
        cols = codes.numel() // max(rows, 1)
    if not (scales.is_cuda and scales.dtype == torch.float32 and scales.is_contiguous() and scales.numel() >= rows):
        raise ValueError("fpemu: scales must be a contiguous float32 CUDA tensor with one entry per row")
    if out is None:
        out = torch.empty(codes.shape, dtype=dtype, device=codes.device)
    else:
        _check_tensor(out, "out")
        if out.numel() < rows * cols:
            raise ValueError("fpemu: out too small")
    with _on_device(codes):
        st = L.lib().fpemu_decode_rowwise(codes.data_ptr(), out.data_ptr(), int(rows), int(cols), _dt(out), fmt,
                                          L.FLAG_UNSCALE_FP32 if unscale_fp32 else 0, scales.data_ptr(), _stream(codes))
    L.check(st, "fpemu_decode_rowwise(%s)" % mode)
    return out


def split_runs(sizes, depth):
    """Cut a list of tensor sizes into `depth` consecutive runs [(first, one past last)] of about equal total size
    (every run gets at least one tensor)."""
    count, total = len(sizes), sum(sizes)
    depth = max(1, min(int(depth), count))
    runs, t0, acc, k = [], 0, 0, 1
    for t, n in enumerate(sizes):
        acc += n
        left = count - (t + 1)
        if t == count - 1 or (k < depth and ((acc * depth >= total * k and left >= depth - k) or left == depth - k)):
            runs.append((t0, t + 1))
            t0, k = t + 1, k + 1
    return runs


def pipelined_sweep(runs, amax_fn, reduce_fn, apply_fn):
    """The schedule of an amax-scaled sweep whose maxima need an exchange step: for every run, reduce locally
    (amax_fn) and start its exchange (reduce_fn may return an object with .wait(), i.e. an asynchronous collective);
    then, run by run, wait for the exchange and apply (apply_fn).  Run i's exchange therefore executes while run
    i+1 is being reduced and run i-1 is being cast."""
    pending = []
    for t0, t1 in runs:
        amax_fn(t0, t1)
        pending.append(reduce_fn(t0, t1) if reduce_fn is not None else None)
    for (t0, t1), work in zip(runs, pending):
        if work is not None:
            work.wait()                  # stream-ordered for NCCL work objects: the host does not block
        apply_fn(t0, t1)


class MultiTensorPlan(object):
    """A fixed list of CUDA tensors (same dtype, same device) that is reduced / cast with ONE launch each.

    Build it once for a parameter list or a set of row slabs (the table holds raw pointers, so the tensors must
    stay alive and must not be re-allocated), then call `amax`, `cast` or `amax_scaled_cast` as often as needed.
    Raises ValueError when a tensor is not 32-byte aligned or an offset is not a multiple of 16 — use the
    per-tensor functions for such lists."""

    def __init__(self, ins, outs=None, elem_offsets=None):
        import ctypes
        ins = list(ins)
        if not ins:
            raise ValueError("fpemu: empty tensor list")
        outs = ins if outs is None else list(outs)
        offs = [0] * len(ins) if elem_offsets is None else [int(o) for o in elem_offsets]
        dev, dt = ins[0].device, ins[0].dtype
        for x, y in zip(ins, outs):
            _check_tensor(x)
            _check_tensor(y, "out")
            if x.device != dev or x.dtype != dt or y.dtype not in (dt, torch.uint8) or y.numel() < x.numel():
                raise ValueError("fpemu: multi-tensor lists must share device and dtype, outputs must match inputs")
        # outputs of dtype uint8 make this a plan for `encode` (one byte code per element) instead of `cast`
        self.packed = all(y.dtype == torch.uint8 for y in outs)
        if not self.packed and any(y.dtype != dt for y in outs):
            raise ValueError("fpemu: outputs must be all of the input dtype (cast) or all uint8 (encode)")
        self.ins, self.outs, self.count, self.device, self.dtype = ins, outs, len(ins), dev, dt
        self.offsets = offs

        class Desc(ctypes.Structure):
            _fields_ = [("inp", ctypes.c_void_p), ("out", ctypes.c_void_p), ("n", ctypes.c_int64),
                        ("off", ctypes.c_uint64)]
        descs = (Desc * self.count)()
        for i, (x, y, o) in enumerate(zip(ins, outs, offs)):
            descs[i].inp, descs[i].out, descs[i].n, descs[i].off = x.data_ptr(), y.data_ptr(), x.numel(), o
        nbytes = int(L.lib().fpemu_multi_table_bytes(self.count))
        self.h_table = torch.empty(nbytes, dtype=torch.uint8)
        st = L.lib().fpemu_multi_table_build(ctypes.addressof(descs), self.count, _DT[dt], self.h_table.data_ptr())
        L.check(st, "fpemu_multi_table_build")
        self.d_table = self.h_table.to(dev)
        self.elements = sum(x.numel() for x in ins)
        self._amax_vec = None          # float32[count] reused by the amax-scaled sweeps
        self._parts = {}               # pipeline depth -> {(first tensor, one past the last): sub-plan}

    def amax(self, out=None):
        """float32[count] vector of max|tensor t| (max-accumulated into `out` when given), one launch."""
        if out is None:
            out = torch.zeros(self.count, dtype=torch.float32, device=self.device)
        with _on_device(self.d_table):
            st = L.lib().fpemu_amax_multi(self.d_table.data_ptr(), self.h_table.data_ptr(), self.count, out.data_ptr(),
                                          _stream(self.d_table))
        L.check(st, "fpemu_amax_multi")
        return out

    def _scales_ptr(self, d_scales):
        if d_scales is None:
            return None
        if not (d_scales.is_cuda and d_scales.dtype == torch.float32 and d_scales.numel() >= self.count):
            raise ValueError("fpemu: d_scales must be a float32 CUDA vector with one entry per tensor")
        return d_scales.data_ptr()

    def encode(self, mode, d_scales=None, seed=None, expect_normals=False):
        """Round every tensor like `cast` and write one-byte codes into the plan's uint8 outputs, one launch."""
        if not self.packed:
            raise ValueError("fpemu: this plan's outputs are not uint8 code buffers")
        fmt, rnd, flags = L.parse_mode(mode)
        if expect_normals:
            flags |= L.FLAG_EXPECT_NORMALS
        if rnd == L.RND_STOCHASTIC and seed is None:
            seed = next_seed()
        with _on_device(self.d_table):
            st = L.lib().fpemu_encode_multi(self.d_table.data_ptr(), self.h_table.data_ptr(), self.count, fmt, rnd, flags,
                                            self._scales_ptr(d_scales), int(seed or 0), _stream(self.d_table))
        L.check(st, "fpemu_encode_multi(%s)" % mode)
        return self.outs

    def cast(self, mode, d_scales=None, seed=None, unscale_fp32=False):
        """Cast every tensor in one launch; tensor t uses d_scales[t] (None: scale 1) and seed + t."""
        if self.packed:
            raise ValueError("fpemu: this plan's outputs are uint8 code buffers: use encode()")
        fmt, rnd, flags = L.parse_mode(mode)
        if unscale_fp32:
            flags |= L.FLAG_UNSCALE_FP32
        if rnd == L.RND_STOCHASTIC and seed is None:
            seed = next_seed()
        with _on_device(self.d_table):
            st = L.lib().fpemu_cast_multi(self.d_table.data_ptr(), self.h_table.data_ptr(), self.count, fmt, rnd, flags,
                                          self._scales_ptr(d_scales), int(seed or 0), _stream(self.d_table))
        L.check(st, "fpemu_cast_multi(%s)" % mode)
        return self.outs

    def _from_amax(self, mode, d_amax, flt_max, d_scales_out, seed, flags_extra=0):
        """The second launch of an amax-scaled sweep: cast (or encode) every tensor with the scale the kernel derives
        from d_amax[t] (no separate scale launch); the scales used land in d_scales_out."""
        fmt, rnd, flags = L.parse_mode(mode)
        fn = L.lib().fpemu_encode_multi_amax if self.packed else L.lib().fpemu_cast_multi_amax
        with _on_device(self.d_table):
            st = fn(self.d_table.data_ptr(), self.h_table.data_ptr(), self.count, fmt, rnd, flags | flags_extra,
                    d_amax.data_ptr(), float(flt_max), d_scales_out.data_ptr(), int(seed) & 0xFFFFFFFFFFFFFFFF,
                    _stream(self.d_table))
        L.check(st, "fpemu_%s_multi_amax(%s)" % ("encode" if self.packed else "cast", mode))

    def _pipeline_parts(self, depth):
        """The tensor list cut into `depth` consecutive runs of about equal size, each with its own table, so that the
        exchange step of one run (an all-reduce of its maxima) overlaps the reduction / cast of its neighbours."""
        depth = max(1, min(int(depth), self.count))
        parts = self._parts.get(depth)
        if parts is None:
            if depth == 1:
                parts = {(0, self.count): self}
            else:
                parts = {(t0, t1): MultiTensorPlan(self.ins[t0:t1], self.outs[t0:t1], self.offsets[t0:t1])
                         for t0, t1 in split_runs([x.numel() for x in self.ins], depth)}
            self._parts[depth] = parts
        return parts

    def _amax_scaled(self, mode, flt_max, seed, reduce_fn, pipeline, flags_extra):
        fmt, rnd, _ = L.parse_mode(mode)
        if flt_max is None:
            flt_max = FLT_MAX[fmt]
        if rnd == L.RND_STOCHASTIC and seed is None:
            seed = next_seed()
        seed = int(seed or 0)
        if self._amax_vec is None:
            self._amax_vec = torch.empty(self.count, dtype=torch.float32, device=self.device)
        vec = self._amax_vec
        scales = torch.empty(self.count, dtype=torch.float32, device=self.device)
        vec.zero_()
        parts = self._pipeline_parts(pipeline if reduce_fn is not None else 1)
        pipelined_sweep(sorted(parts),
                        lambda t0, t1: parts[(t0, t1)].amax(out=vec[t0:t1]),
                        None if reduce_fn is None else (lambda t0, t1: reduce_fn(vec[t0:t1])),
                        lambda t0, t1: parts[(t0, t1)]._from_amax(mode, vec[t0:t1], flt_max, scales[t0:t1], seed + t0,
                                                                  flags_extra))
        return self.outs, scales

    def amax_scaled_encode(self, mode, flt_max=None, seed=None, reduce_fn=None, pipeline=1):
        """Per-tensor amax-scaled 8-bit codes for the whole list: amax launch -> [reduce_fn] -> encode launch (which
        derives the scales itself).  Returns (codes, scales); `decode(codes[t], fmt, scale=scales[t])` gives the
        cast's values.  `reduce_fn` / `pipeline`: see amax_scaled_cast."""
        if not self.packed:
            raise ValueError("fpemu: this plan's outputs are not uint8 code buffers")
        return self._amax_scaled(mode, flt_max, seed, reduce_fn, pipeline, L.FLAG_EXPECT_NORMALS)   # scaled to the maximum

    def amax_scaled_cast(self, mode, flt_max=None, seed=None, reduce_fn=None, unscale_fp32=False, pipeline=1):
        """Per-tensor amax-scaled cast of the whole list in TWO launches: amax -> [reduce_fn(amax_vec)] -> cast (the
        cast kernel derives every tensor's scale from its maximum).  Returns (outs, scales).
        `reduce_fn(vec)` reduces the float32 vector of maxima in place across ranks (an all-reduce MAX); it may return
        an object with `.wait()` (an asynchronous collective).  `pipeline` = k > 1 cuts the list into k consecutive
        runs and issues run i's all-reduce before run i+1's amax launch, so the collective runs behind the
        reductions / casts of the neighbouring runs instead of between two kernels; results are identical."""
        if self.packed:
            raise ValueError("fpemu: this plan's outputs are uint8 code buffers: use amax_scaled_encode()")
        return self._amax_scaled(mode, flt_max, seed, reduce_fn, pipeline, L.FLAG_UNSCALE_FP32 if unscale_fp32 else 0)


class HostCaster(object):
    """Host-buffer casts (pinned in, pinned out) pipelined through device staging buffers.
    Owns the staging buffers and three streams; `run` / `run_multi` are what the end-to-end benchmark times."""

    def __init__(self, dtype=torch.float32, chunk_elems=1 << 24, device=None):
        self.device = torch.device("cuda", torch.cuda.current_device() if device is None else device)
        self.dtype = dtype
        self.chunk = int(chunk_elems) // 16 * 16
        self.stage = [torch.empty(self.chunk, dtype=dtype, device=self.device) for _ in range(2)]
        self.streams = [torch.cuda.Stream(self.device) for _ in range(3)]
        self.resident = None               # device copy of the input for run_multi, grown on demand
        self.work = torch.zeros(2, dtype=torch.float32, device=self.device)

    def run(self, h_in, h_out, mode, scale=1.0, seed=0):
        if h_in.is_cuda or h_out.is_cuda:
            raise ValueError("HostCaster.run takes host tensors")
        fmt, rnd, flags = L.parse_mode(mode)
        with _on_device(self.stage[0]):
            st = L.lib().fpemu_cast_host(h_in.data_ptr(), h_out.data_ptr(), h_in.numel(), _DT[self.dtype], fmt, rnd,
                                         flags, float(scale), int(seed), self.stage[0].data_ptr(),
                                         self.stage[1].data_ptr(), self.chunk, self.streams[0].cuda_stream,
                                         self.streams[1].cuda_stream, self.streams[2].cuda_stream)
        L.check(st, "fpemu_cast_host(%s)" % mode)
        return h_out

    def run_multi(self, h_in, jobs):
        """Several casts of one host tensor with ONE upload: `jobs` is a list of (h_out, mode, scale, seed) where
        `scale` is a float or the string "amax" (per-tensor amax-scaled, mpemu/qutils.py:274-287).  Returns the list
        of output tensors.  Synchronises its three streams before returning."""
        import ctypes

        class Job(ctypes.Structure):
            _fields_ = [("h_out", ctypes.c_void_p), ("format", ctypes.c_int), ("rounding", ctypes.c_int),
                        ("flags", ctypes.c_int), ("scale", ctypes.c_float), ("amax_scaled", ctypes.c_int),
                        ("flt_max", ctypes.c_float), ("seed", ctypes.c_uint64)]
        if h_in.is_cuda or not h_in.is_contiguous() or h_in.dtype != self.dtype:
            raise ValueError("HostCaster.run_multi takes a contiguous host tensor of the caster's dtype")
        n = h_in.numel()
        arr = (Job * len(jobs))()
        for i, (h_out, mode, scale, seed) in enumerate(jobs):
            if h_out.is_cuda or not h_out.is_contiguous() or h_out.dtype != self.dtype or h_out.numel() < n:
                raise ValueError("HostCaster.run_multi: outputs must be contiguous host tensors of the input's dtype and size")
            fmt, rnd, flags = L.parse_mode(mode)
            amax = isinstance(scale, str)
            if amax and scale != "amax":
                raise ValueError("scale must be a number or 'amax'")
            arr[i].h_out, arr[i].format, arr[i].rounding, arr[i].flags = h_out.data_ptr(), fmt, rnd, flags
            arr[i].scale, arr[i].amax_scaled = (1.0 if amax else float(scale)), int(amax)
            arr[i].flt_max, arr[i].seed = FLT_MAX.get(fmt, 0.0), int(seed)
        if self.resident is None or self.resident.numel() < n:
            self.resident = torch.empty(n, dtype=self.dtype, device=self.device)
        with _on_device(self.stage[0]):
            st = L.lib().fpemu_cast_host_multi(h_in.data_ptr(), n, _DT[self.dtype], ctypes.addressof(arr), len(jobs),
                                               self.resident.data_ptr(), self.stage[0].data_ptr(),
                                               self.stage[1].data_ptr(), self.work.data_ptr(), self.chunk,
                                               self.streams[0].cuda_stream, self.streams[1].cuda_stream,
                                               self.streams[2].cuda_stream)
        L.check(st, "fpemu_cast_host_multi")
        return [j[0] for j in jobs]
