"""ctypes front-end of oracle/fpemu_oracle.c (the CPU restatement of the reference fpemu cast).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's CPU-baseline
legs.  The product package never imports it.  See the header of fpemu_oracle.c for the parity pin.
"""
import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "fpemu_oracle.c")
LIB = os.path.join(HERE, "libfpemu_oracle.so")

# numeric values of include/fpemu_b200.h
FMT = {"E5M2": 0, "E4M3": 1, "E3M4": 2, "BF16": 3, "FP16": 4, "E4M3_IEEE": 5, "FP4": 6}
RND = {"RTZ": 0, "RNE": 1, "STOCHASTIC": 2, "RNAZ": 3, "RNTZ": 4, "RPINF": 5, "RNINF": 6}
FLAG_DAZ = 1
FLAG_UNSCALE_FP32 = 2
DT = {"f32": 0, "f16": 1, "bf16": 2}


def build(force=False):
    """Compile the C restatement with gcc (seconds).  -ffp-contract=off: no fused multiply-adds."""
    hdr = os.path.join(HERE, "..", "include", "fpemu_b200.h")
    if (not force and os.path.exists(LIB) and os.path.getmtime(LIB) >= os.path.getmtime(SRC)
            and os.path.getmtime(LIB) >= os.path.getmtime(hdr)):
        return LIB
    cmd = ["gcc", "-O2", "-std=c11", "-ffp-contract=off", "-fno-fast-math", "-shared", "-fPIC", SRC, "-o", LIB, "-lm"]
    subprocess.check_call(cmd)
    return LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = ctypes.CDLL(LIB)
        vp, i64, i32, f32, u64 = ctypes.c_void_p, ctypes.c_int64, ctypes.c_int, ctypes.c_float, ctypes.c_uint64
        L.fpemu_oracle_cast.argtypes = [vp, vp, i64, i32, i32, i32, i32, f32, vp]
        L.fpemu_oracle_cast.restype = i32
        L.fpemu_oracle_cast_blocked.argtypes = [vp, vp, i64, i64, i32, i32, i32, i32, vp]
        L.fpemu_oracle_cast_blocked.restype = i32
        L.fpemu_oracle_amax.argtypes = [vp, i64, i32]
        L.fpemu_oracle_amax.restype = f32
        L.fpemu_oracle_scale_from_amax.argtypes = [f32, f32]
        L.fpemu_oracle_scale_from_amax.restype = f32
        L.fpemu_oracle_scale_from_amax_dtype.argtypes = [f32, f32, i32, i32]
        L.fpemu_oracle_scale_from_amax_dtype.restype = f32
        L.fpemu_oracle_rand_bits.argtypes = [u64, u64, i64, i32, vp]
        L.fpemu_oracle_rand_bits.restype = None
        L.fpemu_oracle_philox4x32_10.argtypes = [u64, u64, vp]
        L.fpemu_oracle_philox4x32_10.restype = None
        L.fpemu_oracle_set_nan_mode.argtypes = [i32]
        L.fpemu_oracle_set_nan_mode.restype = None
        L.fpemu_oracle_float2half.argtypes = [f32]
        L.fpemu_oracle_float2half.restype = ctypes.c_uint16
        L.fpemu_oracle_half2float.argtypes = [ctypes.c_uint16]
        L.fpemu_oracle_half2float.restype = f32
        L.fpemu_oracle_encode.argtypes = [vp, vp, i64, i32, i32, i32, i32, f32, vp]
        L.fpemu_oracle_encode.restype = i32
        L.fpemu_oracle_decode.argtypes = [vp, vp, i64, i32, i32, i32, f32]
        L.fpemu_oracle_decode.restype = i32
        L.fpemu_oracle_code_from_h16.argtypes = [ctypes.c_uint16, i32]
        L.fpemu_oracle_code_from_h16.restype = ctypes.c_uint8
        L.fpemu_oracle_h16_from_code.argtypes = [ctypes.c_uint8, i32]
        L.fpemu_oracle_h16_from_code.restype = ctypes.c_uint16
        _lib = L
    return _lib


def set_nan_mode(x86_f16c):
    """False (default): float->half NaN conversion as CUDA cvt.rn.f16.f32; True: as x86 F16C."""
    lib().fpemu_oracle_set_nan_mode(1 if x86_f16c else 0)


def parse_mode(mode):
    """'E5M2_DAZ_RNE' -> (format, rounding, flags); mirrors the grammar of fpemu_kernels.cu:1058-1364."""
    m = mode.upper()
    flags = 0
    for prefix, fmt in (("E5M2_DAZ_", "E5M2"), ("E4M3_IEEE_", "E4M3_IEEE"), ("FLOAT16_DAZ_", "FP16"),
                        ("E5M2_", "E5M2"), ("E4M3_", "E4M3"), ("E3M4_", "E3M4"),
                        ("BFLOAT16_", "BF16"), ("FLOAT16_", "FP16")):
        if m.startswith(prefix):
            if "DAZ" in prefix:
                flags |= FLAG_DAZ
            return FMT[fmt], RND[m[len(prefix):]], flags
    if m == "FP4_NEAREST":
        return FMT["FP4"], RND["RNE"], 0
    raise ValueError("unknown fpemu mode %r" % mode)


def _np_dtype_code(a):
    if a.dtype == np.float32:
        return DT["f32"]
    if a.dtype == np.float16:
        return DT["f16"]
    if a.dtype == np.uint16:  # bf16 bit patterns (numpy has no bfloat16)
        return DT["bf16"]
    raise TypeError("oracle supports float32 / float16 / uint16(bf16 bits), got %s" % a.dtype)


def rand_bits(seed, elem_offset, n, fmt):
    out = np.empty(n, dtype=np.uint32)
    lib().fpemu_oracle_rand_bits(seed, elem_offset, n, fmt, out.ctypes.data)
    return out


def cast(x, mode, scale=1.0, unscale_fp32=False, seed=0, elem_offset=0, rand=None):
    """QDQ `x` (numpy float32 / float16 / uint16-as-bf16-bits) like fpemu_*.forward(x, mode, n, False, scale)."""
    x = np.ascontiguousarray(x)
    fmt, rnd, flags = parse_mode(mode)
    if unscale_fp32:
        flags |= FLAG_UNSCALE_FP32
    out = np.empty_like(x)
    rptr = None
    if rnd == RND["STOCHASTIC"]:
        if rand is None:
            rand = rand_bits(seed, elem_offset, x.size, fmt)
        rand = np.ascontiguousarray(rand, dtype=np.uint32)
        rptr = rand.ctypes.data
    st = lib().fpemu_oracle_cast(x.ctypes.data, out.ctypes.data, x.size, _np_dtype_code(x), fmt, rnd, flags,
                                 float(scale), rptr)
    if st != 0:
        raise ValueError("fpemu_oracle_cast status %d" % st)
    return out


def cast_f64(x, mode, scale=1.0, unscale_fp32=False, seed=0, elem_offset=0):
    """QDQ of a float64 array as the reference's kernels do it when dispatched for double
    (AT_DISPATCH_FLOATING_TYPES_AND_HALF, cuda/fpemu_kernels.cu:1059 ...): `scalar_t scale = in_scale` is a double
    there (:405), so `float inval = in[gid] * scale` (:420) multiplies in double and narrows once; formats without a
    scale narrow the element (`__anyfloat2float_rn`, :93-105).  The float is rounded like a float tensor's element; the
    result, un-scaled by `float scale_reciprocal = 1.0 / scale` and (CUDA op) passed through half (`__half2anyfloat`,
    :136-148, :461), is widened back."""
    x = np.ascontiguousarray(x, dtype=np.float64)
    fmt, _, _ = parse_mode(mode)
    takes_scale = fmt not in (FMT["BF16"], FMT["FP16"])
    s = np.float32(scale if takes_scale else 1.0)
    with np.errstate(over="ignore", invalid="ignore"):
        x32 = (x * np.float64(s)).astype(np.float32) if s != 1.0 else x.astype(np.float32)
        q = cast(x32, mode, scale=1.0, seed=seed, elem_offset=elem_offset)
        if s != 1.0:
            q = q * np.float32(1.0 / np.float64(s))
            if not unscale_fp32 and fmt != FMT["FP4"]:
                q = q.astype(np.float16).astype(np.float32)
    return q.astype(np.float64)


def encode(x, mode, scale=1.0, seed=0, elem_offset=0, rand=None):
    """The rounding of `cast`, returned as the format's one-byte codes (uint8 array)."""
    x = np.ascontiguousarray(x)
    fmt, rnd, flags = parse_mode(mode)
    out = np.empty(x.shape, dtype=np.uint8)
    rptr = None
    if rnd == RND["STOCHASTIC"]:
        if rand is None:
            rand = rand_bits(seed, elem_offset, x.size, fmt)
        rand = np.ascontiguousarray(rand, dtype=np.uint32)
        rptr = rand.ctypes.data
    st = lib().fpemu_oracle_encode(x.ctypes.data, out.ctypes.data, x.size, _np_dtype_code(x), fmt, rnd, flags,
                                   float(scale), rptr)
    if st != 0:
        raise ValueError("fpemu_oracle_encode status %d" % st)
    return out


def decode(codes, fmt_name, dtype=np.float32, scale=1.0, unscale_fp32=False):
    """Codes -> values (`dtype`: np.float32, np.float16, or np.uint16 for bf16 bit patterns), un-scaled like `cast`."""
    codes = np.ascontiguousarray(codes, dtype=np.uint8)
    out = np.empty(codes.shape, dtype=dtype)
    st = lib().fpemu_oracle_decode(codes.ctypes.data, out.ctypes.data, codes.size, _np_dtype_code(out), FMT[fmt_name],
                                   FLAG_UNSCALE_FP32 if unscale_fp32 else 0, float(scale))
    if st != 0:
        raise ValueError("fpemu_oracle_decode status %d" % st)
    return out


def cast_blocked(x, mode, block_size, unscale_fp32=False, seed=0, elem_offset=0, rand=None):
    x = np.ascontiguousarray(x)
    fmt, rnd, flags = parse_mode(mode)
    if unscale_fp32:
        flags |= FLAG_UNSCALE_FP32
    out = np.empty_like(x)
    rptr = None
    if rnd == RND["STOCHASTIC"]:
        if rand is None:
            rand = rand_bits(seed, elem_offset, x.size, fmt)
        rand = np.ascontiguousarray(rand, dtype=np.uint32)
        rptr = rand.ctypes.data
    st = lib().fpemu_oracle_cast_blocked(x.ctypes.data, out.ctypes.data, x.size, int(block_size),
                                         _np_dtype_code(x), fmt, rnd, flags, rptr)
    if st != 0:
        raise ValueError("fpemu_oracle_cast_blocked status %d" % st)
    return out


def amax(x):
    x = np.ascontiguousarray(x)
    return float(lib().fpemu_oracle_amax(x.ctypes.data, x.size, _np_dtype_code(x)))


def scale_from_amax(a, flt_max, like=None, clamp=True):
    """`flt_max / vmax` + clamp of mpemu/qutils.py:283-285 (torch evaluates it as reciprocal * flt_max in the tensor's
    dtype).  `like`: the tensor (numpy array) whose dtype the arithmetic follows; default float32."""
    dt = DT["f32"] if like is None else _np_dtype_code(like)
    return float(lib().fpemu_oracle_scale_from_amax_dtype(float(a), float(flt_max), dt, 1 if clamp else 0))


def cast_rowwise(x2d, mode, flt_max, unscale_fp32=False, clamp=True):
    """Per-channel amax-scaled cast, the loop of mpemu/qutils.py:289-311 restated over the oracle (clamp=False: the
    fine-grained branch :313-337 over its groups)."""
    x2d = np.ascontiguousarray(x2d)
    out = np.empty_like(x2d)
    scales = np.empty(x2d.shape[0], dtype=np.float32)
    for r in range(x2d.shape[0]):
        s = scale_from_amax(amax(x2d[r]), flt_max, like=x2d, clamp=clamp)
        scales[r] = s
        out[r] = cast(x2d[r], mode, scale=s, unscale_fp32=unscale_fp32)
    return out, scales


def encode_rowwise(x2d, mode, flt_max, seed=0, elem_offset=0):
    """Per-channel codes + scales: the same loop, `encode` instead of `cast` (row r starts at element r * cols)."""
    x2d = np.ascontiguousarray(x2d)
    codes = np.empty(x2d.shape, dtype=np.uint8)
    scales = np.empty(x2d.shape[0], dtype=np.float32)
    for r in range(x2d.shape[0]):
        s = scale_from_amax(amax(x2d[r]), flt_max, like=x2d)
        scales[r] = s
        codes[r] = encode(x2d[r], mode, scale=s, seed=seed, elem_offset=elem_offset + r * x2d.shape[1])
    return codes, scales


def decode_rowwise(codes, scales, fmt_name, dtype=np.float32, unscale_fp32=False):
    out = np.empty(codes.shape, dtype=dtype)
    for r in range(codes.shape[0]):
        out[r] = decode(codes[r], fmt_name, dtype=dtype, scale=float(scales[r]), unscale_fp32=unscale_fp32)
    return out


def philox(seed, ctr):
    out = (ctypes.c_uint32 * 4)()
    lib().fpemu_oracle_philox4x32_10(seed, ctr, out)
    return list(out)


def float2half_bits(x):
    return int(lib().fpemu_oracle_float2half(float(x)))


def same_bits_or_both_nan(a, b):
    """Bit-exact comparison that treats any-NaN == any-NaN (payloads differ between x86 F16C and CUDA cvt)."""
    a = np.ascontiguousarray(a)
    b = np.ascontiguousarray(b)
    assert a.dtype == b.dtype and a.shape == b.shape
    if a.dtype == np.uint16:  # bf16 bits
        an = ((a & 0x7FFF) > 0x7F80)
        bn = ((b & 0x7FFF) > 0x7F80)
        return ((a == b) | (an & bn))
    ai = a.view(np.uint32 if a.dtype == np.float32 else np.uint16)
    bi = b.view(ai.dtype)
    return (ai == bi) | (np.isnan(a) & np.isnan(b))
