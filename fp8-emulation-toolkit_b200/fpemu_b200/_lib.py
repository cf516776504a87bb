"""ctypes binding of libfpemu_b200.so (C ABI declared in include/fpemu_b200.h).

There is no CPU or pure-PyTorch fallback: if the CUDA library is missing this module raises at first use.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(_HERE), "libfpemu_b200.so")

# numeric values of include/fpemu_b200.h
FMT_E5M2, FMT_E4M3, FMT_E3M4, FMT_BF16, FMT_FP16, FMT_E4M3_IEEE, FMT_FP4 = range(7)
RND_RTZ, RND_RNE, RND_STOCHASTIC, RND_RNAZ, RND_RNTZ, RND_RPINF, RND_RNINF = range(7)
FLAG_DAZ, FLAG_UNSCALE_FP32, FLAG_NO_SCALE_CLAMP, FLAG_EXPECT_NORMALS = 1, 2, 4, 8
DT_F32, DT_F16, DT_BF16 = range(3)
ERR_MODE, ERR_DTYPE, ERR_SCALE, ERR_ARG, ERR_ALIGN, ERR_TOO_LARGE = -1, -2, -3, -4, -5, -6

_vp, _i64, _i32, _f32, _u64, _cp = (ctypes.c_void_p, ctypes.c_int64, ctypes.c_int, ctypes.c_float,
                                    ctypes.c_uint64, ctypes.c_char_p)

SIGNATURES = {
    "fpemu_version": (_i32, []),
    "fpemu_strerror": (_cp, [_i32]),
    "fpemu_parse_mode": (_i32, [_cp, ctypes.POINTER(_i32), ctypes.POINTER(_i32), ctypes.POINTER(_i32)]),
    "fpemu_cast": (_i32, [_vp, _vp, _i64, _i32, _i32, _i32, _i32, _f32, _vp, _u64, _u64, _vp]),
    "fpemu_cast_f64": (_i32, [_vp, _vp, _i64, _i32, _i32, _i32, _f32, _vp, _u64, _u64, _vp]),
    "fpemu_cast_blocked": (_i32, [_vp, _vp, _i64, _i64, _i32, _i32, _i32, _i32, _u64, _u64, _vp]),
    "fpemu_amax": (_i32, [_vp, _i64, _i32, _vp, _vp]),
    "fpemu_scale_from_amax": (_i32, [_vp, _i32, _f32, _vp, _vp]),
    "fpemu_scale_from_amax_dtype": (_i32, [_vp, _i32, _f32, _i32, _vp, _vp]),
    "fpemu_cast_amax_scaled": (_i32, [_vp, _vp, _i64, _i32, _i32, _i32, _i32, _f32, _vp, _u64, _u64, _vp]),
    "fpemu_cast_rowwise": (_i32, [_vp, _vp, _i64, _i64, _i32, _i32, _i32, _i32, _f32, _vp, _u64, _u64, _vp]),
    "fpemu_multi_table_bytes": (_i64, [_i32]),
    "fpemu_multi_table_build": (_i32, [_vp, _i32, _i32, _vp]),
    "fpemu_amax_multi": (_i32, [_vp, _vp, _i32, _vp, _vp]),
    "fpemu_cast_multi": (_i32, [_vp, _vp, _i32, _i32, _i32, _i32, _vp, _u64, _vp]),
    "fpemu_cast_amax_scaled_noalloc": (_i32, [_vp, _vp, _i64, _i32, _i32, _i32, _i32, _f32, _u64, _u64, _vp]),
    "fpemu_cast_host": (_i32, [_vp, _vp, _i64, _i32, _i32, _i32, _i32, _f32, _u64, _vp, _vp, _i64, _vp, _vp, _vp]),
    "fpemu_cast_host_multi": (_i32, [_vp, _i64, _i32, _vp, _i32, _vp, _vp, _vp, _vp, _i64, _vp, _vp, _vp]),
    "fpemu_mean_scale_work_bytes": (_i32, []),
    "fpemu_scale_from_abs_mean": (_i32, [_vp, _i64, _i32, _f32, _f32, _vp, _vp, _vp]),
    "fpemu_block_scale_from_amax": (_i32, [_vp, _i32, _i32, _vp, _vp]),
    "fpemu_fused_work_floats": (_i32, []),
    "fpemu_cast_amax_scaled_fused": (_i32, [_vp, _vp, _i64, _i32, _i32, _i32, _i32, _f32, _vp, _i32, _u64, _u64, _vp]),
    "fpemu_cast_rows_scaled": (_i32, [_vp, _vp, _i64, _i64, _i32, _i32, _i32, _i32, _vp, _u64, _u64, _vp]),
    "fpemu_encode_rowwise": (_i32, [_vp, _vp, _i64, _i64, _i32, _i32, _i32, _i32, _f32, _vp, _u64, _u64, _vp]),
    "fpemu_decode_rowwise": (_i32, [_vp, _vp, _i64, _i64, _i32, _i32, _i32, _vp, _vp]),
    "fpemu_set_seed_source": (_i32, [_vp]),
    "fpemu_advance_seed_source": (_i32, [_vp]),
    "fpemu_encode_multi": (_i32, [_vp, _vp, _i32, _i32, _i32, _i32, _vp, _u64, _vp]),
    "fpemu_cast_multi_amax": (_i32, [_vp, _vp, _i32, _i32, _i32, _i32, _vp, _f32, _vp, _u64, _vp]),
    "fpemu_encode_multi_amax": (_i32, [_vp, _vp, _i32, _i32, _i32, _i32, _vp, _f32, _vp, _u64, _vp]),
    "fpemu_exchange_floats": (_i64, [_i32, _i32]),
    "fpemu_amax_multi_push": (_i32, [_vp, _vp, _i32, _vp, _vp, _i32, _i32, ctypes.c_uint32, _vp]),
    "fpemu_cast_multi_xchg": (_i32, [_vp, _vp, _i32, _i32, _i32, _i32, _vp, _i32, ctypes.c_uint32, _f32, _vp, _u64, _vp]),
    "fpemu_encode_multi_xchg": (_i32, [_vp, _vp, _i32, _i32, _i32, _i32, _vp, _i32, ctypes.c_uint32, _f32, _vp, _u64, _vp]),
    "fpemu_encode": (_i32, [_vp, _vp, _i64, _i32, _i32, _i32, _i32, _f32, _vp, _u64, _u64, _vp]),
    "fpemu_decode": (_i32, [_vp, _vp, _i64, _i32, _i32, _i32, _f32, _vp, _vp]),
}

_lib = None


class FpemuError(RuntimeError):
    pass


def lib():
    """The loaded library.  Raises (loudly) when it has not been built: there is no fallback path."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise FpemuError(
                "libfpemu_b200.so not found at %s — build it with `python fp8-emulation-toolkit_b200/build.py` "
                "(there is no CPU / PyTorch fallback for the fpemu cast)" % LIB_PATH)
        L = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def strerror(status):
    return lib().fpemu_strerror(int(status)).decode()


def check(status, what="fpemu"):
    """0 -> ok; <0 -> ValueError (bad argument); >0 -> RuntimeError (CUDA error).  Never exit()."""
    if status == 0:
        return
    msg = "%s: %s (status %d)" % (what, strerror(status), status)
    if status < 0:
        raise ValueError(msg)
    raise FpemuError(msg)


_mode_cache = {}


def parse_mode(mode):
    """'E4M3_RNE' -> (format, rounding, flags).  ValueError for unknown strings (the reference returns zeros)."""
    r = _mode_cache.get(mode)
    if r is None:
        f, rn, fl = _i32(), _i32(), _i32()
        st = lib().fpemu_parse_mode(str(mode).encode(), ctypes.byref(f), ctypes.byref(rn), ctypes.byref(fl))
        if st != 0:
            raise ValueError("fpemu: unknown mode %r" % (mode,))
        r = (f.value, rn.value, fl.value)
        _mode_cache[mode] = r
    return r
