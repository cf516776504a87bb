"""fpemu_b200 — host side of the B200-native fpemu cast (thin ctypes layer over libfpemu_b200.so)."""
from . import _lib
from ._lib import FpemuError, parse_mode
from .ops import (FLT_MAX, HostCaster, MultiTensorPlan, amax, cast, cast_amax_scaled, cast_amax_scaled_noalloc, cast_blocked, cast_rows_scaled, cast_rowwise, decode, decode_rowwise, encode, encode_rowwise,
                  manual_seed, mean_scale, next_seed, scale_from_amax, enable_device_seed, disable_device_seed, advance_device_seed)

__all__ = ["FpemuError", "parse_mode", "FLT_MAX", "HostCaster", "MultiTensorPlan", "amax", "cast", "cast_amax_scaled", "cast_amax_scaled_noalloc",
           "cast_blocked",
           "cast_rows_scaled", "cast_rowwise", "decode", "decode_rowwise", "encode", "encode_rowwise", "manual_seed", "mean_scale", "next_seed", "scale_from_amax", "enable_device_seed",
           "disable_device_seed", "advance_device_seed", "_lib"]
