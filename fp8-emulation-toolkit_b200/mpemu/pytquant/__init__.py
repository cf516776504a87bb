"""mpemu.pytquant — the reference exposes `fpemu_cpp` always and `fpemu_cuda` when CUDA is present
(mpemu/pytquant/__init__.py:4-13).  Here only the CUDA op exists: there is no CPU implementation."""
from . import cuda
from .cuda import fpemu_cuda

__all__ = ["cuda", "fpemu_cuda"]
