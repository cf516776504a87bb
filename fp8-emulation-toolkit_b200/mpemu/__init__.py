"""mpemu — drop-in façade with the entry points of the reference package of the same name
(mpemu/__init__.py:1-3 of the reference), restricted to the fpemu cast path: pytquant (FPEmuOp), qutils
(quantize_tensor + hooks), the per-format emulators and mpt_emu (initialize / quantize_model).
The reference's CPU hardware model (`cmodel`), sparsity and int8 tooling are out of scope."""
from . import pytquant
from . import module_wrappers
from . import qutils
from . import mpt_emu

__all__ = ["pytquant", "module_wrappers", "qutils", "mpt_emu"]
