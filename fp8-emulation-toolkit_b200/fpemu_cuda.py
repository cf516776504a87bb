"""`fpemu_cuda` — drop-in for the reference's compiled extension module of the same name.

Reference interface (mpemu/pytquant/cuda/fpemu_impl.cpp:27-54):
    forward(input, mode, size, inplace, scale, block_norm, block_size)  -> [Tensor]
    backward(grad,  mode, size, inplace, scale, block_norm, block_size) -> [Tensor]   (same function)
Same positional order, same return (a 1-element list holding a new tensor, or `input` itself when
`inplace`).  Differences, all deliberate (DESIGN.md "Boundary"):
  * kernels run on torch's CURRENT stream and the call does not synchronise the device
    (reference: legacy default stream + cudaDeviceSynchronize, fpemu_kernels.cu:1060-1071);
  * `scale` may be a Python float OR a 1-element float32 CUDA tensor (used without a host sync);
  * `size` is 64-bit; bfloat16 tensors are accepted; the output is not pre-zeroed;
  * unknown mode -> ValueError (reference: silently returns zeros / the untouched input, :1375-1380);
    scale > 65504 -> ValueError (reference: prints and exit(1), :1037-1041).
"""
import torch

from fpemu_b200 import _lib as L
from fpemu_b200 import ops

manual_seed = ops.manual_seed


def _scale_args(scale):
    if isinstance(scale, torch.Tensor):
        if scale.is_cuda and scale.dtype == torch.float32 and scale.numel() == 1:
            return 1.0, scale                      # device scalar: no .item(), no sync
        return float(scale), None                  # what pybind11's float conversion does in the reference
    return float(scale), None


def forward(input, mode, size, inplace, scale, block_norm, block_size):
    if not input.is_cuda:
        raise RuntimeError("input must be a CUDA tensor")
    if not input.is_contiguous():
        raise RuntimeError("input must be contiguous")
    size = int(size)
    fmt, rnd, flags = L.parse_mode(mode)
    blocked = bool(block_norm)
    # block_norm is honoured only by the E5M2 / E4M3 / E4M3_IEEE / E3M4 / FP4 kernels (fpemu_kernels.cu:1058-1364)
    if fmt in (L.FMT_BF16, L.FMT_FP16) or (flags & L.FLAG_DAZ):
        blocked = False
    if blocked and int(block_size) != size and size % int(block_size) != 0:
        blocked = False                            # reference: silently disabled, :1048-1051
    x = input
    if input.dtype == torch.float64 and blocked:
        # double tensors have their own kernel (fpemu_cast_f64) except in blocked mode, where narrowing first is exact:
        # the block scale is a power of two, so (float)(x * scale) == (float)x * scale
        x = input.float()
    if inplace and x is input:
        out = x
    elif size < x.numel():
        out = torch.zeros_like(x)          # elements past `size` stay zero like the reference's zeros_like (:1044)
    else:
        out = ops.empty_like_coaligned(x)  # every element is written: no memset pass; a view gets an equally aligned output
    host_scale, d_scale = _scale_args(scale)
    if blocked:
        ops.cast_blocked(x, mode, int(block_size), out=out, n=size)
    else:
        ops.cast(x, mode, out=out, scale=host_scale, d_scale=d_scale, n=size)
    if x is not input:
        if inplace:
            input.copy_(out)
            return [input]
        return [out.to(input.dtype)]
    return [out]


backward = forward
