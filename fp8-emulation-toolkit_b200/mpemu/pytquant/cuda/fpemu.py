"""FPEmuOp — autograd wrapper of the cast with a straight-through backward.

Mirrors the reference's `FPEmuOp` (mpemu/pytquant/cuda/fpemu.py:43-71): same call
`FPEmuOp.apply(input, mode='NONE', inplace=False, scale=1.0, blocknorm=False, blocksize=1)`, `'NONE'` is a
pass-through, sparse tensors have their values cast, `mark_dirty` on in-place, backward returns the incoming
gradient unchanged.  Not reproduced: the global `torch.manual_seed(42)` the reference runs at import (:19)."""
import torch
from torch.autograd import Function

import fpemu_cuda     # the top-level native-module replacement (fp8-emulation-toolkit_b200/fpemu_cuda.py)


class FPEmuOp(Function):
    @staticmethod
    def forward(ctx, input, mode='NONE', inplace=False, scale=1.0, blocknorm=False, blocksize=1):
        if mode == 'NONE':
            ctx.mark_dirty(input)
            return input
        if input.is_sparse:
            input = input.coalesce()
            values = input._values().contiguous()
            outputs = fpemu_cuda.forward(values, mode, values.nelement(), inplace, scale, blocknorm, blocksize)
            if inplace:
                output = input
            else:
                output = torch.sparse_coo_tensor(input.indices(), outputs[0], input.size())
        else:
            src = input.contiguous()
            outputs = fpemu_cuda.forward(src, mode, src.nelement(), inplace, scale, blocknorm, blocksize)
            output = outputs[0]
            if inplace and src is not input:      # a non-contiguous input was copied: write the result back
                input.copy_(output)
                output = input
        if inplace:
            ctx.mark_dirty(input)
        return output

    @staticmethod
    def backward(ctx, output_grad):
        # straight-through estimator; one gradient slot per forward argument
        return output_grad, None, None, None, None, None
