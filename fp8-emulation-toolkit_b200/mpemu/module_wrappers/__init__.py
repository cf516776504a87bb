"""Thin nn.Module wrappers that make functional ops hookable (reference: mpemu/module_wrappers/*.py).
The sparse Linear/Conv wrappers of the reference belong to its pruning feature and are out of scope."""
from .ops import Matmul, BatchMatmul, AddMatmul, EltwiseAdd, EltwiseMul, EltwiseDiv, Norm, Mean

__all__ = ["Matmul", "BatchMatmul", "AddMatmul", "EltwiseAdd", "EltwiseMul", "EltwiseDiv", "Norm", "Mean"]
