"""Hookable stand-ins for torch.matmul / bmm / addmm / elementwise ops / norm / mean
(reference: mpemu/module_wrappers/matmul.py:13-32, eltwise.py, aggregate.py)."""
import torch
import torch.nn as nn


class Matmul(nn.Module):
    def forward(self, a, b):
        return torch.matmul(a, b)


class BatchMatmul(nn.Module):
    def forward(self, a, b):
        return torch.bmm(a, b)


class AddMatmul(nn.Module):
    def forward(self, input, mat1, mat2, beta=1, alpha=1):
        return torch.addmm(input, mat1, mat2, beta=beta, alpha=alpha)


class _Fold(nn.Module):
    def __init__(self, inplace=False):
        super().__init__()
        self.inplace = inplace

    def extra_repr(self):
        return "inplace={}".format(self.inplace)


class EltwiseAdd(_Fold):
    def forward(self, *inputs):
        acc = inputs[0]
        for t in inputs[1:]:
            acc = acc.add_(t) if self.inplace else acc + t
        return acc


class EltwiseMul(_Fold):
    def forward(self, *inputs):
        acc = inputs[0]
        for t in inputs[1:]:
            acc = acc.mul_(t) if self.inplace else acc * t
        return acc


class EltwiseDiv(_Fold):
    def forward(self, x, y):
        return x.div_(y) if self.inplace else x.div(y)


class Norm(nn.Module):
    def forward(self, x, p="fro", dim=None, keepdim=False):
        return torch.norm(x, p=p, dim=dim, keepdim=keepdim)


class Mean(nn.Module):
    def forward(self, x, *args, **kwargs):
        return torch.mean(x, *args, **kwargs)
