from . import fpemu as fpemu_cuda      # mpemu/pytquant/cuda/__init__.py:1-3 of the reference

__all__ = ["fpemu_cuda"]
