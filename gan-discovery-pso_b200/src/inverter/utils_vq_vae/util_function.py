"""`vq` with the reference's call signature (src/inverter/utils_vq_vae/util_function.py:4-34): the
index of the nearest codebook row for every embedding vector, on the device through
`psob_vq_nearest` (no autograd: the reference marks the indices non-differentiable, :23).

Only the forward index search is part of the hot-path scope; the
straight-through estimator (`vq_st`, :36-77) is training code and stays out of scope.
"""
import torch

from psob import engine as _engine


class VectorQuantization(torch.autograd.Function):
    @staticmethod
    def forward(ctx, inputs, codebook):
        with torch.no_grad():
            indices = _engine.vq_nearest(inputs.detach().float(), codebook.detach().float())
        ctx.mark_non_differentiable(indices)
        return indices

    @staticmethod
    def backward(ctx, grad_output):
        raise RuntimeError("Trying to call `.grad()` on graph containing `VectorQuantization`. "
                           "The function `VectorQuantization` is not differentiable. Use `vq_st` instead.")


vq = VectorQuantization.apply
__all__ = ["vq"]
