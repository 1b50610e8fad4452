"""The custom ``nn`` modules of the reference as PyTorch modules (semantics oracles for the fused kernels).

* :class:`Dropout`        -- /root/reference/asyncsgd/dropout.lua:5-29 (inverted dropout: bernoulli(1-p) mask,
                             output and gradient divided by 1-p; the fc1 epilogue of the tensor-core path).
* :class:`DivideConstant` -- /root/reference/BiCNN/DivideConstant.lua:5-25 (``c / x``, gradient ``-c g / x^2``).
* :class:`Normalize`      -- /root/reference/BiCNN/Normalize.lua:11-76 (Lp row normalisation; the reference builds
                             a d x d Jacobian per row, here the backward is the closed form).
"""
from __future__ import annotations

import torch


class Dropout(torch.nn.Module):
    def __init__(self, p: float = 0.5):
        super().__init__()
        if p >= 1 or p < 0:
            raise ValueError("<Dropout> illegal percentage, must be 0 <= p < 1")   # dropout.lua:8-10
        self.p = p
        self.noise = None

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        # the reference module has no train/eval switch: it always draws a mask (dropout.lua:14-22)
        if self.p == 0:
            return x.clone()
        self.noise = torch.bernoulli(torch.full_like(x, 1 - self.p))
        return x * self.noise / (1 - self.p)     # autograd reproduces updateGradInput (mask, then / (1-p))


class DivideConstant(torch.nn.Module):
    def __init__(self, constant_scalar: float):
        super().__init__()
        assert isinstance(constant_scalar, (int, float)), "input is not scalar!"
        self.numerator = float(constant_scalar)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return self.numerator / x


class _NormalizeFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, p, eps):
        norm = x.abs().pow(p).sum(-1, keepdim=True).add(eps).pow(1.0 / p)
        y = x / norm
        ctx.save_for_backward(x, y, norm)
        ctx.p = p
        return y

    @staticmethod
    def backward(ctx, g):
        x, y, norm = ctx.saved_tensors
        p = ctx.p
        # d/dx_j (x_i / ||x||_p) = delta_ij / n - x_i |x_j|^(p-1) sign(x_j) / n^(p+1)
        w = x.abs().pow(p - 1) * x.sign() / norm.pow(p - 1)
        return (g - (g * y).sum(-1, keepdim=True) * w) / norm, None, None


class Normalize(torch.nn.Module):
    def __init__(self, p: float = 2.0, eps: float = 1e-10):
        super().__init__()
        assert p > 0
        self.p, self.eps = float(p), eps

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return _NormalizeFn.apply(x, self.p, self.eps)
