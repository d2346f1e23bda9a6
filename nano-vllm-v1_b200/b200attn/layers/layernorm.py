"""Drop-in for nanovllm/layers/layernorm.py (RMSNorm; fused residual add)."""
import torch
from torch import nn

from .. import ops


class RMSNorm(nn.Module):

    def __init__(self, hidden_size: int, eps: float = 1e-6) -> None:
        super().__init__()
        self.eps = eps
        self.weight = nn.Parameter(torch.ones(hidden_size))

    def forward(self, x: torch.Tensor, residual: torch.Tensor | None = None):
        if x.shape[-1] == ops.HEAD_DIM and residual is None and not x.is_contiguous():
            x = x.contiguous()            # per-head norm on a qkv view when used stand-alone
        if residual is None:
            return ops.rms_norm(x if x.is_contiguous() else x.contiguous(), self.weight, self.eps)
        return ops.rms_norm(x if x.is_contiguous() else x.contiguous(), self.weight, self.eps,
                            residual if residual.is_contiguous() else residual.contiguous())
