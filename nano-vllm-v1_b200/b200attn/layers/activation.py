"""Drop-in for nanovllm/layers/activation.py."""
import torch
from torch import nn

from .. import ops


class SiluAndMul(nn.Module):

    def __init__(self):
        super().__init__()

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return ops.silu_and_mul(x if x.is_contiguous() else x.contiguous())
