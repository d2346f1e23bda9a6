"""Drop-in for nanovllm/layers/rotary_embedding.py.  The cache is built exactly as the
reference builds it (rotary_embedding.py:103-112, fp32 [max_pos, 1, D]); the rotation itself
runs fused with the per-head RMSNorm in `qk_norm_rope` when the caller provides the norm
weights, or with unit weights through `forward` for a stand-alone call."""
from functools import lru_cache

import torch
from torch import nn

from .. import ops


class RotaryEmbedding(nn.Module):

    def __init__(self, head_size: int, rotary_dim: int, max_position_embeddings: int, base: float) -> None:
        super().__init__()
        self.head_size = head_size
        assert rotary_dim == head_size
        inv_freq = 1.0 / (base ** (torch.arange(0, rotary_dim, 2, dtype=torch.float) / rotary_dim))
        t = torch.arange(max_position_embeddings, dtype=torch.float)
        freqs = torch.einsum("i,j -> ij", t, inv_freq)
        cache = torch.cat((freqs.cos(), freqs.sin()), dim=-1).unsqueeze_(1)
        self.register_buffer("cos_sin_cache", cache, persistent=False)

    def forward(self, positions: torch.Tensor, query: torch.Tensor, key: torch.Tensor):
        """Rotation only, same signature as rotary_embedding.py:114-153."""
        return ops.qk_norm_rope(query, key, positions, None, None, self.cos_sin_cache)

    def norm_rope(self, positions, q, k, q_norm, k_norm):
        """q_norm(q), k_norm(k), rotate -- one launch (qwen3.py:191-195)."""
        return ops.qk_norm_rope(q, k, positions, q_norm.weight, k_norm.weight, self.cos_sin_cache, q_norm.eps)


@lru_cache(1)
def _get_rope_cached(head_size: int, rotary_dim: int, max_position: int, base: float):
    return RotaryEmbedding(head_size, rotary_dim, max_position, base)


def get_rope(head_size: int, rotary_dim: int, max_position: int, base: float, rope_scaling: dict | None = None):
    if rope_scaling is not None:
        rope_type = rope_scaling.get("rope_type", "default")
        assert rope_type == "default", f"rope_type '{rope_type}' is not supported yet"
    return _get_rope_cached(head_size, rotary_dim, max_position, base)
