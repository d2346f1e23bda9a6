"""Oracle: per-head RMSNorm, RoPE, fused add+RMSNorm, SiluAndMul (test infrastructure only).

Restates, operation by operation and rounding by rounding:
  * RMSNorm.rms_forward ........ /root/reference/nanovllm/layers/layernorm.py:57-93
  * RMSNorm.add_rms_forward .... layernorm.py:95-141
  * apply_rotary_emb ........... layers/rotary_embedding.py:27-65 (NeoX halves, fp32 math)
  * RotaryEmbedding cache ...... rotary_embedding.py:103-112 (fp32 [max_pos, 1, D] = cos|sin)
  * RotaryEmbedding.forward .... rotary_embedding.py:114-153
  * SiluAndMul.forward ......... layers/activation.py:36-58
  * caller order ............... models/qwen3.py:183-197 (q_norm, k_norm, then rotary)

Pinned (mode="eager") bit-exact against those modules run eagerly on CPU from
/root/reference (tests/golden/pointwise_*.pt, tools/make_golden.py).  On the GPU the
reference wraps them in torch.compile; inductor keeps every explicit `.to(dtype)` but
computes `silu(x) * y` in fp32 with a single rounding, and may contract a*b-c*d into FMA.
mode="compiled" restates the former; the latter is why GPU tolerances are "<= 1 bf16 ulp".
"""
from typing import Optional, Tuple

import torch


def make_cos_sin_cache(head_dim: int, max_position: int, base: float) -> torch.Tensor:
    """rotary_embedding.py:103-112 -> fp32 [max_position, 1, head_dim] = [cos(D/2) | sin(D/2)]."""
    inv_freq = 1.0 / (base ** (torch.arange(0, head_dim, 2, dtype=torch.float) / head_dim))
    t = torch.arange(max_position, dtype=torch.float)
    freqs = torch.einsum("i,j -> ij", t, inv_freq)
    return torch.cat((freqs.cos(), freqs.sin()), dim=-1).unsqueeze_(1)


def rms_norm(x: torch.Tensor, weight: torch.Tensor, eps: float = 1e-6) -> torch.Tensor:
    """layernorm.py:85-93: fp32 normalise, round to x.dtype, THEN multiply by weight in x.dtype."""
    orig = x.dtype
    xf = x.float()
    var = xf.pow(2).mean(dim=-1, keepdim=True)
    xf = xf * torch.rsqrt(var + eps)
    return xf.to(orig) * weight.to(orig)


def add_rms_norm(x: torch.Tensor, residual: torch.Tensor, weight: torch.Tensor,
                 eps: float = 1e-6) -> Tuple[torch.Tensor, torch.Tensor]:
    """layernorm.py:131-141: fp32 x+residual; residual_out = round(sum); norm as rms_norm on the fp32 sum."""
    orig = x.dtype
    xf = x.float() + residual.float()
    res_out = xf.to(orig)
    var = xf.pow(2).mean(dim=-1, keepdim=True)
    xf = xf * torch.rsqrt(var + eps)
    return xf.to(orig) * weight.to(orig), res_out


def rope(x: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor) -> torch.Tensor:
    """rotary_embedding.py:60-65.  x [T,H,D]; cos/sin fp32 [T,1,D/2]."""
    x1, x2 = torch.chunk(x.float(), 2, dim=-1)
    y1 = x1 * cos - x2 * sin
    y2 = x2 * cos + x1 * sin
    return torch.cat((y1, y2), dim=-1).to(x.dtype)


def qk_norm_rope(q: torch.Tensor, k: torch.Tensor, positions: torch.Tensor,
                 q_weight: torch.Tensor, k_weight: torch.Tensor, cos_sin_cache: torch.Tensor,
                 eps: float = 1e-6) -> Tuple[torch.Tensor, torch.Tensor]:
    """qwen3.py:191-195: q = q_norm(q); k = k_norm(k); q, k = rotary_emb(positions, q, k).

    q [T,Hq,D], k [T,Hkv,D] (views of the qkv projection), positions int64 [T].
    The bf16 round-trip between the norm and the rotation is part of the contract."""
    qn = rms_norm(q, q_weight, eps)
    kn = rms_norm(k, k_weight, eps)
    cos_sin = cos_sin_cache[positions]                      # [T,1,D]  (rotary_embedding.py:148)
    cos, sin = cos_sin.chunk(2, dim=-1)
    return rope(qn, cos, sin), rope(kn, cos, sin)


def silu_and_mul(x: torch.Tensor, mode: str = "compiled") -> torch.Tensor:
    """activation.py:56-58.  x [T, 2I] -> [T, I] = silu(x[:, :I]) * x[:, I:].

    mode="eager": silu rounded to x.dtype before the product (two roundings), which is what
    the module does with torch.compile disabled.  mode="compiled": fp32 throughout, one
    rounding -- what inductor emits on the GPU."""
    g, u = x.chunk(2, -1)
    if mode == "eager":
        return torch.nn.functional.silu(g) * u
    assert mode == "compiled", mode
    return (torch.nn.functional.silu(g.float()) * u.float()).to(x.dtype)
