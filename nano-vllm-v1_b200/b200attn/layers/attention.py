"""Drop-in for nanovllm/layers/attention.py: same class, constructor, attributes and
`forward(q, k, v)`; the Triton scatter and both flash-attn calls are replaced by the sm_100a
kernels behind include/b200attn.h.

Dispatch (attention.py:214-249):
    cache allocated            -> KV scatter by slot_mapping                 (piece 1)
    cu_seqlens_q is None       -> paged decode, graph form                   (piece 3)
    else, max_seqlen_q == 1
          and block_tables     -> paged decode, eager form (same kernel)     (piece 3)
    else                       -> step plan (built once per step on the device, cached on the
                                  Context): rows with q_len > 1 -> persistent tcgen05 prefill
                                  (piece 2, paged over block_tables or unpaged on fresh k, v);
                                  rows with q_len == 1 of a paged call -> decode kernel (piece 3)
The scatter is a separate launch, stream-ordered before any read, because sequences scheduled
in the same step may read prefix blocks another sequence writes in that step.
"""
import torch
from torch import nn

from .. import ops

import os

_get_context = None
_PREFILL_VARIANT = {"mma": 1, "tc": 2, "tc1": 3}.get(os.environ.get("B200ATTN_PREFILL", ""))


def _context():
    """The Context the engine set: the patched-in nanovllm one when installed as a drop-in,
    else this package's own mirror."""
    global _get_context
    if _get_context is None:
        from ..utils.context import get_context
        _get_context = get_context
    return _get_context()


def use_context_getter(fn) -> None:
    global _get_context
    _get_context = fn


store_kvcache = ops.store_kvcache


class Attention(nn.Module):

    def __init__(self, num_heads, head_dim, scale, num_kv_heads):
        super().__init__()
        self.num_heads = num_heads
        self.head_dim = head_dim
        self.scale = scale
        self.num_kv_heads = num_kv_heads
        self.k_cache = self.v_cache = torch.tensor([])

    def forward(self, q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, kv_stored: bool = False):
        """`kv_stored=True`: the caller already wrote this step's K/V into the caches (the fused
        norm+rope+store launch, ops.qk_norm_rope_store) -- skip the scatter; `k` may then be None
        when the call reads K/V from the cache (paged or decode)."""
        context = _context()
        k_cache, v_cache = self.k_cache, self.v_cache
        if not kv_stored and k_cache.numel() and v_cache.numel():
            store_kvcache(k, v, k_cache, v_cache, context.slot_mapping)
        if context.cu_seqlens_q is None:
            # CUDA-graph decode form: one query token per row, lengths in context_lens
            return ops.attn_decode(q, k_cache, v_cache, context.context_lens, context.block_tables, self.scale)
        if context.block_tables is not None:
            if context.max_seqlen_q == 1 and context.context_lens is not None \
                    and context.context_lens.numel() == q.shape[0]:
                # eager decode (bs > 512 or enforce_eager): every row has exactly one query token
                return ops.attn_decode(q, k_cache, v_cache, context.context_lens, context.block_tables, self.scale)
            k, v = k_cache, v_cache
        if _PREFILL_VARIANT is not None:                      # B200ATTN_PREFILL=mma|tc: unplanned kernels (debug/bench)
            return ops.attn_varlen(q, k, v, context.cu_seqlens_q, context.cu_seqlens_k,
                                   context.max_seqlen_q, context.max_seqlen_k, self.scale,
                                   block_table=context.block_tables, variant=_PREFILL_VARIANT)
        plan = getattr(context, "_b200_plan", None)
        if plan is None or plan.total_q != q.shape[0]:
            plan = ops.attn_plan(context.cu_seqlens_q, context.cu_seqlens_k, q.shape[0], self.num_heads,
                                 self.num_kv_heads, route_decode=context.block_tables is not None, reuse=True)
            try:
                context._b200_plan = plan                     # every layer of this step reuses it
            except AttributeError:
                pass
        return ops.attn_varlen_planned(q, k, v, context.cu_seqlens_q, context.cu_seqlens_k,
                                       context.max_seqlen_q, context.max_seqlen_k, self.scale,
                                       block_table=context.block_tables, plan=plan)
