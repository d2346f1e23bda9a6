"""Oracle-side mirror of the reference's metadata bag (test infrastructure only).

Follows /root/reference/nanovllm/utils/context.py:22-55 -- same eight fields in the
same positional order (the reference calls `set_context` positionally,
model_runner.py:494), and NO `is_prefill` field.
"""
from dataclasses import dataclass
from typing import Optional

import torch


@dataclass
class Context:
    cu_seqlens_q: Optional[torch.Tensor] = None
    cu_seqlens_k: Optional[torch.Tensor] = None
    max_seqlen_q: int = 0
    max_seqlen_k: int = 0
    slot_mapping: Optional[torch.Tensor] = None
    context_lens: Optional[torch.Tensor] = None
    block_tables: Optional[torch.Tensor] = None
    seq_need_compute_logits: Optional[torch.Tensor] = None
