"""Oracle: paged / unpaged varlen attention and paged decode (test infrastructure only).

CPU fp32 restatement of what the reference's `Attention.forward`
(/root/reference/nanovllm/layers/attention.py:168-249) asks flash-attn to compute:

  * dispatch ................ attention.py:214-249 (scatter first :219-220; decode-graph
                              form iff `cu_seqlens_q is None` :222; paged iff
                              `block_tables is not None` :237-240)
  * which K/V rows are read . slot = block_table[p // 256] * 256 + p % 256
                              (model_runner.py:453-481), cache layout
                              [num_blocks, 256, Hkv, D] (model_runner.py:368-374)
  * padding conventions ..... eager block tables padded with -1 (model_runner.py:394-396),
                              graph tables padded with 0 and context_lens 0 for padding
                              rows (model_runner.py:589-596)
  * the arithmetic itself lives in flash-attn (dependency `flash-attn`, unpinned in
    pyproject.toml:18; 2.8.3 installed): heads map h -> h // (Hq/Hkv); causal mask is
    aligned to the BOTTOM-RIGHT corner (query i sees keys j <= i + Lk - Lq); a row with no
    visible key yields 0; softmax in fp32 on scale * q.k.

Parity status: PINNED.  flash-attn has no CPU code path, so the reference cannot run in the
build container; instead the unmodified reference `Attention.forward` was run on a B200
(tools/make_golden_gpu.py, reference installed offline under baseline/_ref) on seeded cases
covering unpaged prefill, paged chunked/mixed prefill (G = 2, 4, 5) and graph-form decode with
padding rows; the outputs are committed as tests/golden/attn_ref_*.pt and
tests/test_oracle_cpu.py::test_attention_oracle_matches_reference_forward_on_gpu holds this file
to them (max-abs <= 2e-2, cache after the step bit-exact by SHA-256).
tests/test_gpu_vs_flashattn.py re-checks live against flash-attn's entry points on the box, and
tests/test_oracle_cpu.py also checks first principles (dense softmax(QK^T)V).
"""
from typing import Optional

import torch

from .context import Context
from .kv_scatter import kv_scatter

PAGE = 256  # Sequence.block_size is a class constant (sequence.py:73)


def gather_paged_kv(cache: torch.Tensor, block_table_row: torch.Tensor, length: int) -> torch.Tensor:
    """Rows 0..length-1 of one sequence: cache[block_table[p // P], p % P] -> [length, Hkv, D]."""
    page = cache.shape[1]
    if length == 0:
        return cache.new_zeros((0,) + tuple(cache.shape[2:]))
    nblk = (length + page - 1) // page
    blocks = block_table_row[:nblk].to(torch.int64)
    assert (blocks >= 0).all(), "block table entry inside the context is negative"
    rows = cache[blocks].reshape(nblk * page, *cache.shape[2:])
    return rows[:length]


def _attend_one(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, scale: float, causal: bool = True) -> torch.Tensor:
    """q [Lq,Hq,D], k/v [Lk,Hkv,D] (any float dtype) -> fp32 [Lq,Hq,D], bottom-right causal."""
    Lq, Hq, D = q.shape
    Lk, Hkv, _ = k.shape
    out = torch.zeros(Lq, Hq, D, dtype=torch.float32)
    if Lq == 0 or Lk == 0:
        return out
    G = Hq // Hkv
    qf = q.float().permute(1, 0, 2)                                  # [Hq, Lq, D]
    kf = k.float().permute(1, 0, 2).repeat_interleave(G, dim=0)      # [Hq, Lk, D]
    vf = v.float().permute(1, 0, 2).repeat_interleave(G, dim=0)
    s = torch.matmul(qf, kf.transpose(1, 2)) * scale                  # [Hq, Lq, Lk]
    if causal:
        i = torch.arange(Lq).unsqueeze(1)
        j = torch.arange(Lk).unsqueeze(0)
        visible = j <= i + (Lk - Lq)
        s = s.masked_fill(~visible.unsqueeze(0), float("-inf"))
        dead = ~visible.any(dim=1)                                    # rows with no visible key
    else:
        dead = torch.zeros(Lq, dtype=torch.bool)
    # P stays fp32 here.  flash-attn (and both sm_100a kernels) round P to bf16 before P.V: this restatement is
    # deliberately MORE exact than the reference it restates at this one point (the difference is far inside
    # the 2e-2 bar); rounding P here would tie the oracle to one tile order's running maximum.
    p = torch.softmax(s, dim=-1)
    p = torch.nan_to_num(p, nan=0.0)
    o = torch.matmul(p, vf).permute(1, 0, 2)                          # [Lq, Hq, D]
    o[dead] = 0.0
    return o


def attn_varlen(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor,
                cu_seqlens_q: torch.Tensor, cu_seqlens_k: torch.Tensor, scale: float,
                block_tables: Optional[torch.Tensor] = None) -> torch.Tensor:
    """flash_attn_varlen_func(..., causal=True, block_table=...) as called at attention.py:242-248.

    Unpaged (block_tables None): k, v are packed [Tk, Hkv, D] indexed by cu_seqlens_k.
    Paged: k, v are the caches [num_blocks, 256, Hkv, D]; row p of sequence b lives at
    cache[block_tables[b, p // 256], p % 256].
    Returns q.dtype [T, Hq, D].
    """
    cu_q = cu_seqlens_q.tolist()
    cu_k = cu_seqlens_k.tolist()
    B = len(cu_q) - 1
    out = torch.zeros(q.shape, dtype=torch.float32)
    for b in range(B):
        q0, q1 = cu_q[b], cu_q[b + 1]
        Lk = cu_k[b + 1] - cu_k[b]
        if block_tables is None:
            kb, vb = k[cu_k[b]:cu_k[b + 1]], v[cu_k[b]:cu_k[b + 1]]
        else:
            kb = gather_paged_kv(k, block_tables[b], Lk)
            vb = gather_paged_kv(v, block_tables[b], Lk)
        out[q0:q1] = _attend_one(q[q0:q1], kb, vb, scale)
    return out.to(q.dtype)


def attn_decode(q: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor,
                context_lens: torch.Tensor, block_tables: torch.Tensor, scale: float) -> torch.Tensor:
    """flash_attn_with_kvcache(q[B,1,Hq,D], ..., cache_seqlens, block_table, causal=True)
    as called at attention.py:229-232.  q: [B, Hq, D].  Rows with context_len 0 (graph
    padding, model_runner.py:592) give 0."""
    B = q.shape[0]
    out = torch.zeros(q.shape, dtype=torch.float32)
    lens = context_lens.tolist()
    for b in range(B):
        L = lens[b]
        if L == 0:
            continue
        kb = gather_paged_kv(k_cache, block_tables[b], L)
        vb = gather_paged_kv(v_cache, block_tables[b], L)
        out[b:b + 1] = _attend_one(q[b:b + 1], kb, vb, scale)
    return out.to(q.dtype)


def attention_forward(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor,
                      k_cache: torch.Tensor, v_cache: torch.Tensor, ctx: Context,
                      scale: float) -> torch.Tensor:
    """`Attention.forward` (attention.py:168-249): scatter, then one of the three reads."""
    if k_cache.numel() and v_cache.numel():                      # attention.py:219-220
        kv_scatter(k, v, k_cache, v_cache, ctx.slot_mapping)
    if ctx.cu_seqlens_q is None:                                 # attention.py:222-234
        return attn_decode(q, k_cache, v_cache, ctx.context_lens, ctx.block_tables, scale)
    if ctx.block_tables is not None:                             # attention.py:237-240
        return attn_varlen(q, k_cache, v_cache, ctx.cu_seqlens_q, ctx.cu_seqlens_k, scale, ctx.block_tables)
    return attn_varlen(q, k, v, ctx.cu_seqlens_q, ctx.cu_seqlens_k, scale, None)


def causal_pairs(lq: int, lk: int) -> int:
    """Unmasked (query, key) pairs of one sequence, SURVEY.md 8(d): Lq*(Lk-Lq) + Lq*(Lq+1)/2."""
    return lq * (lk - lq) + lq * (lq + 1) // 2
