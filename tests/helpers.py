"""Shared seeded-input builders for the CPU and GPU tests (no oracle logic here)."""
import hashlib
import os
import random

import torch

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def scatter_inputs(meta):
    """Re-create the inputs tools/make_golden.py fed to the reference's store_kvcache."""
    g = torch.Generator().manual_seed(meta["seed"])
    T, Hq, Hkv, D, nb = meta["T"], meta["Hq"], meta["Hkv"], meta["D"], meta["num_blocks"]
    qkv = torch.randn(T, (Hq + 2 * Hkv) * D, generator=g).bfloat16()
    q, k, v = qkv.split([Hq * D, Hkv * D, Hkv * D], dim=-1)
    k = k.reshape(T, Hkv, D).contiguous()
    v = v.view(T, Hkv, D)
    k_cache = torch.randn(nb, 256, Hkv, D, generator=g).bfloat16()
    v_cache = torch.randn(nb, 256, Hkv, D, generator=g).bfloat16()
    return qkv, k, v, k_cache, v_cache


def make_paged_case(seed, lens_q, lens_k, Hq, Hkv, D=128, page=256, extra_blocks=3,
                    pad=-1, table_width=None, dtype=torch.bfloat16, scale_std=1.0):
    """Random paged-KV problem.  Returns dict of CPU tensors.

    Every sequence gets its own shuffled physical blocks; the packed "logical" K/V
    (what an unpaged call would see) is returned too so paged == unpaged can be checked."""
    g = torch.Generator().manual_seed(seed)
    B = len(lens_k)
    nblk = [(l + page - 1) // page for l in lens_k]
    total_blocks = sum(nblk) + extra_blocks
    perm = torch.randperm(total_blocks, generator=g).tolist()
    width = table_width or max(max(nblk), 1)
    block_tables = torch.full((B, width), pad, dtype=torch.int32)
    k_cache = (scale_std * torch.randn(total_blocks, page, Hkv, D, generator=g)).to(dtype)
    v_cache = (scale_std * torch.randn(total_blocks, page, Hkv, D, generator=g)).to(dtype)
    cursor = 0
    k_log, v_log = [], []
    for b in range(B):
        blks = perm[cursor:cursor + nblk[b]]
        cursor += nblk[b]
        if blks:
            block_tables[b, :len(blks)] = torch.tensor(blks, dtype=torch.int32)
            kb = k_cache[blks].reshape(-1, Hkv, D)[:lens_k[b]]
            vb = v_cache[blks].reshape(-1, Hkv, D)[:lens_k[b]]
        else:
            kb = k_cache.new_zeros(0, Hkv, D)
            vb = v_cache.new_zeros(0, Hkv, D)
        k_log.append(kb)
        v_log.append(vb)
    Tq = sum(lens_q)
    q = (scale_std * torch.randn(Tq, Hq, D, generator=g)).to(dtype)
    cu_q = torch.tensor([0] + list(_cumsum(lens_q)), dtype=torch.int32)
    cu_k = torch.tensor([0] + list(_cumsum(lens_k)), dtype=torch.int32)
    return dict(q=q, k_cache=k_cache, v_cache=v_cache, block_tables=block_tables,
                cu_seqlens_q=cu_q, cu_seqlens_k=cu_k,
                context_lens=torch.tensor(lens_k, dtype=torch.int32),
                k_packed=torch.cat(k_log) if k_log else None,
                v_packed=torch.cat(v_log) if v_log else None,
                max_seqlen_q=max(lens_q) if lens_q else 0, max_seqlen_k=max(lens_k) if lens_k else 0)


def _cumsum(xs):
    s = 0
    for x in xs:
        s += x
        yield s


def bf16_ulp(x: torch.Tensor) -> torch.Tensor:
    """Size of one bf16 ulp at |x| (fp32 tensor)."""
    xf = x.float().abs().clamp_min(2.0 ** -126)
    e = torch.floor(torch.log2(xf))
    return torch.pow(2.0, e - 7)


# Cases for tests/golden/attn_ref_*.pt: one call of the reference's Attention.forward each
# (tools/make_golden_gpu.py runs them on the GPU box through the unmodified reference).
#   mode "unpaged" : sub-path A, Lq == Lk, block_tables None (attention.py:242-248)
#   mode "paged"   : chunked prefill / prefix hit / mixed batch over block_tables (attention.py:237-248)
#   mode "decode"  : CUDA-graph form, cu_seqlens_q None, table padded with 0 and `pad_rows` extra
#                    padding rows (context_len 0, slot -1) (attention.py:222-234, model_runner.py:589-596)
FORWARD_CASES = {
    "unpaged_q0p6b_tp4": dict(seed=101, mode="unpaged", Hq=4, Hkv=2, lens_q=[128, 1, 63, 300, 7], lens_k=[128, 1, 63, 300, 7]),
    "paged_mixed_g4": dict(seed=102, mode="paged", Hq=4, Hkv=1, lens_q=[188, 1, 128, 1, 1, 257, 64, 40],
                           lens_k=[700, 129, 128, 1, 257, 513, 64, 1000]),
    "paged_chunk_g5": dict(seed=103, mode="paged", Hq=5, Hkv=1, lens_q=[100, 1, 31, 129], lens_k=[356, 300, 31, 600]),
    "paged_g2_hkv2": dict(seed=104, mode="paged", Hq=4, Hkv=2, lens_q=[256, 3, 1], lens_k=[512, 260, 77]),
    "decode_g2_hkv8": dict(seed=105, mode="decode", Hq=16, Hkv=8, lens_q=[1] * 8, lens_k=[4096, 1, 255, 256, 257, 1300, 33, 777],
                           pad_rows=2, table_width=20),
    "decode_g5_hkv1": dict(seed=106, mode="decode", Hq=5, Hkv=1, lens_q=[1] * 5, lens_k=[513, 2, 1024, 100, 3000], pad_rows=1,
                           table_width=16),
    "decode_g8_hkv1": dict(seed=107, mode="decode", Hq=8, Hkv=1, lens_q=[1] * 3, lens_k=[2049, 640, 5], pad_rows=0, table_width=12),
}


def make_forward_case(meta):
    """CPU tensors for one `Attention.forward(q, k, v)` call plus its Context fields.

    The cache starts with random content everywhere (stale data in the slots the step writes),
    the step's new K/V rows are fresh random values, and slot_mapping addresses the last Lq
    positions of every sequence exactly as model_runner.py:453-481 computes them."""
    Hq, Hkv, D, page = meta["Hq"], meta["Hkv"], 128, 256
    lens_q, lens_k = meta["lens_q"], meta["lens_k"]
    decode = meta["mode"] == "decode"
    c = make_paged_case(meta["seed"], lens_q, lens_k, Hq, Hkv, pad=0 if decode else -1,
                        table_width=meta.get("table_width"))
    g = torch.Generator().manual_seed(meta["seed"] + 7)
    T = sum(lens_q)
    qkv = torch.randn(T, (Hq + 2 * Hkv) * D, generator=g).bfloat16()
    q = qkv[:, :Hq * D].reshape(T, Hq, D).contiguous()                    # rotary output: fresh tensors
    k = qkv[:, Hq * D:(Hq + Hkv) * D].reshape(T, Hkv, D).contiguous()
    v = qkv[:, (Hq + Hkv) * D:].view(T, Hkv, D)                           # strided view of qkv
    slots = []
    for b, (lq, lk) in enumerate(zip(lens_q, lens_k)):
        for p in range(lk - lq, lk):
            slots.append(int(c["block_tables"][b, p // page]) * page + p % page)
    slot_mapping = torch.tensor(slots, dtype=torch.int32)
    block_tables, context_lens = c["block_tables"], c["context_lens"]
    pad_rows = meta.get("pad_rows", 0) if decode else 0
    if pad_rows:
        q = torch.cat([q, torch.zeros(pad_rows, Hq, D, dtype=q.dtype)])
        qkv = torch.cat([qkv, torch.zeros(pad_rows, qkv.shape[1], dtype=qkv.dtype)])
        k = torch.cat([k, torch.zeros(pad_rows, Hkv, D, dtype=k.dtype)])
        v = qkv[:, (Hq + Hkv) * D:].view(T + pad_rows, Hkv, D)
        slot_mapping = torch.cat([slot_mapping, torch.full((pad_rows,), -1, dtype=torch.int32)])
        context_lens = torch.cat([context_lens, torch.zeros(pad_rows, dtype=torch.int32)])
        block_tables = torch.cat([block_tables, torch.zeros(pad_rows, block_tables.shape[1], dtype=torch.int32)])
    if decode:
        ctx = (None, None, 0, 0, slot_mapping, context_lens, block_tables, None)
    elif meta["mode"] == "unpaged":
        ctx = (c["cu_seqlens_q"], c["cu_seqlens_k"], c["max_seqlen_q"], c["max_seqlen_k"], slot_mapping, context_lens, None, None)
    else:
        ctx = (c["cu_seqlens_q"], c["cu_seqlens_k"], c["max_seqlen_q"], c["max_seqlen_k"], slot_mapping, context_lens,
               block_tables, None)
    return dict(q=q, k=k, v=v, qkv=qkv, k_cache=c["k_cache"], v_cache=c["v_cache"], ctx=ctx, num_real=T)


def tensor_sha256(t: torch.Tensor) -> str:
    """Digest of a 16-bit tensor's bytes (same definition as tools/make_golden_gpu.py)."""
    return hashlib.sha256(t.contiguous().view(torch.int16).cpu().numpy().tobytes()).hexdigest()
