"""Shared seeded-input builders for the CPU and GPU tests (no oracle logic here)."""
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
