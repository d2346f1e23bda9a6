#!/usr/bin/env python
"""Bring-up checker for the tcgen05 prefill kernel: runs a ladder of cases from trivial to
ragged against the CPU oracle and prints per-case error statistics (GPU box only)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "nano-vllm-v1_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import torch  # noqa: E402

import oracle  # noqa: E402
from b200attn import ops  # noqa: E402
from helpers import make_paged_case  # noqa: E402

SCALE = 128 ** -0.5
variants = [int(v) for v in (sys.argv[1].split(",") if len(sys.argv) > 1 else ["3", "2"])]


def report(name, got, want):
    g, w = got.float().cpu(), want.float()
    err = (g - w).abs()
    bad = err > 2e-2 * w.abs().clamp_min(1.0)          # 2e-2 abs at unit scale; relative beyond it
    msg = f"{name}: max_err {err.max().item():.4g} nan {int(torch.isnan(g).sum())} bad {int(bad.sum())}/{bad.numel()}"
    if bad.any():
        idx = bad.nonzero()
        rows = sorted(set(idx[:, 0].tolist()))
        heads = sorted(set(idx[:, 1].tolist()))
        cols = sorted(set(idx[:, 2].tolist()))
        msg += f" rows[{rows[0]}..{rows[-1]}] n={len(rows)} heads {heads[:8]} cols[{cols[0]}..{cols[-1]}] n={len(cols)}"
        r, h, c = idx[0].tolist()
        msg += f" first ({r},{h},{c}) got {g[r, h, c]:.4f} want {w[r, h, c]:.4f}"
    print(msg, flush=True)
    return not bad.any() and not torch.isnan(g).any()


def run(name, lens_q, lens_k, Hq, Hkv, variant, paged=True, seed=1, std=1.0):
    c = make_paged_case(seed, lens_q, lens_k, Hq, Hkv, pad=-1, scale_std=std)
    if paged:
        want = oracle.attn_varlen(c["q"], c["k_cache"], c["v_cache"], c["cu_seqlens_q"], c["cu_seqlens_k"], SCALE, c["block_tables"])
        got = ops.attn_varlen(c["q"].cuda(), c["k_cache"].cuda(), c["v_cache"].cuda(), c["cu_seqlens_q"].cuda(),
                              c["cu_seqlens_k"].cuda(), c["max_seqlen_q"], c["max_seqlen_k"], SCALE,
                              block_table=c["block_tables"].cuda(), variant=variant)
    else:
        want = oracle.attn_varlen(c["q"], c["k_packed"], c["v_packed"], c["cu_seqlens_q"], c["cu_seqlens_k"], SCALE)
        got = ops.attn_varlen(c["q"].cuda(), c["k_packed"].cuda(), c["v_packed"].cuda(), c["cu_seqlens_q"].cuda(),
                              c["cu_seqlens_k"].cuda(), c["max_seqlen_q"], c["max_seqlen_k"], SCALE, variant=variant)
    torch.cuda.synchronize()
    return report(f"v{variant} {name}", got, want)


ok = True
for v in variants:
    ok &= run("1seq 128x128 1head", [128], [128], 1, 1, v)
    ok &= run("1seq 128x128 unpaged", [128], [128], 1, 1, v, paged=False)
    ok &= run("1seq 256x256", [256], [256], 2, 1, v)
    ok &= run("1seq 100x100 (tail)", [100], [100], 2, 1, v)
    ok &= run("1seq 384x384", [384], [384], 4, 2, v)
    ok &= run("chunk 64 of 500", [64], [500], 4, 1, v)
    ok &= run("1seq 1000 causal", [1000], [1000], 8, 2, v)
    ok &= run("ragged mixed", [188, 1, 128, 1, 1, 257, 64, 40], [700, 129, 128, 1, 257, 513, 64, 1000], 16, 8, v)
    ok &= run("ragged unpaged", [128, 1, 63, 300, 7], [128, 1, 63, 300, 7], 5, 1, v, paged=False)
    ok &= run("big values", [300], [300], 4, 1, v, std=4.0)
print("ALL OK" if ok else "FAILURES", flush=True)
sys.exit(0 if ok else 1)
