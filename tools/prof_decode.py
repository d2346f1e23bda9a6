#!/usr/bin/env python
"""Tiny driver for ncu: two launches of the paged decode kernel per shape (cfg3 at TP1 and TP8 heads, cfg5 decode, cfg2 decode)."""
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "nano-vllm-v1_b200")):
    sys.path.insert(0, p)
import torch
from b200attn import ops
D, PAGE = 128, 256
for name, B, ctx, Hq, Hkv in (("cfg3_tp1", 512, 4096, 32, 8), ("cfg3_tp8", 512, 4096, 4, 1), ("cfg5_tp8", 64, 32832, 8, 1), ("cfg2", 256, 384, 16, 8)):
    per = (ctx + PAGE - 1) // PAGE
    g = torch.Generator(device="cuda").manual_seed(0)
    kc = torch.randn(B * per, PAGE, Hkv, D, generator=g, device="cuda", dtype=torch.bfloat16)
    vc = torch.randn(B * per, PAGE, Hkv, D, generator=g, device="cuda", dtype=torch.bfloat16)
    q = torch.randn(B, Hq, D, generator=g, device="cuda", dtype=torch.bfloat16)
    bt = torch.randperm(B * per, generator=torch.Generator().manual_seed(1)).to(torch.int32).view(B, per).cuda()
    lens = torch.full((B,), ctx, dtype=torch.int32, device="cuda")
    for _ in range(2):
        o = ops.attn_decode(q, kc, vc, lens, bt, D ** -0.5)
    torch.cuda.synchronize()
    print(name, "ok", o.float().abs().mean().item())
    del kc, vc
