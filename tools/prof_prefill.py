#!/usr/bin/env python
"""Tiny driver for ncu: a few launches of one prefill kernel on 4 x 4096 causal, 32q/8kv.
argv[1]: ps (persistent tcgen05, default) | tc (variant 2) | mma (variant 1)"""
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "nano-vllm-v1_b200")):
    sys.path.insert(0, p)
import torch
from b200attn import ops
which = sys.argv[1] if len(sys.argv) > 1 else "ps"
B, L, Hq, Hkv, D, P = 4, 4096, 32, 8, 128, 256
g = torch.Generator(device="cuda").manual_seed(0)
nb = B * L // P
kc = torch.randn(nb, P, Hkv, D, generator=g, device="cuda", dtype=torch.bfloat16)
vc = torch.randn(nb, P, Hkv, D, generator=g, device="cuda", dtype=torch.bfloat16)
q = torch.randn(B * L, Hq, D, generator=g, device="cuda", dtype=torch.bfloat16)
bt = torch.arange(nb, dtype=torch.int32, device="cuda").view(B, -1)
cu = torch.arange(0, (B + 1) * L, L, dtype=torch.int32, device="cuda")
plan = ops.attn_plan(cu, cu, B * L, Hq, Hkv, route_decode=True)
for _ in range(3):
    if which == "ps":
        o = ops.attn_varlen_planned(q, kc, vc, cu, cu, L, L, D ** -0.5, block_table=bt, plan=plan)
    else:
        o = ops.attn_varlen(q, kc, vc, cu, cu, L, L, D ** -0.5, block_table=bt, variant={"tc": 2, "mma": 1}[which])
torch.cuda.synchronize()
print("ok", o.float().abs().mean().item())
