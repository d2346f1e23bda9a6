#!/usr/bin/env python
"""One (CTA shape, split count) point of the paged decode kernel on the sharded shapes (cfg3 at TP2/4/8 heads, cfg5 decode):
graph-replayed us and GB/s.  Shape and splits come from B200ATTN_DECODE_SHAPE / B200ATTN_DECODE_SPLITS (read once per
process), so tools/gpu_sweep_decode_tp.sh runs one process per point."""
import json
import os
import statistics
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "nano-vllm-v1_b200")):
    sys.path.insert(0, p)
import torch
from b200attn import ops
D, PAGE = 128, 256
cases = {"cfg3_tp8": (512, 4096, 4, 1), "cfg3_tp4": (512, 4096, 8, 2), "cfg3_tp2": (512, 4096, 16, 4), "cfg5_tp8": (64, 32832, 8, 1),
         "cfg2": (256, 384, 16, 8), "bs256_ctx256": (256, 256, 32, 8), "bs256_ctx1280": (256, 1280, 32, 8), "bs64_ctx2048": (64, 2048, 32, 8),
         "bs256_ctx1280_tp4": (256, 1280, 8, 2), "cfg3_tp1": (512, 4096, 32, 8)}
out = {}
for name in sys.argv[1:] or list(cases):
    B, ctx, Hq, Hkv = cases[name]
    per = (ctx + PAGE - 1) // PAGE
    g = torch.Generator(device="cuda").manual_seed(0)
    kc = torch.randn(B * per, PAGE, Hkv, D, generator=g, device="cuda", dtype=torch.bfloat16)
    vc = torch.randn(B * per, PAGE, Hkv, D, generator=g, device="cuda", dtype=torch.bfloat16)
    q = torch.randn(B, Hq, D, generator=g, device="cuda", dtype=torch.bfloat16)
    bt = torch.randperm(B * per, generator=torch.Generator().manual_seed(1)).to(torch.int32).view(B, per).cuda()
    lens = torch.full((B,), ctx, dtype=torch.int32, device="cuda")
    o = torch.empty_like(q)
    flush = torch.zeros(256 * 1024 * 1024 // 4, dtype=torch.int32, device="cuda")
    for _ in range(3):
        ops.attn_decode(q, kc, vc, lens, bt, D ** -0.5, out=o)
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr):
        ops.attn_decode(q, kc, vc, lens, bt, D ** -0.5, out=o)
    ts = []
    for _ in range(15):
        flush.max()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); gr.replay(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ms = statistics.median(ts)
    nbytes = (B * ctx * 2 * Hkv * D + B * 2 * Hq * D) * 2
    out[name] = {"us": round(ms * 1e3, 1), "gbps": round(nbytes / ms / 1e6)}
    del kc, vc
print(os.environ.get("B200ATTN_DECODE_SHAPE", "default"), os.environ.get("B200ATTN_DECODE_SPLITS", "auto"), json.dumps(out), flush=True)
