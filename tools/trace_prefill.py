#!/usr/bin/env python
"""Phase timeline of the persistent prefill kernel (CTA 0) from its clock64 trace build.
GPU box only.  Prints, per softmax warpgroup, the median cycles of every phase of a K/V tile in steady
state, and the MMA issuer's wait/issue times; writes the raw stamps to --out."""
import argparse
import json
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "nano-vllm-v1_b200")):
    sys.path.insert(0, p)
import torch
from b200attn import ops, load_library

ap = argparse.ArgumentParser()
ap.add_argument("--tiles", type=int, default=96)
ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "prefill_trace.json"))
ap.add_argument("--solo", action="store_true", help="one query tile per CTA (odd group, 128-row chunks over a long prefix): "
                "softmax warpgroup 0 runs with no second softmax warp on its SM sub-partitions")
a = ap.parse_args()
B, L, Hq, Hkv, D, P = 4, 4096, 32, 8, 128, 256
if a.solo:
    B, L, Hq, Hkv = 40, 8192, 4, 4
g = torch.Generator(device="cuda").manual_seed(0)
nb = B * L // P
kc = torch.randn(nb, P, Hkv, D, generator=g, device="cuda", dtype=torch.bfloat16)
vc = torch.randn(nb, P, Hkv, D, generator=g, device="cuda", dtype=torch.bfloat16)
q = torch.randn(B * L, Hq, D, generator=g, device="cuda", dtype=torch.bfloat16)
bt = torch.arange(nb, dtype=torch.int32, device="cuda").view(B, -1)
cu = torch.arange(0, (B + 1) * L, L, dtype=torch.int32, device="cuda")
if a.solo:
    LQ = 128
    q = q[:B * LQ].contiguous()
    cuq = torch.arange(0, (B + 1) * LQ, LQ, dtype=torch.int32, device="cuda")
    plan = ops.attn_plan(cuq, cu, B * LQ, Hq, Hkv, route_decode=True)
    run = lambda: ops.attn_varlen_planned(q, kc, vc, cuq, cu, LQ, L, D ** -0.5, block_table=bt, plan=plan)  # noqa: E731
else:
    plan = ops.attn_plan(cu, cu, B * L, Hq, Hkv, route_decode=True)
    run = lambda: ops.attn_varlen_planned(q, kc, vc, cu, cu, L, L, D ** -0.5, block_table=bt, plan=plan)  # noqa: E731
for _ in range(3):
    run()
torch.cuda.synchronize()
lib = load_library()
buf = torch.zeros(6, a.tiles, 8, dtype=torch.int64, device="cuda")
lib.b200_debug_prefill_trace(buf.data_ptr(), a.tiles)
run()
torch.cuda.synchronize()
lib.b200_debug_prefill_trace(None, 0)
t = buf.cpu()
t0 = int(t[t > 0].min())
rel = torch.where(t > 0, t - t0, torch.zeros_like(t))
names = ["wait S", "S ready->loaded", "max (+mask/rescale)", "exp+cvt+st issue", "st wait+fence+arrive"]
lo, hi = a.tiles // 4, a.tiles - 2
for wg in ((0,) if a.solo else (0, 1)):
    r = rel[wg, lo:hi].double()
    d = (r[:, 1:6] - r[:, 0:5])
    per = (r[1:, 1] - r[:-1, 1])
    print(f"softmax WG{wg}: tile period median {per.median().item():.0f} cyc;",
          ", ".join(f"{n} {d[:, i].median().item():.0f}" for i, n in enumerate(names)))
    if (r[:, 6] > 0).all():
        print(f"   inside the max phase: loads issued -> all landed {(r[:, 6] - r[:, 2]).median().item():.0f}, landed -> row max known "
              f"{(r[:, 7] - r[:, 6]).median().item():.0f}, max known -> phase end {(r[:, 3] - r[:, 7]).median().item():.0f}")
for wg in ((0,) if a.solo else (0, 1)):
    r = rel[2 + wg, lo:hi].double()
    print(f"MMA issuer for WG{wg}: waits for P {((r[:, 1] - r[:, 0]).median().item()):.0f} cyc, issue PV+QK {((r[:, 2] - r[:, 1]).median().item()):.0f} cyc,"
          f" P published -> seen {(rel[2 + wg, lo:hi, 1] - rel[wg, lo:hi, 5]).double().median().item():.0f}")
if a.solo:
    sys.exit(0)
r = rel[2, lo:hi].double()
print(f"MMA issuer loop top: wait V(j) {(r[:, 5] - r[:, 4]).median().item():.0f} cyc, wait K(j+1) {(r[:, 6] - r[:, 5]).median().item():.0f} cyc,"
      f" previous tile's last issue -> loop top {(rel[2, lo + 1:hi + 1, 4] - rel[3, lo:hi, 2]).double().median().item():.0f} cyc")
# S-ready latency: MMA issue done (PV j + QK j+1 queued) -> softmax sees S(j+1)
for wg in (0, 1):
    lat = (rel[wg, lo + 1:hi + 1, 1] - rel[2 + wg, lo:hi, 2]).double()
    print(f"WG{wg}: MMA issued -> next S ready {lat.median().item():.0f} cyc")
os.makedirs(os.path.dirname(a.out), exist_ok=True)
json.dump({"shape": "4x4096 causal 32q/8kv, CTA 0", "tiles": a.tiles, "stamps_rel_cycles": rel.tolist()}, open(a.out, "w"))
print("wrote", a.out)
