#!/usr/bin/env python
"""GPU baseline of bench.py, run as a SUBPROCESS so that the bench process itself never loads flash-attn or Triton:
the UNMODIFIED reference `Attention.forward` (baseline/_ref: Triton store_kvcache + flash_attn_with_kvcache, flash-attn
2.8.3) on the identical seeded cfg3 tensors of bench.build_problem, same step definition (scatter + paged decode of one
layer, graph-decode form of the Context), CUDA events, >= 3 warm-ups.  Prints one JSON line.

    python tools/ref_attention_bench.py --rank 0 --world 1 --device 0 --steps 20 --warmup 3
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rank", type=int, default=0)
    ap.add_argument("--world", type=int, default=1)
    ap.add_argument("--device", type=int, default=0)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    a = ap.parse_args()
    ref = os.environ.get("NANOVLLM_REFERENCE", os.path.join(ROOT, "baseline", "_ref"))
    if not os.path.isdir(os.path.join(ref, "nanovllm")):
        print(json.dumps({"unavailable": f"no reference install at {ref}"}))
        return 0
    sys.path.insert(0, ref)
    sys.path.insert(0, ROOT)
    import torch
    import bench                                                   # problem builder only (no kernels of this repo)
    from nanovllm.layers.attention import Attention
    from nanovllm.utils.context import reset_context, set_context
    import flash_attn

    torch.cuda.set_device(a.device)
    device = torch.device("cuda", a.device)
    pr = bench.build_problem(a.rank, a.world, device)
    hq, hkv, D = pr["hq"], pr["hkv"], bench.D
    attn = Attention(hq, D, D ** -0.5, hkv)
    attn.k_cache, attn.v_cache = pr["k_cache"], pr["v_cache"]
    bt, cl, sm = pr["block_tables"].to(device), pr["context_lens"].to(device), pr["slot_mapping"].to(device)
    qkv = pr["qkv"].to(device)
    q, k, v = bench.split_qkv(qkv, hq, hkv)
    q, k = q.contiguous(), k.contiguous()

    def step():
        set_context(None, None, 0, 0, sm, cl, bt, None)           # graph-decode form (attention.py:222)
        return attn(q, k, v)

    with torch.inference_mode():
        for _ in range(max(a.warmup, 3)):
            step()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(a.steps):
            step()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / a.steps
        # The same steps replayed from ONE CUDA graph, the way the reference engine itself runs decode
        # (engine/model_runner.py:640-715) and the way bench.py times this repo's step: no per-step Python / launch cost.
        ms_graph = None
        try:
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                step()
            torch.cuda.current_stream().wait_stream(side)
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                for _ in range(a.steps):
                    step()
            graph.replay()
            torch.cuda.synchronize()
            e0.record()
            graph.replay()
            e1.record()
            torch.cuda.synchronize()
            ms_graph = e0.elapsed_time(e1) / a.steps
        except Exception as exc:                                   # noqa: BLE001 -- the eager number still stands
            ms_graph_error = f"{type(exc).__name__}: {exc}"[:200]
    reset_context()
    dec_b, sc_b = bench.algorithmic_bytes(bench.B, bench.CTX, hq, hkv)
    print(json.dumps({"impl": "reference Attention.forward (Triton store_kvcache + flash_attn_with_kvcache)",
                      "flash_attn": flash_attn.__version__, "ms_per_step": round(ms, 4),
                      "ms_per_step_graph": round(ms_graph, 4) if ms_graph else None,
                      "gbps_per_gpu": round((dec_b + sc_b) / (ms * 1e-3) / 1e9, 1), "rank": a.rank, "world": a.world}))
    return 0


if __name__ == "__main__":
    sys.exit(main())
