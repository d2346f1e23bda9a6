#!/usr/bin/env python
"""Kernel-level benchmark of every piece of the hot path on the BASELINE.json shapes, next to
the reference's own GPU path (flash-attn 2.8.3 + an index_copy stand-in for the Triton scatter)
on IDENTICAL tensors.  GPU box only.  Writes one JSON document (default gpurun_out/kernels.json).

Timing: CUDA events on the current stream, 5 warm-ups, N timed iterations, median and best;
working sets are far larger than the 126 MB L2 or are rotated over distinct buffers.
"""
import argparse
import json
import os
import statistics
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "nano-vllm-v1_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import torch  # noqa: E402

from b200attn import ops  # noqa: E402

try:
    import flash_attn
except Exception:  # pragma: no cover
    flash_attn = None

D, PAGE = 128, 256
SCALE = D ** -0.5
PEAK = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else \
    {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}


def timeit(fn, iters=20, warm=int(os.environ.get("BK_WARM", "5")), flush=None):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        if flush is not None:
            flush.max()                                  # read-only sweep > L2: evicts the working set, leaves no dirty lines
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return statistics.median(ts), min(ts)


def decode_case(name, B, ctx, Hq, Hkv, iters, results, ragged=False):
    nblk_per = (ctx + PAGE - 1) // PAGE
    nblk = B * nblk_per
    g = torch.Generator(device="cuda").manual_seed(0)
    kc = torch.randn(nblk, PAGE, Hkv, D, generator=g, device="cuda", dtype=torch.bfloat16)
    vc = torch.randn(nblk, PAGE, Hkv, D, generator=g, device="cuda", dtype=torch.bfloat16)
    q = torch.randn(B, Hq, D, generator=g, device="cuda", dtype=torch.bfloat16)
    bt = torch.randperm(nblk, generator=torch.Generator().manual_seed(1)).to(torch.int32).view(B, nblk_per).cuda()
    if ragged:
        lens = torch.randint(1, ctx + 1, (B,), generator=torch.Generator().manual_seed(2), dtype=torch.int32).cuda()
    else:
        lens = torch.full((B,), ctx, dtype=torch.int32, device="cuda")
    bytes_alg = int((lens.long().sum().item() * 2 * Hkv * D + B * 2 * Hq * D) * 2)
    # L2 flush buffer only needed when the working set is small
    flush = torch.zeros(256 * 1024 * 1024 // 4, dtype=torch.int32, device="cuda") if bytes_alg < 1e9 else None
    ours = lambda: ops.attn_decode(q, kc, vc, lens, bt, SCALE)  # noqa: E731
    med, best = timeit(ours, iters, flush=flush)
    # the same launch(es) replayed from a CUDA graph (how the engine runs decode): no host launch cost
    out_g = torch.empty_like(q)
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        ops.attn_decode(q, kc, vc, lens, bt, SCALE, out=out_g)
    torch.cuda.current_stream().wait_stream(side)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        ops.attn_decode(q, kc, vc, lens, bt, SCALE, out=out_g)
    gmed, gbest = timeit(graph.replay, iters, flush=flush)
    rec = {"kind": "decode", "graph_ms_median": gmed, "graph_ms_best": gbest, "graph_gbps": bytes_alg / gmed / 1e6,
           "shape_env": os.environ.get("B200ATTN_DECODE_SHAPE", ""), "B": B, "ctx": ctx, "Hq": Hq, "Hkv": Hkv, "ragged": ragged, "alg_bytes": bytes_alg,
           "ours_ms_median": med, "ours_ms_best": best, "ours_gbps": bytes_alg / med / 1e6,
           "ours_frac_measured_hbm": bytes_alg / med / 1e6 / PEAK["hbm_gbs"], "ours_frac_8TBps": bytes_alg / med / 1e6 / 8000,
           "num_splits": ops.load_library().b200_decode_plan(B, Hkv, nblk_per * PAGE)}
    if flash_attn is not None:
        fa = lambda: flash_attn.flash_attn_with_kvcache(q.view(B, 1, Hq, D), kc, vc, cache_seqlens=lens, block_table=bt,  # noqa: E731
                                                        softmax_scale=SCALE, causal=True)
        fmed, fbest = timeit(fa, iters, flush=flush)
        rec.update({"fa2_ms_median": fmed, "fa2_gbps": bytes_alg / fmed / 1e6, "speedup_vs_fa2": fmed / med})
        rec["max_abs_diff_vs_fa2"] = (ours().float() - fa().view(B, Hq, D).float()).abs().max().item()
    results[name] = rec
    print(name, json.dumps(rec), flush=True)
    del kc, vc


PS_ONLY = False    # --ps-only: time just the persistent prefill kernel (A/B sweeps over env knobs)


def pairs(lq, lk):
    return lq * (lk - lq) + lq * (lq + 1) // 2


def prefill_case(name, lens_q, lens_k, Hq, Hkv, iters, results, variants=(2, 1), shared_prefix=0):
    """Paged varlen prefill.  shared_prefix > 0: every sequence's first `shared_prefix` tokens
    map to the same physical blocks (cfg5: 64 branches over one 32k prefix)."""
    B = len(lens_q)
    g = torch.Generator(device="cuda").manual_seed(0)
    nblk_per = [(l + PAGE - 1) // PAGE for l in lens_k]
    shared_blocks = shared_prefix // PAGE
    total_blocks = shared_blocks + sum(n - shared_blocks for n in nblk_per) + 2
    kc = torch.randn(total_blocks, PAGE, Hkv, D, generator=g, device="cuda", dtype=torch.bfloat16)
    vc = torch.randn(total_blocks, PAGE, Hkv, D, generator=g, device="cuda", dtype=torch.bfloat16)
    width = max(nblk_per)
    bt = torch.full((B, width), -1, dtype=torch.int32)
    cursor = shared_blocks
    for b in range(B):
        own = nblk_per[b] - shared_blocks
        bt[b, :shared_blocks] = torch.arange(shared_blocks, dtype=torch.int32)
        bt[b, shared_blocks:nblk_per[b]] = torch.arange(cursor, cursor + own, dtype=torch.int32)
        cursor += own
    bt = bt.cuda()
    T = sum(lens_q)
    q = torch.randn(T, Hq, D, generator=g, device="cuda", dtype=torch.bfloat16)
    cu_q = torch.tensor([0] + list(torch.tensor(lens_q).cumsum(0)), dtype=torch.int32).cuda()
    cu_k = torch.tensor([0] + list(torch.tensor(lens_k).cumsum(0)), dtype=torch.int32).cuda()
    flops = 4 * D * Hq * sum(pairs(a, b) for a, b in zip(lens_q, lens_k))
    rec = {"kind": "prefill", "B": B, "T": T, "max_q": max(lens_q), "max_k": max(lens_k), "Hq": Hq, "Hkv": Hkv, "flops": flops}
    outs = {}
    for v in (() if PS_ONLY else variants):
        fn = lambda: ops.attn_varlen(q, kc, vc, cu_q, cu_k, max(lens_q), max(lens_k), SCALE, block_table=bt, variant=v)  # noqa: E731
        med, best = timeit(fn, iters, warm=2)
        tag = {1: "mma", 2: "tc2", 3: "tc1"}[v]
        rec.update({f"{tag}_ms_median": med, f"{tag}_ms_best": best, f"{tag}_tflops": flops / med / 1e9,
                    f"{tag}_frac_bf16_burst": flops / med / 1e9 / PEAK["bf16_tflops"],
                    f"{tag}_frac_bf16_sustained": flops / med / 1e9 / PEAK["bf16_tflops_sustained"]})
        outs[tag] = fn()
    # planned path: persistent tcgen05 kernel (+ decode kernel for single-token rows); the plan is built
    # once per engine step and shared by all layers, so it is timed separately
    plan = ops.attn_plan(cu_q, cu_k, T, Hq, Hkv, route_decode=True)
    pmed, _ = timeit(lambda: ops.attn_plan(cu_q, cu_k, T, Hq, Hkv, route_decode=True), iters, warm=2)
    fn = lambda: ops.attn_varlen_planned(q, kc, vc, cu_q, cu_k, max(lens_q), max(lens_k), SCALE, block_table=bt, plan=plan)  # noqa: E731
    med, best = timeit(fn, iters, warm=2)
    rec.update({"ps_ms_median": med, "ps_ms_best": best, "ps_tflops": flops / med / 1e9, "plan_ms_median": pmed,
                "ps_frac_bf16_burst": flops / med / 1e9 / PEAK["bf16_tflops"],
                "ps_frac_bf16_sustained": flops / med / 1e9 / PEAK["bf16_tflops_sustained"]})
    outs["ps"] = fn()
    if flash_attn is not None and not PS_ONLY:
        fa = lambda: flash_attn.flash_attn_varlen_func(q, kc, vc, cu_seqlens_q=cu_q, cu_seqlens_k=cu_k, max_seqlen_q=max(lens_q),  # noqa: E731
                                                       max_seqlen_k=max(lens_k), softmax_scale=SCALE, causal=True, block_table=bt)
        fmed, fbest = timeit(fa, iters, warm=2)
        rec.update({"fa2_ms_median": fmed, "fa2_tflops": flops / fmed / 1e9})
        ref = fa()
        for tag, o in outs.items():
            rec[f"{tag}_max_abs_diff_vs_fa2"] = (o.float() - ref.float()).abs().max().item()
            rec[f"{tag}_speedup_vs_fa2"] = fmed / rec[f"{tag}_ms_median"]
    results[name] = rec
    print(name, json.dumps(rec), flush=True)
    del kc, vc


def pointwise_cases(results, iters):
    g = torch.Generator(device="cuda").manual_seed(0)
    # scatter: prefill-sized (cfg2 step 0, 8B shapes)
    for name, T, Hq, Hkv in (("scatter_T16384_hkv8", 16384, 32, 8), ("scatter_T512_hkv8", 512, 32, 8), ("scatter_T36864_hkv1", 36864, 8, 1)):
        qkv = torch.randn(T, (Hq + 2 * Hkv) * D, generator=g, device="cuda", dtype=torch.bfloat16)
        k = qkv[:, Hq * D:(Hq + Hkv) * D].reshape(T, Hkv, D).contiguous()
        v = qkv[:, (Hq + Hkv) * D:].view(T, Hkv, D)
        nb = T // PAGE + 64
        kc = torch.zeros(nb, PAGE, Hkv, D, device="cuda", dtype=torch.bfloat16)
        vc = torch.zeros_like(kc)
        slots = torch.randperm(nb * PAGE, generator=torch.Generator().manual_seed(3))[:T].to(torch.int32).cuda()
        flush = torch.zeros(256 * 1024 * 1024 // 4, dtype=torch.int32, device="cuda")
        med, best = timeit(lambda: ops.store_kvcache(k, v, kc, vc, slots), iters, flush=flush)
        b = T * (4 * Hkv * D * 2 + 4)
        rec = {"kind": "scatter", "T": T, "Hkv": Hkv, "alg_bytes": b, "ours_ms_median": med, "ours_gbps": b / med / 1e6,
               "ours_frac_measured_hbm": b / med / 1e6 / PEAK["hbm_gbs"]}
        idx = slots.long()
        tm, _ = timeit(lambda: (kc.view(-1, Hkv * D).index_copy_(0, idx, k.view(T, -1)), vc.view(-1, Hkv * D).index_copy_(0, idx, v.reshape(T, -1))), iters, flush=flush)
        rec["torch_index_copy_ms_median"] = tm
        results[name] = rec
        print(name, json.dumps(rec), flush=True)
    # norm+rope and silu at 8B TP1 shapes
    for T in (16384, 512):
        Hq, Hkv, I = 32, 8, 12288
        qkv = torch.randn(T, (Hq + 2 * Hkv) * D, generator=g, device="cuda", dtype=torch.bfloat16)
        q, k, _ = qkv.split([Hq * D, Hkv * D, Hkv * D], dim=-1)
        pos = torch.randint(0, 40960, (T,), device="cuda")
        import oracle
        cache = oracle.make_cos_sin_cache(D, 40960, 1e6).view(-1, D).cuda()
        w = torch.ones(D, device="cuda", dtype=torch.bfloat16)
        flush = torch.zeros(256 * 1024 * 1024 // 4, dtype=torch.int32, device="cuda")
        med, _ = timeit(lambda: ops.qk_norm_rope(q.view(T, Hq, D), k.view(T, Hkv, D), pos, w, w, cache, 1e-6), iters, flush=flush)
        b = T * (2 * (Hq + Hkv) * D * 2 + D * 4 + 8)
        results[f"norm_rope_T{T}"] = {"kind": "norm_rope", "T": T, "alg_bytes": b, "ours_ms_median": med, "ours_gbps": b / med / 1e6,
                                      "ours_frac_measured_hbm": b / med / 1e6 / PEAK["hbm_gbs"]}
        print(f"norm_rope_T{T}", json.dumps(results[f"norm_rope_T{T}"]), flush=True)
        # the launch the product path actually runs (install.py): q/k norm + RoPE + KV-cache scatter, k_out skipped
        v = qkv[:, (Hq + Hkv) * D:].view(T, Hkv, D)
        nb = T // PAGE + 64
        kc = torch.zeros(nb, PAGE, Hkv, D, device="cuda", dtype=torch.bfloat16)
        vc = torch.zeros_like(kc)
        slots = torch.randperm(nb * PAGE, generator=torch.Generator().manual_seed(3))[:T].to(torch.int32).cuda()
        for need_k in (False, True):
            med, _ = timeit(lambda: ops.qk_norm_rope_store(q.view(T, Hq, D), k.view(T, Hkv, D), v, pos, w, w, cache, 1e-6, kc, vc, slots,
                                                           need_k_out=need_k), iters, flush=flush)
            # read q,k,v + cos/sin + position + slot; write q_out, K and V cache rows (+ k_out)
            b = T * ((Hq + 2 * Hkv) * D * 2 + D * 4 + 12 + (Hq + 2 * Hkv + (Hkv if need_k else 0)) * D * 2)
            name = f"norm_rope_store_T{T}" + ("_kout" if need_k else "")
            results[name] = {"kind": "norm_rope_store", "T": T, "need_k_out": need_k, "alg_bytes": b, "ours_ms_median": med,
                             "ours_gbps": b / med / 1e6, "ours_frac_measured_hbm": b / med / 1e6 / PEAK["hbm_gbs"]}
            print(name, json.dumps(results[name]), flush=True)
        del kc, vc
        # hidden-size RMSNorm (73 launches per Qwen3-8B step), with and without the fused residual add
        Hd = 4096
        h = torch.randn(T, Hd, generator=g, device="cuda", dtype=torch.bfloat16)
        r = torch.randn(T, Hd, generator=g, device="cuda", dtype=torch.bfloat16)
        wh = torch.ones(Hd, device="cuda", dtype=torch.bfloat16)
        for resid in (False, True):
            med, _ = timeit(lambda: ops.rms_norm(h, wh, 1e-6, r if resid else None), iters, flush=flush)
            b = T * Hd * 2 * (4 if resid else 2)
            name = f"rmsnorm_T{T}" + ("_residual" if resid else "")
            results[name] = {"kind": "rmsnorm", "T": T, "hidden": Hd, "residual": resid, "alg_bytes": b, "ours_ms_median": med,
                             "ours_gbps": b / med / 1e6, "ours_frac_measured_hbm": b / med / 1e6 / PEAK["hbm_gbs"]}
            print(name, json.dumps(results[name]), flush=True)
        x = torch.randn(T, 2 * I, generator=g, device="cuda", dtype=torch.bfloat16)
        med, _ = timeit(lambda: ops.silu_and_mul(x), iters, flush=flush)
        b = T * 6 * I
        rec = {"kind": "silu_and_mul", "T": T, "alg_bytes": b, "ours_ms_median": med, "ours_gbps": b / med / 1e6,
               "ours_frac_measured_hbm": b / med / 1e6 / PEAK["hbm_gbs"]}
        tm, _ = timeit(lambda: torch.nn.functional.silu(x[:, :I]) * x[:, I:], iters, flush=flush)
        rec["torch_eager_ms_median"] = tm
        results[f"silu_T{T}"] = rec
        print(f"silu_T{T}", json.dumps(rec), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--what", default="decode,prefill,pointwise")
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "kernels.json"))
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--ps-only", action="store_true")
    a = ap.parse_args()
    global PS_ONLY
    PS_ONLY = a.ps_only
    results = {"peaks": PEAK, "gpu": torch.cuda.get_device_name(0)}
    what = a.what.split(",")
    if "decode" in what:
        for tp in (1, 2, 4, 8):                                     # cfg3 at TP=1/2/4/8 (per-rank shapes)
            decode_case(f"cfg3_decode_tp{tp}", 512, 4096, 32 // tp, 8 // tp, a.iters, results)
        decode_case("cfg3_decode_tp1_ragged", 512, 4096, 32, 8, a.iters, results, ragged=True)
        decode_case("cfg5_decode_tp8_bs64_ctx32832", 64, 32832, 8, 1, a.iters, results)
        decode_case("cfg2_decode_bs256_ctx384_0p6b", 256, 384, 16, 8, a.iters, results)
        decode_case("decode_bs1_ctx4096_8b", 1, 4096, 32, 8, a.iters, results)
        decode_case("decode_bs8_ctx16384_14b_tp2", 8, 16384, 20, 4, a.iters, results)
        decode_case("decode_bs256_ctx256_8b", 256, 256, 32, 8, a.iters, results)
        decode_case("decode_bs256_ctx1280_8b", 256, 1280, 32, 8, a.iters, results)
        decode_case("decode_bs256_ctx1280_8b_tp4", 256, 1280, 8, 2, a.iters, results)
        decode_case("decode_bs64_ctx2048_8b", 64, 2048, 32, 8, a.iters, results)
        decode_case("decode_bs16_ctx1024_8b", 16, 1024, 32, 8, a.iters, results)
    if "prefill" in what:
        prefill_case("prefill_1x8192_causal_hq8", [8192], [8192], 8, 1, max(a.iters // 2, 3), results, variants=(2, 3, 1))
        prefill_case("prefill_4x4096_causal_8b_tp1", [4096] * 4, [4096] * 4, 32, 8, max(a.iters // 2, 3), results)
        prefill_case("cfg5_prefix32k_64branches_tp8", [32832] + [64] * 63, [32832] * 64, 8, 1, 3, results, shared_prefix=32768)
        prefill_case("cfg2_step0_128x128_0p6b", [128] * 128, [128] * 128, 16, 8, a.iters, results, variants=(2, 3, 1))
        mixed_q = [128] * 8 + [1] * 439
        mixed_k = [128] * 8 + [128 + (i % 100) for i in range(439)]
        prefill_case("cfg4_mixed_447seqs_14b_tp2", mixed_q, mixed_k, 20, 4, a.iters, results, variants=(2, 3, 1))
        prefill_case("chunk_2048_of_16384_8b", [2048] * 2, [16384] * 2, 32, 8, max(a.iters // 2, 3), results)
    if "pointwise" in what:
        pointwise_cases(results, a.iters)
    os.makedirs(os.path.dirname(a.out), exist_ok=True)
    with open(a.out, "w") as f:
        json.dump(results, f, indent=1)
    print("wrote", a.out)


if __name__ == "__main__":
    main()
