#!/usr/bin/env python
"""End-to-end run of the reference's UNTOUCHED `LLM.generate` (offline install under
baseline/_ref), either as shipped (`--arm reference`: Triton + flash-attn) or with the B200
drop-in applied (`--arm dropin`).  GPU box only.  One JSON document per run.

    python tools/e2e_generate.py --arm dropin --size 0.6b --num-seqs 256 --in-len 128 --out-len 256 \
        --chunked-prefill --out gpurun_out/e2e_dropin.json
    python tools/e2e_generate.py --compare gpurun_out/e2e_reference_greedy.json gpurun_out/e2e_dropin_greedy.json

Both arms get the same harness patches (tools/e2e/sitecustomize.py): seeded random weights and,
with --greedy, an argmax sampler that records top-2 logit margins.  Throughput = generated tokens /
wall time of generate(), as the reference's bench.py:26-27 computes it.
"""
import argparse
import json
import os
import random
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def launcher(a):
    env = dict(os.environ)
    paths = [os.path.join(ROOT, "tools", "e2e"), os.path.join(ROOT, "nano-vllm-v1_b200"),
             os.environ.get("NANOVLLM_REFERENCE", os.path.join(ROOT, "baseline", "_ref"))]
    env["PYTHONPATH"] = os.pathsep.join(paths + [env.get("PYTHONPATH", "")])
    env["B200E2E_RANDOM_INIT"] = str(a.seed)
    env["B200E2E_SHM_BYTES"] = str(64 << 20)
    env.pop("B200ATTN_DROPIN", None)
    if a.arm == "dropin":
        env["B200ATTN_DROPIN"] = "1"
    elif a.arm == "dropin-attn":
        env["B200ATTN_DROPIN"] = "attn"
    if a.step_profile:
        env["B200E2E_STEP_PROFILE"] = os.path.abspath(a.step_profile)
        if a.kernel_profile:
            env["B200E2E_KERNEL_PROFILE"] = os.path.abspath(a.kernel_profile)
    if a.fp32_logits:
        env["B200E2E_FP32_LOGITS"] = "1"
    if a.greedy:
        env["B200E2E_GREEDY"] = "1"
        env["B200E2E_MARGINS_OUT"] = a.out + ".margins"
        if a.force_tokens:
            env["B200E2E_FORCE_TOKENS"] = os.path.abspath(a.force_tokens)
    cmd = [sys.executable, os.path.abspath(__file__), "--worker"] + [x for x in sys.argv[1:] if x != "--worker"]
    return subprocess.run(cmd, env=env).returncode


def worker(a):
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    from make_fake_model import make
    model_dir = make(a.size, f"/tmp/b200_fake_qwen3_{a.size}_{a.layers or 'full'}", a.layers)
    import torch
    from nanovllm import LLM, SamplingParams
    rnd = random.Random(0)                                                 # as the reference's bench.py:9
    prompts = [[rnd.randint(0, 10000) for _ in range(a.in_len)] for _ in range(a.num_seqs)]
    if a.shared_prefix:
        prefix = [rnd.randint(0, 10000) for _ in range(a.shared_prefix)]
        prompts = [prefix + p for p in prompts]
    t0 = time.perf_counter()
    llm = LLM(model_dir, enforce_eager=a.enforce_eager, max_model_len=a.max_model_len,
              tensor_parallel_size=a.tp, chunked_prefill=a.chunked_prefill, max_num_seqs=a.max_num_seqs,
              max_num_batched_tokens=a.max_num_batched_tokens, gpu_memory_utilization=a.gpu_memory_utilization)
    t_init = time.perf_counter() - t0
    sp = SamplingParams(temperature=0.6, ignore_eos=True, max_tokens=a.out_len)
    llm.generate([[1, 2, 3]], SamplingParams(temperature=0.6, ignore_eos=True, max_tokens=4), use_tqdm=False)   # warm-up, bench.py:22
    torch.cuda.synchronize()
    from nanovllm.engine.sequence import Sequence
    first_id = next(Sequence.counter) + 1                                  # ids below belong to warm-up sequences
    os.environ["B200E2E_FIRST_SEQ_ID"] = str(first_id)
    times = []
    for _ in range(max(a.repeat, 1)):                                      # first pass includes torch.compile re-traces
        t0 = time.perf_counter()
        outs = llm.generate(prompts, sp, use_tqdm=False)
        torch.cuda.synchronize()
        times.append(time.perf_counter() - t0)
        if a.greedy:
            break                                                          # recorded decisions must be those of ONE pass
    dt = min(times)
    import b200attn.install as dropin_mod
    n_out = sum(len(o["token_ids"]) for o in outs)
    rec = dict(arm=a.arm, size=a.size, layers=a.layers, tp=a.tp, num_seqs=a.num_seqs, in_len=a.in_len, out_len=a.out_len,
               shared_prefix=a.shared_prefix, chunked_prefill=a.chunked_prefill, enforce_eager=a.enforce_eager,
               greedy=a.greedy, seconds=dt, init_seconds=t_init, generated_tokens=n_out, tok_per_s=n_out / dt, seconds_all=times, first_seq_id=first_id,
               total_tok_per_s=(n_out + sum(len(p) for p in prompts)) / dt, dropin_installed=dropin_mod.installed(),
               gpu=torch.cuda.get_device_name(0), token_ids=[o["token_ids"] for o in outs] if a.greedy else None)
    os.makedirs(os.path.dirname(os.path.abspath(a.out)), exist_ok=True)
    with open(a.out, "w") as f:
        json.dump(rec, f)
    print(json.dumps({k: v for k, v in rec.items() if k != "token_ids"}), flush=True)
    return 0


def compare(ref_path, our_path, steps, tie_margin):
    """Greedy parity over `steps` decisions per sequence.  `our_path` is a TEACHER-FORCED run (it was fed
    the reference's tokens), so every decision of both arms was taken on an identical context: the
    arg-max token ids must be identical, except where the top-2 logit margin is within `tie_margin` in BOTH
    arms -- a tie that rounding legitimately flips.
    Prints a JSON verdict; exit status 1 if any decision differs outside a near-tie."""
    ref, our = json.load(open(ref_path)), json.load(open(our_path))
    mref = json.load(open(ref_path + ".margins"))
    mour = json.load(open(our_path + ".margins"))
    n = len(ref["token_ids"])
    total = same = tie = real = 0
    seq_identical = 0
    worst, margins = [], []
    for i in range(n):
        r = {rec[0]: (rec[1], rec[2]) for rec in mref[str(i + ref["first_seq_id"])]}
        o = {rec[0]: (rec[1], rec[2]) for rec in mour[str(i + our["first_seq_id"])]}
        ok_seq = True
        for st in range(steps):
            if st not in r or st not in o:
                continue
            total += 1
            margins.append(r[st][1])
            if r[st][0] == o[st][0]:
                same += 1
                continue
            ok_seq = False
            if max(r[st][1], o[st][1]) <= tie_margin:             # a tie in BOTH arms
                tie += 1
            else:
                real += 1
            worst.append(dict(seq=i, step=st, ref_margin=round(r[st][1], 5), our_margin=round(o[st][1], 5)))
        seq_identical += ok_seq
    # arm-to-arm logit noise on identical contexts: |top-1 logit of the reference - top-1 logit of the drop-in| over the
    # decisions where both picked the same token (records carry the top-1 value as a fourth field)
    noise = []
    for i in range(n):
        rr = {rec[0]: rec for rec in mref[str(i + ref["first_seq_id"])]}
        oo = {rec[0]: rec for rec in mour[str(i + our["first_seq_id"])]}
        for st in range(steps):
            if st in rr and st in oo and len(rr[st]) > 3 and len(oo[st]) > 3 and rr[st][1] == oo[st][1]:
                noise.append(abs(rr[st][3] - oo[st][3]))
    noise.sort()
    nq = lambda f: noise[min(int(len(noise) * f), len(noise) - 1)] if noise else None  # noqa: E731
    margins.sort()
    worst.sort(key=lambda d: -min(d["ref_margin"], d["our_margin"]))
    verdict = dict(sequences=n, steps=steps, decisions=total, identical=same, differ_at_near_tie=tie, differ_for_real=real,
                   top1_logit_abs_diff=dict(p50=nq(0.5), p90=nq(0.9), p99=nq(0.99), max=nq(1.0)),
                   sequences_fully_identical=seq_identical, tie_margin=tie_margin, largest_margin_differences=worst[:8],
                   ref_margin_quantiles=dict(p0=margins[0], p1=margins[len(margins) // 100], p10=margins[len(margins) // 10],
                                             p50=margins[len(margins) // 2]) if margins else None,
                   teacher_forced=True)
    print(json.dumps(verdict))
    return verdict


def compare_free(ref_path, our_path, steps, tie_margin):
    """Greedy parity of two FREE-RUNNING runs (untouched LLM.generate in both arms, argmax sampler).
    Both arms see identical contexts up to a sequence's first differing token, so there every recorded decision
    is comparable: the token ids must be identical, and the first difference of a sequence (if any) must sit on a
    decision whose top-2 logit margin is <= `tie_margin` in BOTH arms (a tie that rounding legitimately flips);
    everything after it is a different context and is not compared.  The verdict also carries the quantiles of the
    reference arm's margins over the compared decisions: a caller must reject p1 <= tie_margin (a degenerate
    model where the tie class could swallow the comparison)."""
    ref, our = json.load(open(ref_path)), json.load(open(our_path))
    mref = json.load(open(ref_path + ".margins"))
    mour = json.load(open(our_path + ".margins"))
    n = len(ref["token_ids"])
    compared = identical_seqs = ties = real = 0
    margins, first_div = [], []
    for i in range(n):
        r = {rec[0]: (rec[1], rec[2]) for rec in mref[str(i + ref["first_seq_id"])]}
        o = {rec[0]: (rec[1], rec[2]) for rec in mour[str(i + our["first_seq_id"])]}
        tr, to = ref["token_ids"][i][:steps], our["token_ids"][i][:steps]
        div = next((t for t in range(min(len(tr), len(to))) if tr[t] != to[t]), None)
        upto = div if div is not None else min(len(tr), len(to))
        for t in range(upto):
            margins.append(r[t][1])
        compared += upto
        if div is None:
            identical_seqs += 1
            continue
        compared += 1
        margins.append(r[div][1])
        rec = dict(seq=i, step=div, ref_margin=round(r[div][1], 6), our_margin=round(o[div][1], 6))
        first_div.append(rec)
        if max(r[div][1], o[div][1]) <= tie_margin:
            ties += 1
        else:
            real += 1
    margins.sort()
    q = lambda f: margins[min(int(len(margins) * f), len(margins) - 1)] if margins else None  # noqa: E731
    first_div.sort(key=lambda d: -max(d["ref_margin"], d["our_margin"]))
    verdict = dict(sequences=n, steps=steps, decisions_compared=compared, decisions_total=n * steps,
                   sequences_fully_identical=identical_seqs, first_divergences=len(first_div),
                   first_divergences_at_tie=ties, first_divergences_for_real=real, tie_margin=tie_margin,
                   largest_margin_first_divergences=first_div[:8],
                   ref_margin_quantiles=dict(p0=q(0.0), p0_1=q(0.001), p1=q(0.01), p10=q(0.1), p50=q(0.5)),
                   teacher_forced=False)
    print(json.dumps(verdict))
    return verdict


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--arm", default="dropin", choices=["reference", "dropin", "dropin-attn"])
    ap.add_argument("--size", default="0.6b")
    ap.add_argument("--layers", type=int, default=None, help="override num_hidden_layers (memory-bounded runs)")
    ap.add_argument("--tp", type=int, default=1)
    ap.add_argument("--num-seqs", type=int, default=256)
    ap.add_argument("--in-len", type=int, default=128)
    ap.add_argument("--out-len", type=int, default=256)
    ap.add_argument("--shared-prefix", type=int, default=0)
    ap.add_argument("--max-model-len", type=int, default=4096)
    ap.add_argument("--max-num-seqs", type=int, default=512)
    ap.add_argument("--max-num-batched-tokens", type=int, default=16384)
    ap.add_argument("--gpu-memory-utilization", type=float, default=0.9)
    ap.add_argument("--chunked-prefill", action="store_true")
    ap.add_argument("--enforce-eager", action="store_true")
    ap.add_argument("--greedy", action="store_true")
    ap.add_argument("--fp32-logits", action="store_true", help="harness (both arms, TP1): LM head in fp32, continuous margins")
    ap.add_argument("--compare-free", nargs=2, help="free-running greedy parity of two --greedy runs (reference, drop-in)")
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "e2e.json"))
    ap.add_argument("--worker", action="store_true")
    ap.add_argument("--compare", nargs=2)
    ap.add_argument("--steps", type=int, default=64)
    ap.add_argument("--tie-margin", type=float, default=0.032, help="2 bf16 ulps at |logit| in [2,4)")
    ap.add_argument("--force-tokens", default=None, help="JSON of the other arm's greedy run to teacher-force")
    ap.add_argument("--repeat", type=int, default=2)
    ap.add_argument("--kernel-profile", default=None, help="with --step-profile: CUPTI kernel table of 4 decode steps (prefix)")
    ap.add_argument("--step-profile", default=None, help="write per-step phase timings here (perturbs tok/s)")
    a = ap.parse_args()
    if a.compare_free:
        v = compare_free(a.compare_free[0], a.compare_free[1], a.steps, a.tie_margin)
        if a.out:
            with open(a.out, "w") as f:
                json.dump(v, f)
        return 1 if v["first_divergences_for_real"] else 0
    if a.compare:
        v = compare(a.compare[0], a.compare[1], a.steps, a.tie_margin)
        if a.out:
            with open(a.out, "w") as f:
                json.dump(v, f)
        return 1 if v["differ_for_real"] else 0
    return worker(a) if a.worker else launcher(a)


if __name__ == "__main__":
    sys.exit(main())
