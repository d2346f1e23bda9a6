#!/usr/bin/env python
"""Online-serving benchmark of BASELINE.json configs[3] through the reference's UNTOUCHED engine API
(`LLM.add_request` / `LLM.step` / `Scheduler`), for both arms on the same box.

It restates the measurement protocol of the reference's own harness (/root/reference/serving_bench.py:61-167, which
is not part of the installed package and therefore not on the GPU box): requests arrive as a Poisson process
(`np.random.exponential(1/rate)` inter-arrival times, seeds 100 as serving_bench.py:13-14), each one is handed to
`add_request` when its arrival time has passed, the engine is stepped in between; TTFT = first step after which the
sequence's prompt is fully cached (serving_bench.py:127-131), TPOT = (completion - first token) / (output_len - 1),
throughput = (input + output tokens) / wall time.  Arms as tools/e2e_generate.py: `reference` = the unmodified
offline install baseline/_ref, `dropin` = the same install with b200attn.install() applied by the import hook;
both with seeded random-init weights of the named Qwen3 size (there are no checkpoints here).

    python tools/serving_bench.py --arm dropin --size 14b --tp 2 --request-rate 10 --num-requests 1024 \
        --max-num-batched-tokens 1024 --chunked-prefill --out gpurun_out/serving_dropin_1024.json
"""
import argparse
import json
import os
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
    cmd = [sys.executable, os.path.abspath(__file__), "--worker"] + [x for x in sys.argv[1:] if x != "--worker"]
    return subprocess.run(cmd, env=env).returncode


def worker(a):
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    from make_fake_model import make
    import numpy as np
    from random import randint, seed
    seed(100)                                                              # serving_bench.py:13-14
    np.random.seed(100)
    model_dir = make(a.size, f"/tmp/b200_fake_qwen3_{a.size}_{a.layers or 'full'}", a.layers)
    import torch
    from nanovllm import LLM, SamplingParams
    t0 = time.perf_counter()
    llm = LLM(model_dir, enforce_eager=a.enforce_eager, max_model_len=a.max_model_len,
              max_num_batched_tokens=a.max_num_batched_tokens, max_num_seqs=a.max_num_seqs,
              tensor_parallel_size=a.tp, chunked_prefill=a.chunked_prefill)
    t_init = time.perf_counter() - t0
    sp = lambda: SamplingParams(temperature=0.6, ignore_eos=True, max_tokens=a.out_len)  # noqa: E731
    warm = [[randint(0, 10000) for _ in range(a.in_len)] for _ in range(50)]             # serving_bench.py:52-57
    llm.generate(warm, [sp() for _ in warm], use_tqdm=False)
    torch.cuda.synchronize()

    prompts = [[randint(0, 10000) for _ in range(a.in_len)] for _ in range(a.num_requests)]
    arrivals = np.cumsum(np.random.exponential(1.0 / a.request_rate, a.num_requests))
    sub, first, done, out_len = {}, {}, {}, {}
    sent, steps = 0, 0
    start = time.perf_counter()
    while sent < a.num_requests or not llm.is_finished():
        now = time.perf_counter()
        while sent < a.num_requests and now - start >= arrivals[sent]:
            llm.add_request(prompts[sent], sp())
            sid = llm.scheduler.waiting[-1].seq_id
            sub[sid] = start + arrivals[sent]
            sent += 1
        if not llm.is_finished():
            finished, _ = llm.step()
            steps += 1
            t = time.perf_counter()
            for seq in list(llm.scheduler.running):
                if seq.seq_id in sub and seq.seq_id not in first and seq.num_cached_tokens == seq.num_prompt_tokens:
                    first[seq.seq_id] = t
            for sid, ids in finished:
                if sid in sub:
                    first.setdefault(sid, t)
                    done[sid] = t
                    out_len[sid] = len(ids)
        else:
            time.sleep(0.01)
    total = time.perf_counter() - start
    ttft = [first[s] - sub[s] for s in first]
    tpot = [(done[s] - first[s]) / (out_len[s] - 1) for s in done if out_len[s] > 1]
    lat = [done[s] - sub[s] for s in done]
    tokens = a.in_len * len(sub) + sum(out_len.values())
    import b200attn.install as dropin_mod
    rec = dict(arm=a.arm, size=a.size, layers=a.layers, tp=a.tp, request_rate=a.request_rate, num_requests=a.num_requests,
               in_len=a.in_len, out_len=a.out_len, max_num_batched_tokens=a.max_num_batched_tokens, max_num_seqs=a.max_num_seqs,
               chunked_prefill=a.chunked_prefill, enforce_eager=a.enforce_eager, dropin_installed=dropin_mod.installed(),
               total_seconds=total, engine_steps=steps, completed=len(done), throughput_tok_per_s=tokens / total,
               ttft_ms_mean=1e3 * float(np.mean(ttft)), ttft_ms_p50=1e3 * float(np.percentile(ttft, 50)),
               ttft_ms_p99=1e3 * float(np.percentile(ttft, 99)),
               tpot_ms_mean=1e3 * float(np.mean(tpot)), tpot_ms_p50=1e3 * float(np.percentile(tpot, 50)),
               tpot_ms_p99=1e3 * float(np.percentile(tpot, 99)),
               latency_s_mean=float(np.mean(lat)), init_seconds=t_init, gpu=torch.cuda.get_device_name(0),
               weights="seeded random init (tools/e2e/sitecustomize.py)", protocol="/root/reference/serving_bench.py:61-167 restated")
    os.makedirs(os.path.dirname(os.path.abspath(a.out)), exist_ok=True)
    with open(a.out, "w") as f:
        json.dump(rec, f)
    print(json.dumps(rec), flush=True)
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--arm", default="dropin", choices=["reference", "dropin"])
    ap.add_argument("--size", default="14b")
    ap.add_argument("--layers", type=int, default=None)
    ap.add_argument("--tp", type=int, default=2)
    ap.add_argument("--request-rate", type=float, default=10.0)
    ap.add_argument("--num-requests", type=int, default=1024)
    ap.add_argument("--in-len", type=int, default=128)
    ap.add_argument("--out-len", type=int, default=100)
    ap.add_argument("--max-model-len", type=int, default=40960)
    ap.add_argument("--max-num-seqs", type=int, default=1024)
    ap.add_argument("--max-num-batched-tokens", type=int, default=1024)
    ap.add_argument("--chunked-prefill", action="store_true")
    ap.add_argument("--enforce-eager", action="store_true")
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "serving.json"))
    ap.add_argument("--worker", action="store_true")
    a = ap.parse_args()
    sys.exit(worker(a) if a.worker else launcher(a))


if __name__ == "__main__":
    main()
