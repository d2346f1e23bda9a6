"""Engine-level parity on a B200 (SURVEY rows a11, a13 and the north_star greedy rule): the reference's UNTOUCHED
`LLM.generate` -- v1 scheduler, BlockManager, chunked prefill, `allocate_kv_cache`, `capture_cudagraph`,
CUDA-graph replay of decode -- is run in subprocesses on the same seeded random-init synthetic Qwen3 (4 layers of the
0.6B shape, 32 prompts x 128 tokens, 64 decode steps, max_num_batched_tokens 1000 so prompts are chunked and
steps mix prefill chunks with decode rows):

  reference arm : the unmodified offline install `baseline/_ref` (Triton store_kvcache + flash-attn 2.8.3 + inductor)
  drop-in arm   : the same install with `b200attn.install.install()` applied by the import hook (every sm_100a kernel
                  of this repo on the path, ModelRunner hooks included), once free-running and once teacher-forced
                  (fed the reference arm's tokens, so all 2048 decisions of both arms are taken on identical contexts)

Both arms get the same harness patches (tools/e2e/sitecustomize.py): seeded N(0, 0.02) weights (per-rank streams), an
arg-max sampler that records the top-1 logit and the top-2 margin (the reference's sampler forbids greedy)
and an fp32 LM head, so margins are continuous instead of quantised to bf16 ulps of the logits.

What can and cannot be asserted.  The logits of this model are ~N(0, 0.64^2) over 151,936 tokens; the gap between the
two largest falls below g with probability ~ 4.9 g / sigma, so among 2048 decisions the smallest margins are ~1e-5 and
the 1 % quantile is ~3e-3 -- whatever the seed, and ~12 % of all margins are below 0.03.  Two correct bf16
implementations do not produce the same bits: measured on a B200 (profiles/r2_engine_parity_calibration.json) the
top-1 logit of the two arms differs by 0.0037 (median) to 0.0195 (max of 2048) on IDENTICAL contexts -- 0.6 % to 3 %
of sigma, the size of one bf16 ulp of the final hidden state -- with 2 layers as with 4.  A decision whose margin is
inside that noise is a coin flip for ANY pair of implementations (flash-attn 2 vs 3 included), so "identical over 64
steps" is asserted outside a tie class, and the tie class is kept honest by four guards instead of by the
unattainable "1 % margin quantile above the noise":

  * LOGIT parity, all decisions, teacher-forced: |top-1 logit difference| <= LOGIT_TOL = 0.03 (an attention bug --
    wrong page, mask, scale, stale cache row -- moves logits by O(sigma) = 0.6);
  * TOKEN parity, all decisions, teacher-forced: identical wherever either arm's margin exceeds TIE_MARGIN = 0.03;
    at most 3 % of the decisions may differ at all;
  * the model is not degenerate: no exact ties (p0 > 0), median margin > 3 x TIE_MARGIN, tie class <= 20 %;
  * free-running: a sequence leaves the reference trajectory only at a decision inside the tie class, and at least
    half of the sequences never leave it.
"""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "baseline", "_ref", "nanovllm")
TOOL = os.path.join(ROOT, "tools", "e2e_generate.py")

LOGIT_TOL = 0.03          # max |top-1 logit(reference) - top-1 logit(drop-in)| on identical contexts (measured max 0.0195)
TIE_MARGIN = 0.03         # a decision with a top-2 margin below this in BOTH arms may legitimately flip (measured max 0.0157)
STEPS = 64
SEQS = 32


def _run_arm(arm, out, extra=(), mnbt=1000):
    cmd = [sys.executable, TOOL, "--arm", arm, "--size", "0.6b", "--layers", "4", "--num-seqs", str(SEQS), "--in-len", "128",
           "--out-len", str(STEPS), "--chunked-prefill", "--max-model-len", "1024", "--max-num-batched-tokens", str(mnbt),
           "--gpu-memory-utilization", "0.3", "--greedy", "--fp32-logits", "--out", out, *extra]
    env = dict(os.environ)
    env.pop("B200ATTN_DROPIN", None)
    r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=1500)
    assert r.returncode == 0, f"{arm} arm failed:\n{r.stdout[-3000:]}\n{r.stderr[-6000:]}"
    with open(out) as f:
        return json.load(f)


@pytest.mark.gpu
def test_llm_generate_greedy_tokens_identical_to_reference_arm(tmp_path):
    if not os.path.isdir(REF):
        pytest.skip("baseline/_ref (offline install of the reference, __graft_entry__.build()) is not in this tree")
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import e2e_generate

    ref_p, our_p, tf_p = (str(tmp_path / n) for n in ("ref.json", "our.json", "tf.json"))
    ref = _run_arm("reference", ref_p)
    our = _run_arm("dropin", our_p)
    forced = _run_arm("dropin", tf_p, ("--force-tokens", ref_p))
    assert ref["dropin_installed"] is False and our["dropin_installed"] is True and forced["dropin_installed"] is True
    assert not ref["enforce_eager"] and not our["enforce_eager"] and our["chunked_prefill"]      # graphs on, chunked
    for rec in (ref, our, forced):
        assert len(rec["token_ids"]) == SEQS and all(len(t) == STEPS for t in rec["token_ids"])
    assert forced["token_ids"] == ref["token_ids"]                 # teacher forcing fed the reference's tokens

    tf = e2e_generate.compare(ref_p, tf_p, STEPS, TIE_MARGIN)
    free = e2e_generate.compare_free(ref_p, our_p, STEPS, TIE_MARGIN)
    with open(ref_p + ".margins") as f:
        margins = sorted(rec[2] for recs in json.load(f).values() for rec in recs if rec[0] < STEPS)
    tie_class = sum(1 for m in margins if m <= TIE_MARGIN) / len(margins)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "engine_greedy_parity.json"), "w") as f:
        json.dump(dict(teacher_forced=tf, free_running=free, tie_class_fraction=tie_class, logit_tol=LOGIT_TOL,
                       ref_tok_per_s=ref["tok_per_s"], our_tok_per_s=our["tok_per_s"]), f)

    # (1) logits agree on every decision taken on an identical context
    noise = tf["top1_logit_abs_diff"]
    assert tf["decisions"] == SEQS * STEPS and noise["max"] is not None and noise["max"] <= LOGIT_TOL, tf
    # (2) tokens identical outside the tie class, and almost everywhere
    assert tf["differ_for_real"] == 0, tf
    assert tf["identical"] >= 0.97 * tf["decisions"], tf
    # (3) the model is not degenerate and the tie class is small
    q = tf["ref_margin_quantiles"]
    assert q["p0"] > 0.0 and q["p50"] > 3 * TIE_MARGIN and tie_class <= 0.20, (q, tie_class)
    # (4) free-running generate: trajectories identical up to a tie, most of them to the end
    assert free["first_divergences_for_real"] == 0, free
    assert free["sequences_fully_identical"] >= 0.5 * SEQS, free
    assert free["decisions_compared"] >= 0.6 * free["decisions_total"], free


@pytest.mark.gpu
def test_reference_vs_reference_noise_floor_control(tmp_path):
    """VERDICT r2 (missing #5): the tie class above was calibrated from drop-in-vs-reference differences only.  This is
    the control: the UNMODIFIED reference against ITSELF on identical contexts (teacher-forced with its own tokens) under
    two perturbations every deployment is free to make -- another max_num_batched_tokens (other chunk boundaries and
    batch compositions: other cuBLAS / flash-attn tilings and split counts) and enforce_eager (no CUDA graphs:
    flash_attn_varlen_func instead of flash_attn_with_kvcache for decode, attention.py:229-248).  No kernel of this repo
    runs here.  The reference's own top-1 logit noise and own token flips are recorded next to the drop-in's
    (gpurun_out/engine_ref_vs_ref.json -> profiles/); asserted: the control really ran on the same contexts, and the
    reference's own flips stay inside the same tie class the drop-in is held to (if the reference itself flipped outside
    it, TIE_MARGIN would be a property of this repo's kernels rather than of bf16 inference, and too generous or too tight)."""
    if not os.path.isdir(REF):
        pytest.skip("baseline/_ref (offline install of the reference, __graft_entry__.build()) is not in this tree")
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import e2e_generate

    base_p, mnbt_p, eager_p = (str(tmp_path / n) for n in ("base.json", "mnbt.json", "eager.json"))
    base = _run_arm("reference", base_p)
    runs = {"max_num_batched_tokens_4096": _run_arm("reference", mnbt_p, ("--force-tokens", base_p), mnbt=4096),
            "enforce_eager": _run_arm("reference", eager_p, ("--force-tokens", base_p, "--enforce-eager"))}
    record = {"tie_margin": TIE_MARGIN, "logit_tol": LOGIT_TOL}
    for (name, rec), path in zip(runs.items(), (mnbt_p, eager_p)):
        assert rec["dropin_installed"] is False and base["dropin_installed"] is False
        assert rec["token_ids"] == base["token_ids"]               # fed the baseline run's tokens: identical contexts
        v = e2e_generate.compare(base_p, path, STEPS, TIE_MARGIN)
        record[name] = v
        assert v["decisions"] == SEQS * STEPS, v
        assert v["differ_for_real"] == 0, v                         # the reference's own flips are ties too
    assert runs["enforce_eager"]["enforce_eager"] and not base["enforce_eager"]
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "engine_ref_vs_ref.json"), "w") as f:
        json.dump(record, f)
