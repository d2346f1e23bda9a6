"""E2E HARNESS ONLY (not product code).  tools/e2e_generate.py puts this directory first on
PYTHONPATH so that every process of a run -- including the reference's spawned TP workers --
gets the same harness patches, applied identically to BOTH arms (reference and drop-in):

  B200E2E_RANDOM_INIT=<seed>  replace `load_model` (no checkpoints offline) by a seeded
                              N(0, 0.02) init of every parameter, norm weights = 1
  B200E2E_GREEDY=1            replace Sampler.forward by argmax and record (step, argmax, top-2 logit
                              margin) of every decision per seq_id (the reference's sampler forbids
                              greedy, F12)
  B200E2E_FP32_LOGITS=1       TP1 only: the LM head multiplies in fp32 (same last-token selection as
                              embed_head.py:194-204), so top-2 margins are continuous instead of quantised to
                              bf16 ulps of the logits (one ulp at |logit| in [2,4) is 0.0156: with bf16 logits more
                              than 1 % of all decisions are EXACT ties in the reference arm itself)
  B200E2E_FORCE_TOKENS=<json> with GREEDY: still record this arm's own argmax, but FEED the token ids
                              of the other arm's run (teacher forcing), so every decision of the two
                              arms is taken on an identical context
  B200E2E_SHM_BYTES=<n>       size of the rank-0 -> worker shared-memory mailbox (F11)
  B200E2E_STEP_PROFILE=<json> rank 0: wall time of every phase of every engine step (schedule,
                              prepare_model_input, forward with a device sync either side, sampler +
                              .tolist(), postprocess) -- perturbs throughput, so never combined with a
                              tok/s measurement

and then chains to the product hook nano-vllm-v1_b200/dropin/sitecustomize.py (B200ATTN_DROPIN)."""
import atexit
import importlib.abc
import importlib.machinery
import json
import os
import runpy
import sys

_SEED = os.environ.get("B200E2E_RANDOM_INIT")
_GREEDY = os.environ.get("B200E2E_GREEDY") == "1"
_TRIGGER = "nanovllm.engine.model_runner"
_MARGINS = []
_PER_SEQ = {}


def _patch(module):
    import torch
    if _SEED is not None:
        def seeded_init(model, path):
            import torch.distributed as dist
            rank = dist.get_rank() if dist.is_initialized() else 0
            # parameters are per-rank shards: every rank draws from its OWN stream (seed + 1000 * rank), the same
            # in both arms.  One shared stream would make the shards of a vocab- or head-parallel weight identical
            # copies, i.e. duplicate logits and exact top-2 ties (a degenerate model).
            g = torch.Generator(device="cuda").manual_seed(int(_SEED) + 1000 * rank)
            with torch.no_grad():
                for name, p in sorted(model.named_parameters(), key=lambda kv: kv[0]):
                    if name.endswith("norm.weight"):
                        p.fill_(1.0)
                    else:
                        p.copy_(torch.randn(p.shape, generator=g, device="cuda", dtype=torch.float32).mul_(0.02).to(p.dtype))
        module.load_model = seeded_init
    if os.environ.get("B200E2E_FP32_LOGITS") == "1":
        head = importlib.import_module("nanovllm.layers.embed_head")
        ctx_mod0 = importlib.import_module("nanovllm.utils.context")

        def fp32_head(self, x):
            assert self.tp_size == 1, "B200E2E_FP32_LOGITS is a TP1 harness option"
            context = ctx_mod0.get_context()
            last = context.cu_seqlens_q[1:] - 1                     # embed_head.py:194-200
            if context.seq_need_compute_logits.numel():
                last = last[context.seq_need_compute_logits]
            return torch.nn.functional.linear(x[last].float(), self.weight.float())
        head.ParallelLMHead.forward = fp32_head
    if _GREEDY:
        sampler = importlib.import_module("nanovllm.layers.sampler")
        ctx_mod = importlib.import_module("nanovllm.utils.context")
        forced = None
        if os.environ.get("B200E2E_FORCE_TOKENS"):                    # teacher forcing: feed the other arm's tokens
            with open(os.environ["B200E2E_FORCE_TOKENS"]) as f:
                forced = json.load(f)["token_ids"]
        cur = {}

        def greedy(self, logits, temperatures):
            top2 = torch.topk(logits.float(), 2, dim=-1)
            choice = top2.indices[:, 0]
            margins = (top2.values[:, 0] - top2.values[:, 1]).tolist()
            tops = top2.values[:, 0].tolist()
            picks = choice.tolist()
            seqs = cur.get("seqs")
            if seqs is not None:
                need = ctx_mod.get_context().seq_need_compute_logits
                rows = need.tolist() if need is not None and need.numel() else list(range(len(seqs)))
                out = list(picks)
                for row, si in enumerate(rows):
                    sq = seqs[si]
                    sid = int(getattr(sq, "seq_id", -1))
                    step = len(sq) - sq.num_prompt_tokens
                    _PER_SEQ.setdefault(sid, []).append((step, picks[row], margins[row], tops[row]))
                    k = sid - int(os.environ.get("B200E2E_FIRST_SEQ_ID", "1"))   # earlier ids: warm-up sequences
                    if forced is not None and 0 <= k < len(forced) and step < len(forced[k]):
                        out[row] = forced[k][step]
                if forced is not None:
                    return torch.tensor(out, dtype=choice.dtype, device=choice.device)
            return choice
        sampler.Sampler.forward = greedy
        orig_run = module.ModelRunner.run

        def run(self, seqs):
            cur["seqs"] = seqs if self.rank == 0 else None
            try:
                return orig_run(self, seqs)
            finally:
                cur["seqs"] = None
        module.ModelRunner.run = run
        out = os.environ.get("B200E2E_MARGINS_OUT")
        if out:
            def dump():
                if _PER_SEQ:
                    with open(out, "w") as f:
                        json.dump({str(k): v for k, v in _PER_SEQ.items()}, f)
            atexit.register(dump)
    prof_out = os.environ.get("B200E2E_STEP_PROFILE")
    if prof_out:
        _install_step_profile(module, prof_out)
    shm = os.environ.get("B200E2E_SHM_BYTES")
    if shm:
        real = module.SharedMemory

        def bigger(name=None, create=False, size=0):
            return real(name=name, create=create, size=int(shm) if create else size)
        module.SharedMemory = bigger


def _install_step_profile(module, out_path):
    import time
    import torch
    sched = importlib.import_module("nanovllm.engine.scheduler")
    steps, cur = [], {}

    def timed(cls, name, key, sync=False):
        orig = getattr(cls, name)

        def wrapper(self, *a, **kw):
            if sync:
                torch.cuda.synchronize()
            t0 = time.perf_counter()
            try:
                return orig(self, *a, **kw)
            finally:
                if sync:
                    torch.cuda.synchronize()
                cur[key] = cur.get(key, 0.0) + time.perf_counter() - t0
        setattr(cls, name, wrapper)

    kern_out = os.environ.get("B200E2E_KERNEL_PROFILE")
    if kern_out:
        # CUPTI kernel table of a few mid-run decode steps (graph replays): which kernels the step is made of
        orig_rm = module.ModelRunner.run_model
        calls = {"n": 0}
        first = int(os.environ.get("B200E2E_KERNEL_PROFILE_STEP", "120"))

        def run_model(self, input_ids, positions):
            calls["n"] += 1
            if self.rank != 0 or not (first <= calls["n"] < first + 4):
                return orig_rm(self, input_ids, positions)
            from torch.profiler import ProfilerActivity, profile
            torch.cuda.synchronize()
            with profile(activities=[ProfilerActivity.CUDA]) as prof:
                out = orig_rm(self, input_ids, positions)
                torch.cuda.synchronize()
            rows = [dict(name=e.key, calls=e.count, cuda_us=e.device_time_total) for e in prof.key_averages()]
            rows.sort(key=lambda r: -r["cuda_us"])
            with open(f"{kern_out}.{calls['n']}.json", "w") as f:
                json.dump(dict(tokens=int(input_ids.numel()), rows=rows), f)
            return out
        module.ModelRunner.run_model = run_model
    timed(module.ModelRunner, "prepare_model_input", "prepare")
    timed(module.ModelRunner, "run_model", "forward_synced", sync=True)
    timed(module.ModelRunner, "run", "run_total")
    timed(sched.Scheduler, "schedule", "schedule")
    orig_post = sched.Scheduler.postprocess

    def post(self, seqs, token_ids, need):
        t0 = time.perf_counter()
        try:
            return orig_post(self, seqs, token_ids, need)
        finally:
            cur["postprocess"] = time.perf_counter() - t0
            cur["num_seqs"] = len(seqs)
            cur["t_end"] = time.perf_counter()
            steps.append(dict(cur))
            cur.clear()
    sched.Scheduler.postprocess = post

    def dump():
        if steps:
            with open(out_path, "w") as f:
                json.dump(steps, f)
    atexit.register(dump)


class _Loader(importlib.abc.Loader):
    def __init__(self, inner):
        self.inner = inner

    def create_module(self, spec):
        return self.inner.create_module(spec)

    def exec_module(self, module):
        self.inner.exec_module(module)
        _patch(module)


class _Finder(importlib.abc.MetaPathFinder):
    def find_spec(self, name, path, target=None):
        if name != _TRIGGER:
            return None
        sys.meta_path.remove(self)
        spec = importlib.util.find_spec(name)        # lets the product hook (if any) wrap first
        if spec is not None and spec.loader is not None:
            spec.loader = _Loader(spec.loader)
        return spec


import importlib.util  # noqa: E402

_product = os.path.join(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))),
                        "nano-vllm-v1_b200", "dropin", "sitecustomize.py")
if os.environ.get("B200ATTN_DROPIN") and os.path.exists(_product):
    runpy.run_path(_product)                         # product hook first ...
if _SEED is not None or _GREEDY or os.environ.get("B200E2E_SHM_BYTES") or os.environ.get("B200E2E_STEP_PROFILE"):
    sys.meta_path.insert(0, _Finder())               # ... so the harness finder sits in front and wraps it
