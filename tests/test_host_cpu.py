"""CPU tests of the host-side logic above the C ABI (no GPU, no CUDA calls)."""
import glob
import importlib
import os
import sys

import numpy as np
import pytest
import torch

from helpers import GOLDEN

REF = "/root/reference"


class _Seq:
    """The slice of nanovllm.engine.sequence.Sequence that prepare_model_input touches."""
    block_size = 256

    def __init__(self, st):
        self.token_ids = st["token_ids"]
        self.num_cached_tokens = st["num_cached_tokens"]
        self.num_new_tokens = st["num_new_tokens"]
        self.block_table = st["block_table"]

    def __len__(self):
        return len(self.token_ids)

    def __getitem__(self, key):
        return self.token_ids[key]


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLDEN, "prepare_*.pt"))))
def test_prepare_model_input_matches_reference(path):
    """Fixtures: outputs of the reference's own ModelRunner.prepare_model_input
    (model_runner.py:400-495) on seeded sequences (tools/make_golden.py)."""
    from b200attn.engine.model_runner import build_model_input
    m = torch.load(path)
    arrays, max_q, max_k = build_model_input([_Seq(s) for s in m["states"]], 256)
    assert max_q == m["max_seqlen_q"] and max_k == m["max_seqlen_k"]
    for name in ("input_ids", "positions", "cu_seqlens_q", "cu_seqlens_k", "slot_mapping", "context_lens",
                 "seq_need_compute_logits", "block_tables"):
        want = m[name]
        got = arrays[name]
        if want is None:
            assert got is None, name
            continue
        assert got is not None, name
        t = torch.from_numpy(np.ascontiguousarray(got))
        assert t.dtype == want.dtype, (name, t.dtype, want.dtype)
        assert t.shape == want.shape and torch.equal(t, want), name


@pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "nanovllm")), reason="reference tree not present")
def test_install_patches_reference_classes_in_place():
    """install() swaps methods on the reference's own classes (no class replacement), keeps the
    attribute names the engine discovers (`k_cache`, `v_cache`) and is idempotent."""
    os.environ.setdefault("TORCHDYNAMO_DISABLE", "1")
    sys.path.insert(0, REF)
    try:
        import b200attn.install as dropin
        att = importlib.import_module("nanovllm.layers.attention")
        before = att.Attention
        dropin.install()
        dropin.install()
        assert dropin.installed()
        assert att.Attention is before
        a = att.Attention(4, 128, 128 ** -0.5, 1)
        assert hasattr(a, "k_cache") and hasattr(a, "v_cache") and a.num_kv_heads == 1
        assert att.Attention.forward.__module__ == "b200attn.install"
        from b200attn import ops
        assert att.store_kvcache is ops.store_kvcache
        mr = importlib.import_module("nanovllm.engine.model_runner")
        assert mr.ModelRunner.prepare_model_input.__module__ == "b200attn.install"
        assert mr.ModelRunner.run.__module__ == "b200attn.install"
        # CPU tensors must be refused, not silently computed (no CPU fallback on the product path)
        from nanovllm.utils.context import reset_context, set_context
        set_context(torch.tensor([0, 2], dtype=torch.int32), torch.tensor([0, 2], dtype=torch.int32), 2, 2,
                    torch.zeros(0, dtype=torch.int32), None, None, None)
        with pytest.raises(RuntimeError):
            a(torch.zeros(2, 4, 128, dtype=torch.bfloat16), torch.zeros(2, 1, 128, dtype=torch.bfloat16),
              torch.zeros(2, 1, 128, dtype=torch.bfloat16))
        reset_context()
    finally:
        sys.path.remove(REF)


def test_run_returns_host_indices_for_the_untouched_postprocess():
    """ours.run keeps ModelRunner.run's pipeline and return shape (model_runner.py:602-638) but hands
    Scheduler.postprocess (scheduler.py:247-250) the indices as a host list: same values, same order,
    len() and zip() behave, and an index is usable as `seqs[i]`."""
    from types import SimpleNamespace
    from b200attn.engine import model_runner as ours
    calls = []

    class FakeRunner:
        rank = 0
        world_size = 1

        def prepare_model_input(self, seqs):
            self._b200_need_logits_host = [0, 2]
            calls.append("prepare")
            return torch.zeros(3, dtype=torch.int64), torch.zeros(3, dtype=torch.int64)

        def prepare_sample(self, seqs):
            calls.append("sample_prep")
            return torch.ones(2)

        def run_model(self, ids, pos):
            calls.append("forward")
            return torch.zeros(2, 8)

        def sampler(self, logits, temps):
            calls.append("sampler")
            return torch.tensor([7, 9])

    ctx = SimpleNamespace(seq_need_compute_logits=torch.tensor([0, 2], dtype=torch.int32))
    reset = []
    toks, need = ours.run(FakeRunner(), ["a", "b", "c"], lambda: ctx, lambda: reset.append(1))
    assert calls == ["prepare", "sample_prep", "forward", "sampler"] and reset == [1]
    assert toks == [7, 9] and need == [0, 2] and isinstance(need, list)
    assert [["a", "b", "c"][i] for i, _ in zip(need, toks)] == ["a", "c"]

    class ForeignPrepare(FakeRunner):                      # someone else's prepare_model_input: fall back to the Context
        def prepare_model_input(self, seqs):
            return torch.zeros(3, dtype=torch.int64), torch.zeros(3, dtype=torch.int64)
    _, need2 = ours.run(ForeignPrepare(), ["a", "b", "c"], lambda: ctx, lambda: None)
    assert need2 is ctx.seq_need_compute_logits


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLDEN, "prepare_*.pt"))))
def test_tp_control_plane_ships_the_packed_arena(path, monkeypatch):
    """under TP rank 0 sends ("run_packed", layout, nbytes, max_q, max_k, bytes) through the
    reference's own mailbox instead of pickled Sequences.  Round-trip the message through pickle (as
    write_shm/read_shm do, model_runner.py:212-264) and check that the worker sets a Context with exactly
    the values the reference's prepare_model_input produces (the golden fixtures)."""
    import pickle
    from b200attn.engine import model_runner as ours
    m = torch.load(path)
    seqs = [_Seq(s) for s in m["states"]]
    sent, ran = [], []

    class Rank0:
        world_size, rank, block_size = 2, 0, 256
        shm = type("Shm", (), {"size": 1 << 20})()

        def write_shm(self, method_name, *args):
            sent.append(pickle.dumps([method_name, *args]))

        def run(self, s):
            ran.append(s)
            return "tokens"

    r0 = Rank0()
    assert ours.call(r0, "run", seqs, _reference_call=None) == "tokens" and ran == [seqs]
    assert r0._b200_prebuilt[0] is seqs                       # rank 0 reuses the arena it just packed
    method, *args = pickle.loads(sent[0])
    assert method == "run_packed" and len(sent[0]) < 1 << 20

    got = {}

    def fake_upload(self, layout, nbytes, host, max_q, max_k, _set_context):
        raw = host.numpy()[:nbytes]
        for name, off, nb, ch, shape in layout:
            dt = torch.int64 if ch == "q" else torch.int32
            got[name] = torch.from_numpy(raw[off:off + nb].copy()).view(dt).view(shape) if nb else torch.zeros(shape, dtype=dt)
        got["max_q"], got["max_k"] = max_q, max_k
        return got["input_ids"], got["positions"]
    monkeypatch.setattr(ours, "_upload_and_set_context", fake_upload)

    class Worker:
        world_size, rank = 2, 1
        forwards = []

        def run_model(self, ids, pos):
            self.forwards.append((ids, pos))

    w = Worker()
    assert ours.run_packed(w, *args, _set_context=None, _reset_context=lambda: None) is None
    assert len(w.forwards) == 1
    assert got["max_q"] == m["max_seqlen_q"] and got["max_k"] == m["max_seqlen_k"]
    for name in ("input_ids", "positions", "cu_seqlens_q", "cu_seqlens_k", "slot_mapping", "context_lens",
                 "seq_need_compute_logits", "block_tables"):
        if m[name] is None:
            assert name not in got, name
        else:
            assert got[name].dtype == m[name].dtype and torch.equal(got[name], m[name]), name

    # a step too large for the reference's 1 MiB mailbox travels through a side shared-memory segment: only the
    # segment's name and the layout go through the mailbox, and the worker ends up with the same Context
    r0.shm.size = 64
    sent.clear(); ran.clear(); got.clear()
    try:
        assert ours.call(r0, "run", seqs, _reference_call=None) == "tokens" and ran == [seqs]
        method, *args2 = pickle.loads(sent[0])
        assert method == "run_packed" and isinstance(args2[-1], tuple) and len(sent[0]) < 4096
        w2 = Worker()
        w2.forwards = []
        assert ours.run_packed(w2, *args2, _set_context=None, _reset_context=lambda: None) is None
        assert len(w2.forwards) == 1 and got["max_q"] == m["max_seqlen_q"]
        for name in ("input_ids", "positions", "cu_seqlens_q", "cu_seqlens_k", "slot_mapping", "context_lens"):
            assert torch.equal(got[name], m[name]), name
        ours.release_side_arena(w2)
    finally:
        ours.release_side_arena(r0)
    # other methods are forwarded untouched; "exit" tells the workers first, tears the peer-memory
    # communicators down (none here), runs the reference's exit and releases the side segment
    fell = []
    sent.clear()
    r0.exit = lambda: fell.append("exit-method")
    ours.call(r0, "exit", _reference_call=lambda self, name, *a: fell.append(name))
    ours.call(r0, "loop", _reference_call=lambda self, name, *a: fell.append(name))
    assert fell == ["exit-method", "loop"] and pickle.loads(sent[0]) == ["exit"]


def test_allreduce_form_policy_and_no_cpu_fallback():
    """Host logic of the f3 all-reduce (b200attn/ops.py): the form is a pure function of sizes (every rank
    must pick the same), follows the measured policy, and the entry point refuses CPU tensors."""
    import torch
    from b200attn import ops
    MB = 1 << 20
    pick = lambda world, nbytes: ops.ar_choose_form(world, nbytes, 4 * MB, ops._ar_two_shot_min_bytes(world))  # noqa: E731
    assert pick(2, 8192) == 3 and pick(2, 4 * MB) == 3                    # TP2: one-shot push always fits
    assert pick(4, 2 * MB) == 3 and pick(4, 4 * MB) == 2                  # TP4: 4 x 4 MB no longer fits a slot
    assert pick(8, 512 << 10) == 3 and pick(8, MB) == 2 and pick(8, 2 * MB) == 2
    assert ops.ar_choose_form(2, 8 * MB, 4 * MB, 16 * MB) == 0            # a 2-rank message too big to push: pull
    with pytest.raises(RuntimeError):
        ops.tp_all_reduce_(torch.zeros(8, dtype=torch.bfloat16))


def test_bench_reference_arm_prints_the_contract_line():
    """`bench.py --impl reference` (the CPU oracle port timed on host cores; the one place besides tests/ and smoke()
    that may execute oracle/) prints ONE JSON line with the keys the driver reads, and loads no native library."""
    import json
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "1"],
                       capture_output=True, text=True, timeout=600, cwd=root)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "paged_attn_decode_hbm_gbps" and d["unit"] == "GB/s"
    assert d["higher_is_better"] is True and d["n_gpus"] == 1 and d["steps"] == 1 and d["value"] > 0
    assert d["config"]["workload"].startswith("qwen3-8b-decode-paged-attn-b512-ctx4096")
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == d["value"] and "sample" in cb
    assert d["e2e"] == {"value": d["value"], "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["gpu_launches"] == 0


def test_nvls_form_policy_is_a_pure_function_of_group_and_bytes():
    """NvlsAllReduce.choose_form / pays / form_fits (ops.py) decide on (world, message bytes) only, so every rank picks
    the same form without talking: flag-in-data forms for small messages, in-switch reduction beyond ~1 MB, and at TP2
    the fused launch only pays up to 1 MB (profiles/r2c_bench_n2.jsonl: peer-store + norm kernel 19.0 us vs fused 21.3 us
    at 2 MB; at world 8 fused two-shot 27.3 us vs 34.8 us)."""
    from b200attn import ops
    MB = 1 << 20

    def comm(world):
        c = object.__new__(ops.NvlsAllReduce)
        c.world, c.max_bytes, c.ll_max_bytes = world, 8 * MB, 4 * MB
        c.ll1_budget, c.ll2_max, c.one_shot_budget, c.fused_max_pair, c.forced_form = MB, MB, 3 << 19, MB, None
        return c

    for world in (2, 4, 8):
        c = comm(world)
        forms = [c.choose_form(n) for n in (4096, 64 * 1024, 512 * 1024, 2 * MB, 4 * MB)]
        assert forms == [comm(world).choose_form(n) for n in (4096, 64 * 1024, 512 * 1024, 2 * MB, 4 * MB)]
        assert forms[0] == 2                                        # tiny: LL one-shot, no barrier
        assert forms[2] == 3                                        # <= 1 MB: LL two-shot
        assert forms[3] in (0, 1) and forms[4] in (0, 1)            # beyond: the switch reduces
        for n, f in zip((4096, 64 * 1024, 512 * 1024, 2 * MB, 4 * MB), forms):
            assert c.form_fits(f, n)
    assert comm(8).choose_form(2 * MB) == 0 and comm(2).choose_form(2 * MB) == 1   # world 8: reduce-scatter + all-gather
    x2 = type("T", (), {})()
    for world, nbytes, want in ((2, 512 * 1024, True), (2, 2 * MB, False), (4, 2 * MB, True), (8, 4 * MB, True)):
        c = comm(world)
        c.fits = lambda x: True
        x2.numel = lambda n=nbytes: n // 2
        assert c.pays(x2) is want


def test_pending_all_reduce_is_settled_before_any_copy_of_its_tensor(monkeypatch):
    """RMSNorm.forward may be handed a NON-contiguous view of the RowParallelLinear output whose
    all-reduce was deferred.  The reduce must hit the original BEFORE the view is copied, or the norm would run on
    un-reduced partial sums (reference order: dist.all_reduce at layers/linear.py:407-410, then layernorm.py:95-141).
    Host logic only: the collective and the norm kernel are replaced by recording stand-ins."""
    import torch
    from b200attn import install, ops
    log = []

    def fake_all_reduce(t):
        log.append("reduce")
        t.mul_(2)                                   # "sum over 2 ranks" of identical partials
        return t

    def fake_rms_norm(x, weight, eps, residual=None):
        assert x.is_contiguous() and (residual is None or residual.is_contiguous())
        log.append("norm")
        return x.clone(), residual

    monkeypatch.setattr(ops, "tp_all_reduce_", fake_all_reduce)
    monkeypatch.setattr(ops, "rms_norm", fake_rms_norm)
    y = torch.ones(4, 16)                           # un-reduced partial sums, pending
    monkeypatch.setattr(ops, "_PENDING", y)
    view = y[:, ::2]                                # what a model that slices between o_proj and the norm would pass
    assert not view.is_contiguous()

    class Norm:
        weight, eps = torch.ones(8), 1e-6

    out, _ = install._rmsnorm_forward(Norm(), view, torch.zeros(4, 8))
    assert log == ["reduce", "norm"]
    assert ops._PENDING is None
    assert torch.equal(out, torch.full((4, 8), 2.0))          # the norm saw the REDUCED values
    # same for the residual-free form
    log.clear()
    y2 = torch.ones(4, 16)
    monkeypatch.setattr(ops, "_PENDING", y2)
    out2 = install._rmsnorm_forward(Norm(), y2[:, ::2])
    assert log == ["reduce", "norm"] and torch.equal(out2[0], torch.full((4, 8), 2.0))


def test_a_timed_out_collective_is_an_exception_not_wrong_tokens(monkeypatch):
    """after a bounded wait gives up, the status word is sticky and all later all-reduces are garbage.
    The engine polls it where it has just synchronised (rank 0, after sampling) and at shutdown (every rank)."""
    from b200attn import ops

    class Comm:
        def __init__(self, st):
            self.st, self.destroyed = st, False

        def status(self):
            return self.st

        def destroy(self):
            self.destroyed = True

    monkeypatch.setattr(ops, "_TP_AR", Comm(0))
    monkeypatch.setattr(ops, "_TP_NVLS", False)          # NVLS unavailable on this group
    assert ops.collectives_status() == 0
    ops.raise_if_collectives_failed()
    bad = Comm(1)
    monkeypatch.setattr(ops, "_TP_NVLS", bad)
    assert ops.collectives_status() == 1
    with pytest.raises(RuntimeError, match="timed out"):
        ops.raise_if_collectives_failed()
    import torch
    monkeypatch.setattr(torch.cuda, "synchronize", lambda *a, **k: None)
    with pytest.raises(RuntimeError, match="timed out"):
        ops.shutdown_collectives()
    assert bad.destroyed and ops._TP_AR is None and ops._TP_NVLS is None     # torn down before the error is raised
