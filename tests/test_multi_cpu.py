"""World-size-2 `gloo` tests (CPU, `-m "not gpu"`) of the N > 1 host logic:

* head sharding: rank r owns q-heads [r*Hq/tp, ...) and kv-heads [r*Hkv/tp, ...) and its own KV-cache shard
  (linear.py:297-332, model_runner.py:360-368); metadata is identical on all ranks; the per-rank results
  concatenated over heads are bit-identical to the unsharded result, with no exchange on the attention path;
* RowParallelLinear drop-in (b200attn/layers/linear.py): partial sums + bias on rank 0 only + one in-place
  all-reduce equal the unsharded linear; the CUDA carrier is replaced by gloo here, the policy that picks
  the carrier's form must agree on all ranks.
"""
import os
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
SCALE = 128 ** -0.5


def _worker(rank, world, port, tmp):
    for p in (ROOT, os.path.join(ROOT, "nano-vllm-v1_b200"), HERE):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import oracle
        from helpers import make_paged_case
        from b200attn import ops
        from b200attn.layers import linear as ours_linear

        # ---- head sharding of a mixed paged batch (prefill chunk rows + decode rows) ----------------
        Hq, Hkv = 8, 2
        lens_q, lens_k = [70, 1, 1, 300], [326, 513, 5, 300]
        c = make_paged_case(11, lens_q, lens_k, Hq, Hkv)             # same seed on every rank => same metadata
        full = oracle.attn_varlen(c["q"], c["k_cache"], c["v_cache"], c["cu_seqlens_q"], c["cu_seqlens_k"], SCALE,
                                  c["block_tables"])
        hq, hkv = Hq // world, Hkv // world
        q_r = c["q"][:, rank * hq:(rank + 1) * hq].contiguous()
        kc_r = c["k_cache"][:, :, rank * hkv:(rank + 1) * hkv].contiguous()     # this rank's cache shard
        vc_r = c["v_cache"][:, :, rank * hkv:(rank + 1) * hkv].contiguous()
        mine = oracle.attn_varlen(q_r, kc_r, vc_r, c["cu_seqlens_q"], c["cu_seqlens_k"], SCALE, c["block_tables"])
        parts = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(parts, mine)
        assert torch.equal(torch.cat(parts, dim=1), full), "head-sharded attention differs from the unsharded result"

        # decode form (graph metadata): context_lens + full-width block table
        cd = make_paged_case(12, [1] * 3, [700, 0, 31], Hq, Hkv, pad=0, table_width=4)
        full_d = oracle.attn_decode(cd["q"], cd["k_cache"], cd["v_cache"], cd["context_lens"], cd["block_tables"], SCALE)
        mine_d = oracle.attn_decode(cd["q"][:, rank * hq:(rank + 1) * hq].contiguous(),
                                    cd["k_cache"][:, :, rank * hkv:(rank + 1) * hkv].contiguous(),
                                    cd["v_cache"][:, :, rank * hkv:(rank + 1) * hkv].contiguous(),
                                    cd["context_lens"], cd["block_tables"], SCALE)
        parts = [torch.empty_like(mine_d) for _ in range(world)]
        dist.all_gather(parts, mine_d)
        assert torch.equal(torch.cat(parts, dim=1), full_d)

        # ---- RowParallelLinear drop-in: bias once, one in-place all-reduce ---------------------------
        g = torch.Generator().manual_seed(3)
        T, IN, OUT = 5, 64, 48
        x = torch.randn(T, IN, generator=g)
        w = torch.randn(OUT, IN, generator=g)
        b = torch.randn(OUT, generator=g)
        shard = IN // world

        class Layer:                                                 # the attributes linear.py:335-411 keeps
            weight = w[:, rank * shard:(rank + 1) * shard].contiguous()
            bias = b
            tp_rank, tp_size = rank, world

        calls = []

        def gloo_all_reduce(t):
            calls.append(tuple(t.shape))
            dist.all_reduce(t)
            return t
        real = ops.tp_all_reduce_
        ops.tp_all_reduce_ = gloo_all_reduce                         # CPU stand-in for the NVLink carrier only
        try:
            y = ours_linear.row_parallel_forward(Layer(), x[:, rank * shard:(rank + 1) * shard].contiguous())
        finally:
            ops.tp_all_reduce_ = real
        assert calls == [(T, OUT)]
        want = torch.nn.functional.linear(x, w, b)
        assert torch.allclose(y, want, atol=1e-4, rtol=1e-4)
        ys = [torch.empty_like(y) for _ in range(world)]
        dist.all_gather(ys, y)
        assert torch.equal(ys[0], ys[1]), "ranks must hold identical activations after the all-reduce"

        # ---- the carrier's form is a pure function of sizes: all ranks agree --------------------------
        forms = [ops.ar_choose_form(world, n, 4 << 20, ops._ar_two_shot_min_bytes(world)) for n in (16, 8192, 2 << 20, 4 << 20)]
        got = [None] * world
        dist.all_gather_object(got, forms)
        assert got[0] == got[1]
        with open(os.path.join(tmp, f"ok{rank}"), "w") as f:
            f.write("ok")
    finally:
        dist.destroy_process_group()


def test_world2_gloo_head_sharding_and_row_parallel_linear(tmp_path):
    port = 29640 + os.getpid() % 200
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    assert all(os.path.exists(os.path.join(str(tmp_path), f"ok{r}")) for r in range(2))
