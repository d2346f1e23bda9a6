"""GPU parity tests (-m gpu): every kernel, called through the C ABI (b200attn.ops ->
libb200attn.so), against the CPU oracle on identical seeded inputs.

Bars: KV scatter bit-exact; attention max-abs <= 2e-2 (BASELINE.json north_star); pointwise
<= 1 bf16 ulp of the oracle (the reference's own compiled form is only defined to that)."""
import glob
import os

import pytest
import torch

import oracle
from helpers import GOLDEN, bf16_ulp, make_forward_case, make_paged_case, scatter_inputs, tensor_sha256

pytestmark = pytest.mark.gpu
ATOL = 2e-2
# prefill kernel variants of b200_attn_prefill: 0 = what the drop-in uses (auto), 1 = mma.sync,
# 2 = tcgen05 two query tiles in flight, 3 = tcgen05 one tile
VARIANTS = [0, 1, 2, 3]
SCALE = 128 ** -0.5


@pytest.fixture(scope="module")
def ops():
    assert torch.cuda.is_available()
    import b200attn
    b200attn.load_library()
    return b200attn.ops


def dev(t):
    return t.cuda() if isinstance(t, torch.Tensor) else t


def maxerr(a, b):
    return (a.float().cpu() - b.float().cpu()).abs().max().item()


# ---------------------------------------------------------------------------- scatter
@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLDEN, "scatter_*.pt"))))
def test_scatter_matches_reference_golden(ops, path):
    meta = torch.load(path)
    qkv, k, v, kc, vc = scatter_inputs(meta)
    Hq, Hkv, D, T = meta["Hq"], meta["Hkv"], meta["D"], meta["T"]
    qkv_d = qkv.cuda()
    v_d = qkv_d[:, (Hq + Hkv) * D:].view(T, Hkv, D)              # strided view, like the reference
    kc_d, vc_d = kc.cuda(), vc.cuda()
    ops.store_kvcache(k.cuda(), v_d, kc_d, vc_d, meta["slot_mapping"].cuda())
    row = Hkv * D
    assert torch.equal(kc_d.cpu().view(-1, row)[meta["touched"]], meta["k_rows"])
    assert torch.equal(vc_d.cpu().view(-1, row)[meta["touched"]], meta["v_rows"])
    oracle.kv_scatter(k, v, kc, vc, meta["slot_mapping"])
    assert torch.equal(kc_d.cpu(), kc) and torch.equal(vc_d.cpu(), vc)


@pytest.mark.parametrize("Hkv", [8, 4, 2, 1, 5])
def test_scatter_random_bit_exact(ops, Hkv):
    g = torch.Generator().manual_seed(Hkv)
    T, D, nb, Hq = 3000, 128, 20, 2 * Hkv
    qkv = torch.randn(T, (Hq + 2 * Hkv) * D, generator=g).bfloat16()
    k = qkv[:, Hq * D:(Hq + Hkv) * D].reshape(T, Hkv, D).contiguous()
    slots = torch.randperm(nb * 256, generator=g)[:T].to(torch.int32)
    slots[torch.rand(T, generator=g) < 0.1] = -1
    kc = torch.randn(nb, 256, Hkv, D, generator=g).bfloat16()
    vc = torch.randn(nb, 256, Hkv, D, generator=g).bfloat16()
    qkv_d = qkv.cuda()
    kc_d, vc_d = kc.cuda(), vc.cuda()
    ops.store_kvcache(k.cuda(), qkv_d[:, (Hq + Hkv) * D:].view(T, Hkv, D), kc_d, vc_d, slots.cuda())
    oracle.kv_scatter(k, qkv[:, (Hq + Hkv) * D:].view(T, Hkv, D), kc, vc, slots)
    assert torch.equal(kc_d.cpu(), kc) and torch.equal(vc_d.cpu(), vc)


def test_scatter_empty_and_all_padding(ops):
    kc = torch.zeros(1, 256, 2, 128, dtype=torch.bfloat16, device="cuda")
    vc = torch.zeros_like(kc)
    e = torch.zeros(0, 2, 128, dtype=torch.bfloat16, device="cuda")
    ops.store_kvcache(e, e, kc, vc, torch.zeros(0, dtype=torch.int32, device="cuda"))
    k = torch.ones(5, 2, 128, dtype=torch.bfloat16, device="cuda")
    ops.store_kvcache(k, k, kc, vc, torch.full((5,), -1, dtype=torch.int32, device="cuda"))
    assert kc.abs().sum().item() == 0 and vc.abs().sum().item() == 0


# ---------------------------------------------------------------------------- decode
RAGGED = [1, 15, 16, 17, 255, 256, 257, 0, 511, 1024, 1300, 33]


@pytest.mark.parametrize("Hq,Hkv", [(32, 8), (16, 4), (8, 2), (4, 1), (16, 8), (5, 1), (8, 1), (40, 8), (16, 1)])
@pytest.mark.parametrize("splits", [None, 1, 3, 8])
def test_decode_ragged_vs_oracle(ops, Hq, Hkv, splits):
    c = make_paged_case(100 + Hq, [1] * len(RAGGED), RAGGED, Hq, Hkv, pad=0, table_width=8)
    want = oracle.attn_decode(c["q"], c["k_cache"], c["v_cache"], c["context_lens"], c["block_tables"], SCALE)
    got = ops.attn_decode(dev(c["q"]), dev(c["k_cache"]), dev(c["v_cache"]), dev(c["context_lens"]),
                          dev(c["block_tables"]), SCALE, num_splits=splits)
    assert maxerr(got, want) <= ATOL
    assert got.cpu()[RAGGED.index(0)].abs().sum() == 0            # graph padding row -> zeros


def test_decode_minus_one_padding_and_nan_garbage_beyond_context(ops):
    """Eager tables are padded with -1 and the cache tail holds whatever torch.empty left:
    neither may leak into the result."""
    lens = [300, 5, 256]
    c = make_paged_case(9, [1] * 3, lens, 8, 2, pad=-1, table_width=4)
    kc, vc = c["k_cache"].clone(), c["v_cache"].clone()
    for b, L in enumerate(lens):                                   # poison every row past the context
        nblk = (L + 255) // 256
        last = int(c["block_tables"][b, nblk - 1])
        if L % 256:
            kc[last, L % 256:] = float("nan")
            vc[last, L % 256:] = float("nan")
    want = oracle.attn_decode(c["q"], c["k_cache"], c["v_cache"], c["context_lens"], c["block_tables"], SCALE)
    got = ops.attn_decode(dev(c["q"]), dev(kc), dev(vc), dev(c["context_lens"]), dev(c["block_tables"]), SCALE)
    assert torch.isfinite(got.float()).all()
    assert maxerr(got, want) <= ATOL


def test_decode_cuda_graph_replay_with_changing_lengths(ops):
    """model_runner.py:640-715 semantics: capture with zero lengths/tables, replay with real ones."""
    Hq, Hkv, B = 16, 8, 8
    c = make_paged_case(21, [1] * B, [700, 3, 0, 256, 90, 0, 1000, 511], Hq, Hkv, pad=0, table_width=6)
    q = torch.zeros(B, Hq, 128, dtype=torch.bfloat16, device="cuda")
    lens = torch.zeros(B, dtype=torch.int32, device="cuda")
    bt = torch.zeros(B, 6, dtype=torch.int32, device="cuda")
    kc, vc = dev(c["k_cache"]), dev(c["v_cache"])
    out = torch.zeros_like(q)
    ops.reserve_decode_workspace(B, Hq, Hkv, 6 * 256, q.device)
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        ops.attn_decode(q, kc, vc, lens, bt, SCALE, out=out)
    torch.cuda.current_stream().wait_stream(s)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        ops.attn_decode(q, kc, vc, lens, bt, SCALE, out=out)
    q.copy_(c["q"]); lens.copy_(c["context_lens"]); bt.copy_(c["block_tables"])
    g.replay()
    want = oracle.attn_decode(c["q"], c["k_cache"], c["v_cache"], c["context_lens"], c["block_tables"], SCALE)
    assert maxerr(out, want) <= ATOL


def test_decode_full_size_cfg3_properties(ops):
    """BASELINE configs[2] at full size (512 seqs x 4096 ctx, 32q/8kv): (a) a random subset of
    sequences against the oracle, (b) V == const => output == const (softmax weights sum to 1),
    (c) split-K result == unsplit result."""
    B, L, Hq, Hkv, D = 512, 4096, 32, 8, 128
    nblk = B * (L // 256)
    g = torch.Generator(device="cuda").manual_seed(0)
    kc = torch.randn(nblk, 256, Hkv, D, generator=g, device="cuda", dtype=torch.bfloat16)
    vc = torch.randn(nblk, 256, Hkv, D, generator=g, device="cuda", dtype=torch.bfloat16)
    q = torch.randn(B, Hq, D, generator=g, device="cuda", dtype=torch.bfloat16)
    perm = torch.randperm(nblk, generator=torch.Generator().manual_seed(1)).to(torch.int32)
    bt = perm.view(B, L // 256).cuda()
    lens = torch.full((B,), L, dtype=torch.int32, device="cuda")
    out1 = ops.attn_decode(q, kc, vc, lens, bt, SCALE, num_splits=1)
    out4 = ops.attn_decode(q, kc, vc, lens, bt, SCALE, num_splits=4)
    assert maxerr(out1, out4) <= 4e-3
    idx = torch.tensor([0, 17, 255, 511])
    sub_blocks = bt[idx.cuda()].cpu()
    uniq, inv = torch.unique(sub_blocks, return_inverse=True)
    want = oracle.attn_decode(q[idx.cuda()].cpu(), kc[uniq.cuda().long()].cpu(), vc[uniq.cuda().long()].cpu(),
                              lens[idx.cuda()].cpu(), inv.to(torch.int32), SCALE)
    assert maxerr(out1[idx.cuda()], want) <= ATOL
    const = torch.randn(Hkv, D, generator=torch.Generator().manual_seed(2)).bfloat16().cuda()
    vc.copy_(const.view(1, 1, Hkv, D).expand_as(vc))
    outc = ops.attn_decode(q, kc, vc, lens, bt, SCALE)
    expect = const.repeat_interleave(Hq // Hkv, dim=0).unsqueeze(0).expand(B, Hq, D)
    assert maxerr(outc, expect) <= 2 ** -7 * const.abs().max().item() + 1e-6


# ---------------------------------------------------------------------------- prefill
@pytest.mark.parametrize("variant", VARIANTS)
@pytest.mark.parametrize("Hq,Hkv", [(16, 8), (4, 1), (5, 1), (8, 2)])
def test_prefill_unpaged_vs_oracle(ops, Hq, Hkv, variant):
    """Sub-paths A/B: Lq == Lk, fresh k (contiguous) and v (strided view of qkv)."""
    lens = [128, 1, 63, 64, 65, 300, 7]
    g = torch.Generator().manual_seed(5)
    T, D = sum(lens), 128
    qkv = torch.randn(T, (Hq + 2 * Hkv) * D, generator=g).bfloat16()
    q = qkv[:, :Hq * D].reshape(T, Hq, D).contiguous()
    k = qkv[:, Hq * D:(Hq + Hkv) * D].reshape(T, Hkv, D).contiguous()
    v = qkv[:, (Hq + Hkv) * D:].view(T, Hkv, D)
    cu = torch.tensor([0] + torch.tensor(lens).cumsum(0).tolist(), dtype=torch.int32)
    want = oracle.attn_varlen(q, k, v, cu, cu, SCALE)
    qkv_d = qkv.cuda()
    got = ops.attn_varlen(q.cuda(), k.cuda(), qkv_d[:, (Hq + Hkv) * D:].view(T, Hkv, D), cu.cuda(), cu.cuda(),
                          max(lens), max(lens), SCALE, variant=variant)
    assert maxerr(got, want) <= ATOL


@pytest.mark.parametrize("variant", VARIANTS)
@pytest.mark.parametrize("Hq,Hkv", [(16, 8), (4, 1), (5, 1), (32, 8)])
def test_prefill_paged_chunked_mixed_vs_oracle(ops, Hq, Hkv, variant):
    """Sub-path C: chunked prefill, prefix hits, decode rows mixed with prefill rows, -1 padding."""
    lens_k = [700, 129, 128, 1, 257, 513, 64, 1000]
    lens_q = [188, 1, 128, 1, 1, 257, 64, 40]
    c = make_paged_case(31 + Hq, lens_q, lens_k, Hq, Hkv, pad=-1)
    want = oracle.attn_varlen(c["q"], c["k_cache"], c["v_cache"], c["cu_seqlens_q"], c["cu_seqlens_k"], SCALE,
                              c["block_tables"])
    got = ops.attn_varlen(dev(c["q"]), dev(c["k_cache"]), dev(c["v_cache"]), dev(c["cu_seqlens_q"]),
                          dev(c["cu_seqlens_k"]), c["max_seqlen_q"], c["max_seqlen_k"], SCALE,
                          block_table=dev(c["block_tables"]), variant=variant)
    assert maxerr(got, want) <= ATOL


@pytest.mark.parametrize("variant", VARIANTS)
def test_prefill_long_causal_row_sampled_vs_oracle(ops, variant):
    """One 4,160-token causal sequence (Hq 8 / Hkv 1, the TP=8 shape): full oracle compare."""
    L = 4160
    c = make_paged_case(77, [L], [L], 8, 1)
    want = oracle.attn_varlen(c["q"], c["k_cache"], c["v_cache"], c["cu_seqlens_q"], c["cu_seqlens_k"], SCALE,
                              c["block_tables"])
    got = ops.attn_varlen(dev(c["q"]), dev(c["k_cache"]), dev(c["v_cache"]), dev(c["cu_seqlens_q"]),
                          dev(c["cu_seqlens_k"]), L, L, SCALE, block_table=dev(c["block_tables"]), variant=variant)
    assert maxerr(got, want) <= ATOL


# ---------------------------------------------------------------------------- planned (persistent) path
@pytest.mark.parametrize("Hq,Hkv", [(16, 8), (4, 1), (5, 1), (32, 8)])
def test_planned_paged_mixed_vs_oracle(ops, Hq, Hkv):
    """Step plan: multi-token rows -> persistent tcgen05 kernel, single-token rows -> decode kernel."""
    lens_k = [700, 129, 128, 1, 257, 513, 64, 1000, 300, 2, 1025]
    lens_q = [188, 1, 128, 1, 1, 257, 64, 40, 300, 2, 513]
    c = make_paged_case(61 + Hq, lens_q, lens_k, Hq, Hkv, pad=-1)
    want = oracle.attn_varlen(c["q"], c["k_cache"], c["v_cache"], c["cu_seqlens_q"], c["cu_seqlens_k"], SCALE,
                              c["block_tables"])
    cu_q, cu_k = dev(c["cu_seqlens_q"]), dev(c["cu_seqlens_k"])
    plan = ops.attn_plan(cu_q, cu_k, sum(lens_q), Hq, Hkv, route_decode=True)
    n_items, n_dec = plan.counts()
    tile = 128 if (Hq // Hkv) % 2 == 0 else 256            # head-pair vs row-pair units
    assert n_dec == sum(1 for x in lens_q if x == 1)
    assert n_items == sum((x + tile - 1) // tile for x in lens_q if x > 1)
    got = torch.full_like(dev(c["q"]), float("nan"))
    for _ in range(2):                                       # the same plan serves every layer of a step
        ops.attn_varlen_planned(dev(c["q"]), dev(c["k_cache"]), dev(c["v_cache"]), cu_q, cu_k, c["max_seqlen_q"],
                                c["max_seqlen_k"], SCALE, block_table=dev(c["block_tables"]), plan=plan, out=got)
    assert torch.isfinite(got.float()).all()
    assert maxerr(got, want) <= ATOL


@pytest.mark.parametrize("Hq,Hkv", [(16, 8), (8, 1), (2, 1), (1, 1), (5, 1), (6, 2)])   # units per item 8, 4, 1, 1, 5, 3
def test_planned_unpaged_vs_oracle(ops, Hq, Hkv):
    lens = [128, 1, 63, 64, 65, 300, 7, 1, 600]
    g = torch.Generator().manual_seed(9)
    T, D = sum(lens), 128
    qkv = torch.randn(T, (Hq + 2 * Hkv) * D, generator=g).bfloat16()
    q = qkv[:, :Hq * D].reshape(T, Hq, D).contiguous()
    k = qkv[:, Hq * D:(Hq + Hkv) * D].reshape(T, Hkv, D).contiguous()
    v = qkv[:, (Hq + Hkv) * D:].view(T, Hkv, D)
    cu = torch.tensor([0] + torch.tensor(lens).cumsum(0).tolist(), dtype=torch.int32)
    want = oracle.attn_varlen(q, k, v, cu, cu, SCALE)
    qkv_d = qkv.cuda()
    got = ops.attn_varlen_planned(q.cuda(), k.cuda(), qkv_d[:, (Hq + Hkv) * D:].view(T, Hkv, D), cu.cuda(), cu.cuda(),
                                  max(lens), max(lens), SCALE)
    assert maxerr(got, want) <= ATOL


def test_planned_long_causal_and_many_units(ops):
    """More work units than SMs (persistent loop, barrier phases across units) + a long causal row."""
    lens = [4160] + [300] * 40 + [1] * 30
    c = make_paged_case(78, lens, lens, 8, 1)
    want = oracle.attn_varlen(c["q"], c["k_cache"], c["v_cache"], c["cu_seqlens_q"], c["cu_seqlens_k"], SCALE,
                              c["block_tables"])
    got = ops.attn_varlen_planned(dev(c["q"]), dev(c["k_cache"]), dev(c["v_cache"]), dev(c["cu_seqlens_q"]),
                                  dev(c["cu_seqlens_k"]), max(lens), max(lens), SCALE, block_table=dev(c["block_tables"]))
    assert maxerr(got, want) <= ATOL


def test_planned_chunked_prefix_rows(ops):
    """Chunked prefill over a long cached prefix: Lq << Lk, bottom-right causal offset, ragged tails."""
    lens_q = [64, 64, 200, 513, 1, 31]
    lens_k = [2112, 2111, 1224, 1537, 2048, 31]
    c = make_paged_case(79, lens_q, lens_k, 8, 2)
    want = oracle.attn_varlen(c["q"], c["k_cache"], c["v_cache"], c["cu_seqlens_q"], c["cu_seqlens_k"], SCALE,
                              c["block_tables"])
    kc, vc = dev(c["k_cache"]), dev(c["v_cache"])
    got = ops.attn_varlen_planned(dev(c["q"]), kc, vc, dev(c["cu_seqlens_q"]), dev(c["cu_seqlens_k"]), max(lens_q),
                                  max(lens_k), SCALE, block_table=dev(c["block_tables"]))
    assert maxerr(got, want) <= ATOL


def test_prefill_full_size_cfg5_properties(ops):
    """BASELINE configs[4] at full per-rank size (TP8: 8q/1kv; one 32,832-token row + 63 rows of 64
    new tokens over the same 32,768-token cached prefix).  Too big for the CPU oracle, so:
    (a) sampled query rows against the oracle restricted to those rows, (b) V == const => output ==
    const, (c) linearity in V, (d) branches given identical queries and tails give identical outputs."""
    Hq, Hkv, D, P = 8, 1, 128, 256
    prefix, tail, nbr = 32768, 64, 63
    lens_q = [prefix + tail] + [tail] * nbr
    lens_k = [prefix + tail] * (nbr + 1)
    shared = prefix // P
    nblk = shared + (nbr + 1) + 1
    g = torch.Generator(device="cuda").manual_seed(5)
    kc = torch.randn(nblk, P, Hkv, D, generator=g, device="cuda", dtype=torch.bfloat16)
    vc = torch.randn(nblk, P, Hkv, D, generator=g, device="cuda", dtype=torch.bfloat16)
    bt = torch.empty(nbr + 1, shared + 1, dtype=torch.int32)
    bt[:, :shared] = torch.arange(shared, dtype=torch.int32)
    bt[:, shared] = torch.arange(shared, shared + nbr + 1, dtype=torch.int32)
    # branches 1 and 2 get the same tail K/V and the same queries (property d)
    kc[shared + 2].copy_(kc[shared + 1]); vc[shared + 2].copy_(vc[shared + 1])
    T = sum(lens_q)
    q = torch.randn(T, Hq, D, generator=g, device="cuda", dtype=torch.bfloat16)
    cu_q = torch.tensor([0] + list(torch.tensor(lens_q).cumsum(0)), dtype=torch.int32)
    cu_k = torch.tensor([0] + list(torch.tensor(lens_k).cumsum(0)), dtype=torch.int32)
    q[cu_q[2]:cu_q[3]].copy_(q[cu_q[1]:cu_q[2]])
    run = lambda v: ops.attn_varlen_planned(q, kc, v, cu_q.cuda(), cu_k.cuda(), max(lens_q), max(lens_k), SCALE,  # noqa: E731
                                            block_table=bt.cuda())
    out = run(vc)
    assert torch.isfinite(out.float()).all()
    # (d)
    assert torch.equal(out[cu_q[1]:cu_q[2]], out[cu_q[2]:cu_q[3]])
    # (a) rows of the long sequence and of one branch, exact attention in fp32 on the CPU
    kflat = kc[bt[0].long().cuda()].reshape(-1, D).float().cpu()
    vflat = vc[bt[0].long().cuda()].reshape(-1, D).float().cpu()
    for r in (0, 127, 128, 4095, 20000, prefix + tail - 1):
        qr = q[r].float().cpu()                                       # [Hq, D]
        sc = (qr @ kflat[: r + 1].T) * SCALE
        want = torch.softmax(sc, dim=-1) @ vflat[: r + 1]
        assert maxerr(out[r], want) <= ATOL
    k5 = kc[bt[5].long().cuda()].reshape(-1, D).float().cpu()
    v5 = vc[bt[5].long().cuda()].reshape(-1, D).float().cpu()
    for i in (0, 63):
        qr = q[cu_q[5] + i].float().cpu()
        vis = prefix + i + 1
        want = torch.softmax((qr @ k5[:vis].T) * SCALE, dim=-1) @ v5[:vis]
        assert maxerr(out[cu_q[5] + i], want) <= ATOL
    # (c) out(V1 + V2) == out(V1) + out(V2) up to bf16 rounding of inputs/outputs
    v2 = torch.randn(vc.shape, generator=g, device="cuda", dtype=torch.bfloat16)
    vs = (vc.float() + v2.float()).bfloat16()
    lin = run(vs).float() - (out.float() + run(v2).float())
    assert lin.abs().max().item() <= 3e-2
    # (b)
    const = torch.randn(Hkv, D, generator=torch.Generator().manual_seed(6)).bfloat16().cuda()
    vconst = const.view(1, 1, Hkv, D).expand_as(vc).contiguous()
    outc = run(vconst)
    expect = const.repeat_interleave(Hq // Hkv, dim=0).unsqueeze(0).expand(T, Hq, D)
    assert maxerr(outc, expect) <= 2 ** -7 * const.abs().max().item() + 1e-6


def test_prefill_full_size_warmup_call_properties(ops):
    """The largest attention call of a run: the reference's warm-up (model_runner.py:311-316), one
    unpaged 40,960-token causal row (here at the TP8 head count 4q/1kv to bound time).  Sampled
    rows against exact fp32 attention; V == const => output == const."""
    T, Hq, Hkv, D = 40960, 4, 1, 128
    g = torch.Generator(device="cuda").manual_seed(7)
    q = torch.randn(T, Hq, D, generator=g, device="cuda", dtype=torch.bfloat16)
    k = torch.randn(T, Hkv, D, generator=g, device="cuda", dtype=torch.bfloat16)
    v = torch.randn(T, Hkv, D, generator=g, device="cuda", dtype=torch.bfloat16)
    cu = torch.tensor([0, T], dtype=torch.int32, device="cuda")
    out = ops.attn_varlen_planned(q, k, v, cu, cu, T, T, SCALE)
    kf, vf = k[:, 0].float().cpu(), v[:, 0].float().cpu()
    for r in (0, 1, 129, 8191, 40959):
        want = torch.softmax((q[r].float().cpu() @ kf[: r + 1].T) * SCALE, dim=-1) @ vf[: r + 1]
        assert maxerr(out[r], want) <= ATOL
    const = torch.randn(Hkv, D, generator=torch.Generator().manual_seed(8)).bfloat16().cuda()
    outc = ops.attn_varlen_planned(q, k, const.view(1, Hkv, D).expand(T, Hkv, D).contiguous(), cu, cu, T, T, SCALE)
    assert maxerr(outc, const.expand(T, Hq, D)) <= 2 ** -7 * const.abs().max().item() + 1e-6


# ---------------------------------------------------------------------------- Attention.forward
def test_attention_forward_prefix_sharing_same_step(ops):
    """a branch scheduled in the same step reads prefix blocks that another sequence
    writes in that very forward pass; the scatter must be ordered before every read."""
    from b200attn.layers.attention import Attention
    from b200attn.utils.context import reset_context, set_context
    Hq, Hkv, D, P = 8, 2, 128, 256
    g = torch.Generator().manual_seed(3)
    prefix, tail = 512, 40
    lens_q = [prefix + tail, tail, tail]                 # seq 0 computes the prefix, 1 and 2 only their tails
    lens_k = [prefix + tail] * 3
    T = sum(lens_q)
    q = torch.randn(T, Hq, D, generator=g).bfloat16()
    k = torch.randn(T, Hkv, D, generator=g).bfloat16()
    v = torch.randn(T, Hkv, D, generator=g).bfloat16()
    bt = torch.tensor([[5, 2, 7], [5, 2, 1], [5, 2, 0]], dtype=torch.int32)   # shared blocks 5, 2
    slots = []
    for b in range(3):
        start = 0 if b == 0 else prefix
        for pos in range(start, prefix + tail):
            slots.append(int(bt[b, pos // P]) * P + pos % P)
    slot_mapping = torch.tensor(slots, dtype=torch.int32)
    cu_q = torch.tensor([0, lens_q[0], lens_q[0] + tail, T], dtype=torch.int32)
    cu_k = torch.tensor([0, lens_k[0], 2 * lens_k[0], 3 * lens_k[0]], dtype=torch.int32)
    ctx = (cu_q, cu_k, max(lens_q), max(lens_k), slot_mapping, torch.tensor(lens_k, dtype=torch.int32), bt, None)
    kc = torch.full((8, P, Hkv, D), float("nan")).bfloat16()      # stale garbage everywhere
    vc = kc.clone()
    attn = Attention(Hq, D, SCALE, Hkv)
    attn.k_cache, attn.v_cache = kc.cuda(), vc.cuda()
    set_context(*[dev(t) for t in ctx])
    got = attn(q.cuda(), k.cuda(), v.cuda())
    reset_context()
    want = oracle.attention_forward(q, k, v, kc, vc, oracle.Context(*ctx), SCALE)
    assert torch.isfinite(got.float()).all()
    assert maxerr(got, want) <= ATOL
    written = ~torch.isnan(kc.float()).any(-1).any(-1)
    assert torch.equal(attn.k_cache.cpu()[written], kc[written])


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLDEN, "attn_ref_*.pt"))))
def test_attention_forward_matches_reference_golden(ops, path):
    """The drop-in `Attention.forward` against the output of the UNMODIFIED reference
    `Attention.forward` on the same seeded inputs (fixtures from tools/make_golden_gpu.py):
    attention within 2e-2, cache after the step bit-exact."""
    from b200attn.layers.attention import Attention
    from b200attn.utils.context import reset_context, set_context
    m = torch.load(path)
    c = make_forward_case(m)
    Hq, Hkv = m["Hq"], m["Hkv"]
    attn = Attention(Hq, 128, SCALE, Hkv)
    attn.k_cache, attn.v_cache = c["k_cache"].cuda(), c["v_cache"].cuda()
    qkv = c["qkv"].cuda()
    n = qkv.shape[0]
    set_context(*[dev(t) for t in c["ctx"]])
    got = attn(c["q"].cuda(), c["k"].cuda(), qkv[:, (Hq + Hkv) * 128:].view(n, Hkv, 128))
    reset_context()
    r = c["num_real"]
    assert maxerr(got[:r], m["out"][:r]) <= ATOL
    assert tensor_sha256(attn.k_cache) == m["k_cache_sha256"]
    assert tensor_sha256(attn.v_cache) == m["v_cache_sha256"]


def test_attention_forward_config0_prefill_then_decode(ops):
    import __graft_entry__ as g
    g.smoke()


def test_attention_eager_decode_routes_to_decode_kernel(ops):
    """enforce_eager / bs > 512: varlen call with every q_len == 1."""
    from b200attn.layers.attention import Attention
    from b200attn.utils.context import reset_context, set_context
    lens = [5, 300, 256, 1]
    c = make_paged_case(13, [1] * 4, lens, 16, 8, pad=-1)
    attn = Attention(16, 128, SCALE, 8)
    attn.k_cache, attn.v_cache = dev(c["k_cache"]), dev(c["v_cache"])
    set_context(dev(c["cu_seqlens_q"]), dev(c["cu_seqlens_k"]), 1, max(lens),
                torch.full((4,), -1, dtype=torch.int32, device="cuda"), dev(c["context_lens"]),
                dev(c["block_tables"]), None)
    knew = torch.zeros(4, 8, 128, dtype=torch.bfloat16, device="cuda")
    got = attn(dev(c["q"]), knew, knew)
    reset_context()
    want = oracle.attn_varlen(c["q"], c["k_cache"], c["v_cache"], c["cu_seqlens_q"], c["cu_seqlens_k"], SCALE,
                              c["block_tables"])
    assert maxerr(got, want) <= ATOL


# ---------------------------------------------------------------------------- pointwise
def _ulp_ok(got, want, n=1):
    got, want = got.float().cpu(), want.float()
    return ((got - want).abs() <= n * bf16_ulp(want) + 1e-30).all().item()


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLDEN, "pointwise_*.pt"))))
def test_pointwise_vs_reference_golden(ops, path):
    m = torch.load(path)
    T, Hq, Hkv, D = m["T"], m["Hq"], m["Hkv"], m["D"]
    qkv = m["qkv"].cuda()
    q, k, _ = qkv.split([Hq * D, Hkv * D, Hkv * D], dim=-1)
    cache = oracle.make_cos_sin_cache(D, m["max_pos"], m["base"]).view(-1, D).cuda()
    qo, ko = ops.qk_norm_rope(q.view(T, Hq, D), k.view(T, Hkv, D), m["positions"].cuda(), m["q_weight"].cuda(),
                              m["k_weight"].cuda(), cache, m["eps"])
    assert _ulp_ok(qo, m["q_out"]) and _ulp_ok(ko, m["k_out"])
    assert (qo.cpu() == m["q_out"]).float().mean() > 0.995       # almost always identical bits
    act = ops.silu_and_mul(m["act_in"].cuda())
    assert _ulp_ok(act, oracle.silu_and_mul(m["act_in"], "compiled"))
    assert _ulp_ok(act, m["act_out_eager"], 2)
    y = ops.rms_norm(m["h"].cuda(), m["h_weight"].cuda(), m["eps"])
    assert _ulp_ok(y, m["h_norm"])
    r = m["r"].clone().cuda()
    y2, r2 = ops.rms_norm(m["h"].cuda(), m["h_weight"].cuda(), m["eps"], r)
    assert _ulp_ok(y2, m["hr_norm"]) and torch.equal(r2.cpu(), m["hr_res"])


def test_rotation_only_and_large_shapes(ops):
    g = torch.Generator().manual_seed(8)
    T, Hq, Hkv, D = 1000, 32, 8, 128
    qkv = torch.randn(T, (Hq + 2 * Hkv) * D, generator=g).bfloat16()
    q, k, _ = qkv.split([Hq * D, Hkv * D, Hkv * D], dim=-1)
    q, k = q.view(T, Hq, D), k.view(T, Hkv, D)
    pos = torch.randint(0, 40960, (T,), generator=g)
    cache = oracle.make_cos_sin_cache(D, 40960, 1e6)
    cos, sin = cache[pos].chunk(2, dim=-1)
    wq, wk = oracle.rope(q, cos, sin), oracle.rope(k, cos, sin)
    qkv_d = qkv.cuda()
    qd, kd, _ = qkv_d.split([Hq * D, Hkv * D, Hkv * D], dim=-1)
    gq, gk = ops.qk_norm_rope(qd.view(T, Hq, D), kd.view(T, Hkv, D), pos.cuda(), None, None, cache.view(-1, D).cuda())
    assert _ulp_ok(gq, wq) and _ulp_ok(gk, wk)
    x = (3 * torch.randn(777, 2 * 12288, generator=g)).bfloat16()
    assert _ulp_ok(ops.silu_and_mul(x.cuda()), oracle.silu_and_mul(x))
    h = torch.randn(300, 5120, generator=g).bfloat16()
    w = (1 + 0.1 * torch.randn(5120, generator=g)).bfloat16()
    assert _ulp_ok(ops.rms_norm(h.cuda(), w.cuda(), 1e-6), oracle.rms_norm(h, w))


@pytest.mark.parametrize("Hq,Hkv,T,need_k", [(16, 8, 300, True), (32, 8, 513, False), (8, 1, 77, True), (5, 1, 1, True)])
def test_fused_norm_rope_store_equals_unfused_then_scatter(ops, Hq, Hkv, T, need_k):
    """one launch = qk_norm_rope (4a) + kv_scatter (1).  q'/k' must carry the same bits as the
    unfused kernel (which is held to the oracle above) and the caches the same bits as the oracle's scatter
    of (k', v) -- including rows with slot -1 left untouched."""
    g = torch.Generator().manual_seed(100 + T)
    D, P, nblk = 128, 256, 6
    qkv = torch.randn(T, (Hq + 2 * Hkv) * D, generator=g).bfloat16()
    pos = torch.randint(0, 40960, (T,), generator=g)
    wq = (1 + 0.1 * torch.randn(D, generator=g)).bfloat16()
    wk = (1 + 0.1 * torch.randn(D, generator=g)).bfloat16()
    cache = oracle.make_cos_sin_cache(D, 40960, 1e6).view(-1, D)
    slots = torch.randperm(nblk * P, generator=g)[:T].to(torch.int32)
    slots[::7] = -1
    kc0 = torch.randn(nblk, P, Hkv, D, generator=g).bfloat16()
    vc0 = torch.randn(nblk, P, Hkv, D, generator=g).bfloat16()
    qkv_d = qkv.cuda()
    qd, kd, vd = (t.view(T, -1, D) for t in qkv_d.split([Hq * D, Hkv * D, Hkv * D], dim=-1))
    q_ref, k_ref = ops.qk_norm_rope(qd, kd, pos.cuda(), wq.cuda(), wk.cuda(), cache.cuda(), 1e-6)
    kc_want, vc_want = kc0.clone(), vc0.clone()
    oracle.kv_scatter(k_ref.cpu(), vd.cpu(), kc_want, vc_want, slots)
    kc, vc = kc0.cuda(), vc0.cuda()
    q_got, k_got = ops.qk_norm_rope_store(qd, kd, vd, pos.cuda(), wq.cuda(), wk.cuda(), cache.cuda(), 1e-6,
                                          kc, vc, slots.cuda(), need_k_out=need_k)
    assert torch.equal(q_got, q_ref)
    assert (k_got is None) == (not need_k)
    if need_k:
        assert torch.equal(k_got, k_ref)
    assert torch.equal(kc.cpu(), kc_want) and torch.equal(vc.cpu(), vc_want)
    # rotation-only form (qkv_bias models): NULL weights
    q2, k2 = ops.qk_norm_rope_store(qd, kd, vd, pos.cuda(), None, None, cache.cuda(), 1e-6, kc, vc, slots.cuda())
    q2r, k2r = ops.qk_norm_rope(qd, kd, pos.cuda(), None, None, cache.cuda())
    assert torch.equal(q2, q2r) and torch.equal(k2, k2r)


@pytest.mark.parametrize("Hq,Hkv,splits", [(8, 2, 1), (8, 2, 3), (20, 4, 1), (16, 1, 2)])
def test_decode_many_items_ragged_with_padding_rows(ops, Hq, Hkv, splits):
    """Several waves of (sequence, kv head, split) CTAs; ragged lengths with padding rows scattered through
    the batch; launched three times back to back on one workspace: identical bits every time."""
    import random
    rnd = random.Random(Hq * 31 + splits)
    B = 1200 // Hkv + 7
    lens = [rnd.choice([0, 1, 16, 40, 255, 256, 300, 600, 700]) for _ in range(B)]
    c = make_paged_case(500 + Hq, [1] * B, lens, Hq, Hkv, pad=0, table_width=3)
    want = oracle.attn_decode(c["q"], c["k_cache"], c["v_cache"], c["context_lens"], c["block_tables"], SCALE)
    args = (dev(c["q"]), dev(c["k_cache"]), dev(c["v_cache"]), dev(c["context_lens"]), dev(c["block_tables"]), SCALE)
    outs = [ops.attn_decode(*args, num_splits=splits).clone() for _ in range(3)]
    assert maxerr(outs[0], want) <= ATOL
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[1], outs[2])
    for b, L in enumerate(lens):
        if L == 0:
            assert outs[0][b].abs().sum().item() == 0


# ---------------------------------------------------------------------------- (f3) TP all-reduce
def test_peer_memory_allreduce_world1_and_argument_errors(ops):
    """Single-GPU slice of csrc/allreduce.cu (the multi-GPU parity run is tools/ar_check.py, results
    under profiles/): a one-rank communicator must leave the data unchanged bit for bit over repeated
    calls (slot alternation, device-side epoch), also under CUDA-graph replay, and reject messages
    that do not fit."""
    ar = ops.TpAllReduce(0, 1, 64 << 10, lambda mine: [mine])
    x = torch.randn(3 * 4096, device="cuda").bfloat16()
    want = x.clone()
    for _ in range(5):
        ar.all_reduce_(x)
    for form in (0, 3, 1, 2):                                   # 1 and 2 degrade to one-shot when world <= 2
        ar.all_reduce_(x, two_shot=form)
    assert torch.equal(x, want)
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        ar.all_reduce_(x)
    torch.cuda.current_stream().wait_stream(s)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        ar.all_reduce_(x)
        ar.all_reduce_(x)
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    assert torch.equal(x, want)
    assert not ar.fits(torch.zeros(64 << 10, device="cuda").bfloat16())        # 128 KB > 64 KB
    assert not ar.fits(torch.zeros(12, device="cuda").bfloat16())              # not a multiple of 8 elements
    with pytest.raises(RuntimeError):
        ar.all_reduce_(torch.zeros(12, device="cuda").bfloat16())
    assert ar.status() == 0                                     # no wait timed out
    ar.destroy()                                                # b200_ar_destroy: frees the buffers, closes IPC maps


# ------------------------------------------------------------------ round 2 additions
@pytest.mark.parametrize("lens", [[1], [1, 1, 1], [1, 5, 1]])
def test_planned_unpaged_single_token_prompts(ops, lens):
    """A step of 1-token prompts without a prefix hit: block_tables is None (sum k == sum q, model_runner.py:483)
    and max_seqlen_q == 1.  The reference calls flash_attn_varlen_func and gets out = v; unpaged 1-token rows are
    NOT routed to the decode kernel, so the prefill kernel must run them (they used to be left uninitialised)."""
    Hq, Hkv, D = 16, 8, 128
    g = torch.Generator().manual_seed(3)
    T = sum(lens)
    q = torch.randn(T, Hq, D, generator=g).bfloat16()
    k = torch.randn(T, Hkv, D, generator=g).bfloat16()
    v = torch.randn(T, Hkv, D, generator=g).bfloat16()
    cu = torch.tensor([0] + torch.tensor(lens).cumsum(0).tolist(), dtype=torch.int32)
    want = oracle.attn_varlen(q, k, v, cu, cu, SCALE)
    out = torch.full((T, Hq, D), float("nan"), dtype=torch.bfloat16, device="cuda")
    got = ops.attn_varlen_planned(q.cuda(), k.cuda(), v.cuda(), cu.cuda(), cu.cuda(), max(lens), max(lens), SCALE, out=out)
    assert torch.isfinite(got.float()).all()
    assert maxerr(got, want) <= ATOL
    if max(lens) == 1:                                           # one visible key per row: softmax weight 1 => out == v
        assert torch.equal(got.cpu(), v.repeat_interleave(Hq // Hkv, dim=1))


def test_decode_fused_split_merge_leaves_tickets_zero(ops):
    """The last split of a (row, kv head) to arrive merges all of them inside the decode launch; the arrival
    tickets at the head of the workspace re-arm themselves, so every launch (and every CUDA-graph replay)
    finds zeros.  Ragged rows: some use every split, some fewer, some are padding."""
    Hq, Hkv = 8, 2
    lens = [4096, 300, 0, 1, 257, 2000, 16, 4095]
    c = make_paged_case(77, [1] * len(lens), lens, Hq, Hkv, pad=0, table_width=16)
    want = oracle.attn_decode(c["q"], c["k_cache"], c["v_cache"], c["context_lens"], c["block_tables"], SCALE)
    args = [dev(c[n]) for n in ("q", "k_cache", "v_cache", "context_lens", "block_tables")]
    outs = [ops.attn_decode(*args, SCALE, num_splits=s) for s in (8, 8, 3, 8)]
    for o in outs:
        assert maxerr(o, want) <= ATOL
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[0], outs[3])      # same split count => same bits
    ws = ops._WS_CACHE[torch.cuda.current_device()]
    assert int(ws[:64 * 1024].view(torch.int32).abs().sum()) == 0


def test_nvls_fused_allreduce_add_rmsnorm_world1(ops):
    """Single-GPU slice of csrc/allreduce_nvls.cu (multi-GPU parity: bench.py --gpus N `allreduce` leg and
    tools/ar_check.py): with one rank the sum is the input itself, so the launch must give the residual of
    add_rms_forward bit for bit and its normed output within 1 bf16 ulp of the oracle (the sum of squares is
    accumulated in a different order than in b200_rmsnorm, so the two kernels may differ in a last bit) -- both
    forms, repeated calls (slot alternation), CUDA-graph replay; weight None is a plain all-reduce (identity here)."""
    nv = ops.NvlsAllReduce(None, 1 << 20)
    g = torch.Generator(device="cuda").manual_seed(5)
    for T, H in [(1, 1024), (7, 4096), (64, 5120), (33, 1024)]:
        x = torch.randn(T, H, generator=g, device="cuda").bfloat16()
        r = torch.randn(T, H, generator=g, device="cuda").bfloat16()
        w = (1 + 0.1 * torch.randn(H, generator=g, device="cuda")).bfloat16()
        want_y, want_r = ops.rms_norm(x, w, 1e-6, r.clone())
        oy, orr = oracle.add_rms_norm(x.cpu(), r.cpu(), w.cpu(), 1e-6)
        assert _ulp_ok(want_y, oy) and torch.equal(want_r.cpu(), orr)
        for form in (1, 0, 2, 3, None, 2, 0, 3):              # every form, interleaved: the flag counters stay in step
            r2 = r.clone()
            y, r3 = nv.all_reduce_add_rms_norm(x, r2, w, 1e-6, form=form)
            assert r3 is r2 and torch.equal(r2, want_r) and _ulp_ok(y, oy), (T, H, form)
            assert (y == want_y).float().mean() > 0.99, (T, H, form)
        assert torch.equal(nv.all_reduce_add_rms_norm(x, None, None, 0.0), x)          # plain all-reduce
        assert _ulp_ok(nv.all_reduce_add_rms_norm(x, None, w, 1e-6), oracle.rms_norm(x.cpu(), w.cpu(), 1e-6))
    x = torch.randn(16, 4096, generator=g, device="cuda").bfloat16()
    w = torch.ones(4096, device="cuda").bfloat16()
    r0 = torch.randn(16, 4096, generator=g, device="cuda").bfloat16()
    r, y = r0.clone(), torch.empty_like(x)
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        nv.all_reduce_add_rms_norm(x, r, w, 1e-6, out=y)
    torch.cuda.current_stream().wait_stream(s)
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr):
        nv.all_reduce_add_rms_norm(x, r, w, 1e-6, out=y)
        nv.all_reduce_add_rms_norm(x, r, w, 1e-6, out=y, form=3)
    r.copy_(r0)
    gr.replay()
    rr = r0.clone()
    ops.rms_norm(x, w, 1e-6, rr)
    want_y, _ = ops.rms_norm(x, w, 1e-6, rr)
    torch.cuda.synchronize()
    assert torch.equal(r, rr) and (y == want_y).float().mean() > 0.99 and _ulp_ok(y, want_y.cpu(), 2)
    assert not nv.fits(torch.zeros(1 << 20, 8, device="cuda").bfloat16())                 # 16 MB > 1 MB
    assert nv.status() == 0
    nv.destroy()


def test_deferred_all_reduce_is_never_left_pending(ops):
    """Host logic of the fused path on one GPU: outside `deferred_all_reduce` RowParallelLinear's output is reduced
    at once; nothing stays pending after the context exits."""
    y = torch.randn(4, 1024, device="cuda").bfloat16()
    assert ops._PENDING is None
    with ops.deferred_all_reduce():
        assert ops._DEFER_DEPTH == 1
    assert ops._DEFER_DEPTH == 0 and ops._PENDING is None
    w = torch.ones(1024, device="cuda").bfloat16()
    out, res = ops.tp_add_rms_norm(y, y.clone(), w, 1e-6)       # no pending sum: plain fused add + RMSNorm
    assert out.shape == y.shape and res.shape == y.shape


def test_cfg4_traced_mixed_batch_vs_oracle(ops):
    """BASELINE configs[3] at its own traced shape: serving_bench, Qwen3-14B at TP=2 (20 q /
    4 kv heads per rank, GQA group 5), max-num-batched-tokens 1024: B = 447 rows = 8 x 128-token prompt chunks +
    439 x 1-token decode rows, Lk 128..228.  Step plan + persistent prefill + routed decode vs the oracle."""
    Hq, Hkv = 20, 4
    rnd = torch.Generator().manual_seed(44)
    dec_k = torch.randint(129, 229, (439,), generator=rnd).tolist()
    lens_q = [128] * 8 + [1] * 439
    lens_k = [128] * 8 + dec_k
    order = torch.randperm(447, generator=rnd).tolist()          # running-first scheduling interleaves them
    lens_q = [lens_q[i] for i in order]
    lens_k = [lens_k[i] for i in order]
    c = make_paged_case(45, lens_q, lens_k, Hq, Hkv, pad=-1)
    want = oracle.attn_varlen(c["q"], c["k_cache"], c["v_cache"], c["cu_seqlens_q"], c["cu_seqlens_k"], SCALE,
                              c["block_tables"])
    cu_q, cu_k = dev(c["cu_seqlens_q"]), dev(c["cu_seqlens_k"])
    plan = ops.attn_plan(cu_q, cu_k, sum(lens_q), Hq, Hkv, route_decode=True)
    n_items, n_dec = plan.counts()
    assert n_dec == 439 and n_items == 8                          # 256-row tiles at an odd GQA group
    got = torch.full_like(dev(c["q"]), float("nan"))
    ops.attn_varlen_planned(dev(c["q"]), dev(c["k_cache"]), dev(c["v_cache"]), cu_q, cu_k, 128, max(lens_k), SCALE,
                            block_table=dev(c["block_tables"]), plan=plan, out=got)
    assert torch.isfinite(got.float()).all()
    assert maxerr(got, want) <= ATOL


@pytest.mark.parametrize("Hq,Hkv", [(16, 4), (8, 2), (4, 1)])
def test_decode_full_size_cfg3_tp_shapes(ops, Hq, Hkv):
    """BASELINE configs[2] at the per-rank head counts of TP=2/4/8, full size (512 seqs x 4096 ctx): sampled rows vs
    the oracle, automatic split count (what the engine uses) == unsplit within rounding."""
    B, L, D = 512, 4096, 128
    nblk = B * (L // 256)
    g = torch.Generator(device="cuda").manual_seed(10 + Hkv)
    kc = torch.randn(nblk, 256, Hkv, D, generator=g, device="cuda", dtype=torch.bfloat16)
    vc = torch.randn(nblk, 256, Hkv, D, generator=g, device="cuda", dtype=torch.bfloat16)
    q = torch.randn(B, Hq, D, generator=g, device="cuda", dtype=torch.bfloat16)
    perm = torch.randperm(nblk, generator=torch.Generator().manual_seed(1)).to(torch.int32)
    bt = perm.view(B, L // 256).cuda()
    lens = torch.full((B,), L, dtype=torch.int32, device="cuda")
    out_auto = ops.attn_decode(q, kc, vc, lens, bt, SCALE)
    out1 = ops.attn_decode(q, kc, vc, lens, bt, SCALE, num_splits=1)
    assert maxerr(out_auto, out1) <= 4e-3
    idx = torch.tensor([0, 100, 301, 511])
    sub_blocks = bt[idx.cuda()].cpu()
    uniq, inv = torch.unique(sub_blocks, return_inverse=True)
    want = oracle.attn_decode(q[idx.cuda()].cpu(), kc[uniq.cuda().long()].cpu(), vc[uniq.cuda().long()].cpu(),
                              lens[idx.cuda()].cpu(), inv.to(torch.int32), SCALE)
    assert maxerr(out_auto[idx.cuda()], want) <= ATOL


@pytest.mark.parametrize("env", [{"B200ATTN_PREFILL_WGS": "2"}, {"B200ATTN_PREFILL_EMU": "1"}, {"B200ATTN_PREFILL_WGS": "2", "B200ATTN_PREFILL_EMU": "2"},
                                 {"B200ATTN_DECODE_FUSED_REDUCE": "0", "B200ATTN_DECODE_SPLITS": "3"},
                                 # tensor-map cache of 3 entries: the 4th look-up of EVERY prefill launch empties it while the
                                 # first three maps are in use (they used to be pointers into the cache)
                                 {"B200ATTN_MAP_CACHE_CAP": "3"}, {"B200ATTN_MAP_CACHE_CAP": "1"}])
def test_measured_and_not_adopted_variants_still_match_the_oracle(env):
    """The A/B knobs kept in the shipped library (two softmax warpgroups per query tile, polynomial exp2, the separate
    reduce launch) are read once per process, so each runs in a subprocess: a mixed paged batch (prefill rows through the
    persistent kernel, 1-token rows through the decode kernel) against the oracle."""
    import subprocess
    import sys
    code = r'''
import sys, torch
sys.path[:0] = [%r, %r, %r]
import oracle
from helpers import make_paged_case
from b200attn import ops
lens_q = [300, 1, 128, 1, 700, 64, 1, 257]
lens_k = [300, 900, 640, 17, 700, 1000, 256, 257]
for Hq, Hkv in ((16, 8), (8, 1), (5, 1)):
    c = make_paged_case(31 + Hq, lens_q, lens_k, Hq, Hkv)
    want = oracle.attn_varlen(c["q"], c["k_cache"], c["v_cache"], c["cu_seqlens_q"], c["cu_seqlens_k"], 128 ** -0.5, c["block_tables"])
    d = lambda t: t.cuda()
    got = ops.attn_varlen_planned(d(c["q"]), d(c["k_cache"]), d(c["v_cache"]), d(c["cu_seqlens_q"]), d(c["cu_seqlens_k"]),
                                  c["max_seqlen_q"], c["max_seqlen_k"], 128 ** -0.5, block_table=d(c["block_tables"]))
    err = (got.float().cpu() - want.float()).abs().max().item()
    assert torch.isfinite(got.float()).all() and err <= 2e-2, (Hq, Hkv, err)
print("ok")
''' % (os.path.dirname(os.path.abspath(__file__)), os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
       os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "nano-vllm-v1_b200"))
    r = subprocess.run([sys.executable, "-c", code], env={**os.environ, **env}, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "ok" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]


@pytest.mark.parametrize("Hq,Hkv", [(8, 1), (32, 8), (16, 8), (20, 4)])
def test_step_plan_unit_table_is_complete_and_heaviest_first(ops, Hq, Hkv):
    """The plan's per-unit table (csrc/launch.h): every (sequence, query tile, sub-unit) exactly once, packed tiles
    (<= 64 valid rows, group divisible by 4) with half the units and the pack flag, records carrying q0/Lq/k0/Lk, and the
    whole list in non-increasing order of K/V tiles (what the static snake over the CTAs relies on)."""
    lens_q = [64, 300, 1, 128, 40, 1, 513, 65, 2]
    lens_k = [700, 300, 900, 128, 1000, 5, 513, 65, 258]
    B, T = len(lens_q), sum(lens_q)
    cu_q = torch.tensor([0] + torch.tensor(lens_q).cumsum(0).tolist(), dtype=torch.int32, device="cuda")
    cu_k = torch.tensor([0] + torch.tensor(lens_k).cumsum(0).tolist(), dtype=torch.int32, device="cuda")
    plan = ops.attn_plan(cu_q, cu_k, T, Hq, Hkv, route_decode=True)
    w = plan.buf.view(torch.int32).cpu()
    G = Hq // Hkv
    tile = 128 if G % 2 == 0 else 256
    per_item = Hq // 2 if tile == 128 else Hq
    pack_ok = tile == 128 and G % 4 == 0
    max_items = T // 128 + B
    off = 16 + ((B + 3) & ~3) + 4 * max_items
    n_units, n_dec, tile_rows, n_items = (int(x) for x in w[:4])
    assert tile_rows == tile and n_dec == sum(1 for x in lens_q if x == 1)
    want = {}
    for b, (lq, lk) in enumerate(zip(lens_q, lens_k)):
        if lq <= 1:
            continue
        for mt in range((lq + tile - 1) // tile):
            packed = pack_ok and lq - mt * 128 <= 64
            for sub in range(per_item // 2 if packed else per_item):
                want[(b, mt, sub)] = 2 if packed else 1
    assert n_items == len({(b, mt) for b, mt, _ in want}) and n_units == len(want)
    recs = w[off:off + 8 * n_units].view(n_units, 8).tolist()
    seen, keys = {}, []
    for b, mt, q0, lq, k0, lk, sub, pack in recs:
        assert (q0, lq, k0, lk) == (int(cu_q[b]), lens_q[b], int(cu_k[b]), lens_k[b])
        seen[(b, mt, sub)] = pack
        keys.append((min(lk, (mt + 1) * tile + lk - lq) + 127) // 128)
    assert seen == want
    assert all(a >= b for a, b in zip(keys, keys[1:])), keys
