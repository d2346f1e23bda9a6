"""CPU tests (-m "not gpu"): the oracle against the reference-generated golden vectors and
against first-principles dense attention.  Nothing here touches the CUDA library."""
import glob
import math
import os

import pytest
import torch

import oracle
from helpers import GOLDEN, make_forward_case, make_paged_case, scatter_inputs, tensor_sha256


# ----------------------------------------------------------------- scatter (pinned, bit-exact)
@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLDEN, "scatter_*.pt"))))
def test_scatter_oracle_matches_reference_triton_kernel(path):
    meta = torch.load(path)
    _, k, v, k_cache, v_cache = scatter_inputs(meta)
    k0, v0 = k_cache.clone(), v_cache.clone()
    oracle.kv_scatter(k, v, k_cache, v_cache, meta["slot_mapping"])
    row = meta["Hkv"] * meta["D"]
    touched = meta["touched"]
    assert torch.equal(k_cache.view(-1, row)[touched], meta["k_rows"])
    assert torch.equal(v_cache.view(-1, row)[touched], meta["v_rows"])
    assert meta["k_untouched_equal"] and meta["v_untouched_equal"]
    mask = torch.ones(k_cache.shape[0] * 256, dtype=torch.bool)
    mask[touched] = False
    assert torch.equal(k_cache.view(-1, row)[mask], k0.view(-1, row)[mask])
    assert torch.equal(v_cache.view(-1, row)[mask], v0.view(-1, row)[mask])


def test_scatter_all_padding_is_noop():
    k = torch.randn(4, 2, 128).bfloat16()
    kc = torch.zeros(1, 256, 2, 128).bfloat16()
    vc = torch.zeros_like(kc)
    oracle.kv_scatter(k, k, kc, vc, torch.full((4,), -1, dtype=torch.int32))
    assert kc.abs().sum() == 0 and vc.abs().sum() == 0


# ----------------------------------------------------------------- attention (pinned on the GPU box)
@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLDEN, "attn_ref_*.pt"))))
def test_attention_oracle_matches_reference_forward_on_gpu(path):
    """Fixtures hold the output of the UNMODIFIED reference `Attention.forward` (Triton scatter +
    flash-attn 2.8.3) run on a B200 by tools/make_golden_gpu.py; inputs are re-created from the seed."""
    m = torch.load(path)
    c = make_forward_case(m)
    o = oracle.attention_forward(c["q"], c["k"], c["v"], c["k_cache"], c["v_cache"], oracle.Context(*c["ctx"]), 128 ** -0.5)
    n = c["num_real"]
    assert (o[:n].float() - m["out"][:n].float()).abs().max().item() <= 2e-2      # bf16 tolerance of north_star
    # the scatter half of the step is bit-exact: digest of the whole cache after the call
    assert tensor_sha256(c["k_cache"]) == m["k_cache_sha256"]
    assert tensor_sha256(c["v_cache"]) == m["v_cache_sha256"]


# ----------------------------------------------------------------- pointwise (pinned, eager form)
@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLDEN, "pointwise_*.pt"))))
def test_pointwise_oracle_matches_reference_modules(path):
    m = torch.load(path)
    T, Hq, Hkv, D = m["T"], m["Hq"], m["Hkv"], m["D"]
    q, k, v = m["qkv"].split([Hq * D, Hkv * D, Hkv * D], dim=-1)
    q, k = q.view(T, Hq, D), k.view(T, Hkv, D)
    cache = oracle.make_cos_sin_cache(D, m["max_pos"], m["base"])
    assert torch.equal(cache[m["cos_sin_cache_sample_pos"]], m["cos_sin_cache_sample"])
    assert torch.equal(oracle.rms_norm(q, m["q_weight"], m["eps"]), m["q_normed"])
    assert torch.equal(oracle.rms_norm(k, m["k_weight"], m["eps"]), m["k_normed"])
    qo, ko = oracle.qk_norm_rope(q, k, m["positions"], m["q_weight"], m["k_weight"], cache, m["eps"])
    assert torch.equal(qo, m["q_out"]) and torch.equal(ko, m["k_out"])
    assert torch.equal(oracle.silu_and_mul(m["act_in"], mode="eager"), m["act_out_eager"])
    # compiled form differs from eager by at most one bf16 rounding
    c = oracle.silu_and_mul(m["act_in"], mode="compiled").float()
    e = m["act_out_eager"].float()
    assert ((c - e).abs() <= 2.0 ** -7 * e.abs() + 1e-30).all()
    assert torch.equal(oracle.rms_norm(m["h"], m["h_weight"], m["eps"]), m["h_norm"])
    y, r = oracle.add_rms_norm(m["h"], m["r"], m["h_weight"], m["eps"])
    assert torch.equal(y, m["hr_norm"]) and torch.equal(r, m["hr_res"])


# ----------------------------------------------------------------- attention vs dense math
def _dense(q, k, v, scale, offset):
    """Independent formulation: explicit loops over heads with a [Lq, Lk] mask."""
    Lq, Hq, D = q.shape
    Lk, Hkv, _ = k.shape
    G = Hq // Hkv
    out = torch.zeros(Lq, Hq, D, dtype=torch.float64)
    for h in range(Hq):
        s = (q[:, h].double() @ k[:, h // G].double().T) * scale
        for i in range(Lq):
            s[i, i + offset + 1:] = -math.inf
        p = torch.softmax(s, -1)
        out[:, h] = p @ v[:, h // G].double()
    return out


@pytest.mark.parametrize("Hq,Hkv", [(16, 8), (4, 1), (5, 1), (8, 1), (2, 1)])
def test_paged_equals_unpaged_and_dense(Hq, Hkv):
    lens_k = [1, 37, 256, 300, 513]
    lens_q = [1, 5, 256, 1, 200]
    c = make_paged_case(7, lens_q, lens_k, Hq, Hkv)
    scale = 128 ** -0.5
    o_paged = oracle.attn_varlen(c["q"], c["k_cache"], c["v_cache"], c["cu_seqlens_q"],
                                 c["cu_seqlens_k"], scale, c["block_tables"])
    # chunked prefill on packed K/V needs per-seq Lk != Lq, so compare per sequence with dense math
    cq, ck = c["cu_seqlens_q"].tolist(), c["cu_seqlens_k"].tolist()
    for b in range(len(lens_k)):
        q = c["q"][cq[b]:cq[b + 1]]
        k = c["k_packed"][ck[b]:ck[b + 1]]
        v = c["v_packed"][ck[b]:ck[b + 1]]
        ref = _dense(q, k, v, scale, lens_k[b] - lens_q[b])
        got = o_paged[cq[b]:cq[b + 1]].double()
        assert (got - ref).abs().max() < 2e-2


def test_unpaged_prefill_equals_paged_when_lq_eq_lk():
    lens = [128, 3, 257]
    c = make_paged_case(3, lens, lens, 16, 8)
    scale = 128 ** -0.5
    a = oracle.attn_varlen(c["q"], c["k_packed"], c["v_packed"], c["cu_seqlens_q"], c["cu_seqlens_k"], scale)
    b = oracle.attn_varlen(c["q"], c["k_cache"], c["v_cache"], c["cu_seqlens_q"], c["cu_seqlens_k"], scale,
                           c["block_tables"])
    assert torch.equal(a, b)


def test_decode_equals_varlen_with_single_query_rows_and_padding():
    lens_k = [4096 // 8, 1, 0, 300]
    c = make_paged_case(5, [1, 1, 1, 1], lens_k, 8, 2, pad=0, table_width=6)
    scale = 128 ** -0.5
    d = oracle.attn_decode(c["q"], c["k_cache"], c["v_cache"], c["context_lens"], c["block_tables"], scale)
    assert d[2].abs().sum() == 0                       # graph padding row -> zeros
    keep = [0, 1, 3]
    c2 = make_paged_case(5, [1, 1, 1, 1], lens_k, 8, 2, pad=0, table_width=6)
    vl = oracle.attn_varlen(c2["q"], c2["k_cache"], c2["v_cache"], c2["cu_seqlens_q"],
                            c2["cu_seqlens_k"], scale, c2["block_tables"])
    assert torch.equal(d[keep], vl[keep])


def test_config0_single_layer_prefill_then_decode():
    """BASELINE.json configs[0]: Qwen3-0.6B attention shapes, 1 seq, 128-token prefill + 32
    decode steps through `attention_forward` (scatter + read), block_table [2] of a 4-block
    cache; the incremental result must equal one causal pass over all 160 tokens."""
    torch.manual_seed(0)
    Hq, Hkv, D, P = 16, 8, 128, 256
    scale = D ** -0.5
    total = 160
    q = torch.randn(total, Hq, D).bfloat16()
    k = torch.randn(total, Hkv, D).bfloat16()
    v = torch.randn(total, Hkv, D).bfloat16()
    kc = torch.zeros(4, P, Hkv, D).bfloat16()
    vc = torch.zeros_like(kc)
    bt = torch.tensor([[2]], dtype=torch.int32)
    outs = []
    ctx = oracle.Context(torch.tensor([0, 128], dtype=torch.int32), torch.tensor([0, 128], dtype=torch.int32),
                         128, 128, torch.arange(2 * P, 2 * P + 128, dtype=torch.int32),
                         torch.tensor([128], dtype=torch.int32), None, None)
    outs.append(oracle.attention_forward(q[:128], k[:128], v[:128], kc, vc, ctx, scale))
    for t in range(128, total):
        ctx = oracle.Context(None, None, 0, 0, torch.tensor([2 * P + t], dtype=torch.int32),
                             torch.tensor([t + 1], dtype=torch.int32), bt, None)
        outs.append(oracle.attention_forward(q[t:t + 1], k[t:t + 1], v[t:t + 1], kc, vc, ctx, scale))
    inc = torch.cat(outs)
    full = oracle.attn_varlen(q, k, v, torch.tensor([0, total], dtype=torch.int32),
                              torch.tensor([0, total], dtype=torch.int32), scale)
    assert torch.equal(kc[2, :total], k) and kc[[0, 1, 3]].abs().sum() == 0
    assert (inc.float() - full.float()).abs().max() <= 2 ** -8   # same fp32 math, different blocking


def test_causal_pairs_counts():
    assert oracle.causal_pairs(32832, 32832) == 538_986_528
    assert oracle.causal_pairs(32832, 32832) + 63 * oracle.causal_pairs(64, 32832) == 671_238_144
