"""GPU differential tests (-m gpu): the new kernels against flash-attn's own entry points,
called with exactly the arguments the reference passes (attention.py:229-232 and :242-248).
flash-attn is the arithmetic the reference delegates to (dependency `flash-attn`, 2.8.3 in
this image), so this is the live pin of the attention oracle.  Skipped if flash_attn is not
importable; never part of the product path."""
import pytest
import torch

import oracle
from helpers import make_paged_case

pytestmark = pytest.mark.gpu
SCALE = 128 ** -0.5
ATOL = 2e-2

flash_attn = pytest.importorskip("flash_attn")


@pytest.fixture(scope="module")
def ops():
    import b200attn
    b200attn.load_library()
    return b200attn.ops


def maxerr(a, b):
    return (a.float().cpu() - b.float().cpu()).abs().max().item()


@pytest.mark.parametrize("Hq,Hkv", [(32, 8), (4, 1), (5, 1), (16, 8)])
def test_decode_matches_flash_attn_with_kvcache(ops, Hq, Hkv):
    lens = [4096, 1, 255, 256, 257, 1300, 33, 777]
    c = make_paged_case(5 + Hq, [1] * len(lens), lens, Hq, Hkv, pad=0, table_width=16)
    q, kc, vc = c["q"].cuda(), c["k_cache"].cuda(), c["v_cache"].cuda()
    cl, bt = c["context_lens"].cuda(), c["block_tables"].cuda()
    ref = flash_attn.flash_attn_with_kvcache(q.view(len(lens), 1, Hq, 128), kc, vc, cache_seqlens=cl, block_table=bt,
                                             softmax_scale=SCALE, causal=True).view(-1, Hq, 128)
    got = ops.attn_decode(q, kc, vc, cl, bt, SCALE)
    want = oracle.attn_decode(c["q"], c["k_cache"], c["v_cache"], c["context_lens"], c["block_tables"], SCALE)
    assert maxerr(got, ref) <= ATOL
    assert maxerr(ref, want) <= ATOL            # pins the oracle to flash-attn as well


@pytest.mark.parametrize("Hq,Hkv", [(16, 8), (4, 1), (5, 1)])
def test_paged_varlen_matches_flash_attn_varlen(ops, Hq, Hkv):
    lens_k = [700, 129, 128, 1, 257, 513, 64, 1000]
    lens_q = [188, 1, 128, 1, 1, 257, 64, 40]
    c = make_paged_case(41 + Hq, lens_q, lens_k, Hq, Hkv, pad=-1)
    a = {k: (v.cuda() if isinstance(v, torch.Tensor) else v) for k, v in c.items()}
    ref = flash_attn.flash_attn_varlen_func(a["q"], a["k_cache"], a["v_cache"], cu_seqlens_q=a["cu_seqlens_q"],
                                            cu_seqlens_k=a["cu_seqlens_k"], max_seqlen_q=c["max_seqlen_q"],
                                            max_seqlen_k=c["max_seqlen_k"], softmax_scale=SCALE, causal=True,
                                            block_table=a["block_tables"])
    got = ops.attn_varlen(a["q"], a["k_cache"], a["v_cache"], a["cu_seqlens_q"], a["cu_seqlens_k"],
                          c["max_seqlen_q"], c["max_seqlen_k"], SCALE, block_table=a["block_tables"])
    want = oracle.attn_varlen(c["q"], c["k_cache"], c["v_cache"], c["cu_seqlens_q"], c["cu_seqlens_k"], SCALE,
                              c["block_tables"])
    assert maxerr(got, ref) <= ATOL
    assert maxerr(ref, want) <= ATOL


def test_unpaged_varlen_matches_flash_attn_varlen(ops):
    lens = [128, 1, 63, 300, 7]
    g = torch.Generator().manual_seed(2)
    T, Hq, Hkv, D = sum(lens), 16, 8, 128
    qkv = torch.randn(T, (Hq + 2 * Hkv) * D, generator=g).bfloat16().cuda()
    q = qkv[:, :Hq * D].reshape(T, Hq, D).contiguous()
    k = qkv[:, Hq * D:(Hq + Hkv) * D].reshape(T, Hkv, D).contiguous()
    v = qkv[:, (Hq + Hkv) * D:].view(T, Hkv, D)
    cu = torch.tensor([0] + torch.tensor(lens).cumsum(0).tolist(), dtype=torch.int32).cuda()
    ref = flash_attn.flash_attn_varlen_func(q, k, v, cu_seqlens_q=cu, cu_seqlens_k=cu, max_seqlen_q=max(lens),
                                            max_seqlen_k=max(lens), softmax_scale=SCALE, causal=True, block_table=None)
    got = ops.attn_varlen(q, k, v, cu, cu, max(lens), max(lens), SCALE)
    assert maxerr(got, ref) <= ATOL
