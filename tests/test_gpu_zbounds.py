"""Out-of-bounds evidence without compute-sanitizer (the tool is closed on this GPU pool: "runs under it have left
GPUs needing a reset"), for every kernel behind the C ABI:

  WRITES  every output (and the split-K workspace) lives inside a larger allocation whose guard bands, and whose
          interior elements the call must not touch (unmapped cache slots, rows past a sequence), carry a canary bit
          pattern; after the call the canaries are intact.
  READS   everything outside the logical inputs is hostile: cache rows past a context and unused cache blocks hold
          NaN, block-table entries past the pages a context needs point ~2^30 blocks away (dereferencing one is an
          illegal address, which fails the test loudly), q/k/v views sit inside NaN-filled parents.  The result must
          be finite and BIT-identical to the same call on clean inputs (the kernels are deterministic).

The file sorts after the other GPU tests on purpose: a kernel that does fault poisons the CUDA context.
"""
import pytest
import torch

import oracle
from helpers import make_paged_case

pytestmark = pytest.mark.gpu
SCALE = 128 ** -0.5
CANARY = 0x5A5B                  # as bf16: 2.4e16 -- never produced by the kernels on these inputs
GUARD = 4096                     # elements on either side
FAR_BLOCK = 0x3FFFFFF0           # a block index whose address is ~2^50 bytes past the cache


@pytest.fixture(scope="module")
def ops():
    assert torch.cuda.is_available()
    import b200attn
    b200attn.load_library()
    return b200attn.ops


class Guarded:
    """A bf16 (or raw) tensor of `shape` carved out of a canary-filled parent."""

    def __init__(self, shape, dtype=torch.bfloat16, fill=None):
        n = 1
        for s in shape:
            n *= s
        self.n = n
        self.parent = torch.full((n + 2 * GUARD,), CANARY, dtype=torch.int16, device="cuda")
        self.t = self.parent[GUARD:GUARD + n].view(dtype).view(*shape) if dtype == torch.bfloat16 else None
        if fill is not None:
            self.t.copy_(fill)

    def guards_intact(self):
        p = self.parent
        return bool((p[:GUARD] == CANARY).all().item() and (p[GUARD + self.n:] == CANARY).all().item())

    def untouched(self, mask=None):
        """Interior elements still holding the canary (mask: bool tensor of the interior shape)."""
        same = self.t.view(torch.int16) == CANARY
        return bool(same.all().item()) if mask is None else bool(same[mask].all().item())


def hostile_paged(c, lens_k):
    """Caches with NaN everywhere outside the contexts and far-away table entries past the needed pages."""
    kc, vc = c["k_cache"].clone(), c["v_cache"].clone()
    bt = c["block_tables"].clone()
    used = torch.zeros(kc.shape[0], dtype=torch.bool)
    for b, L in enumerate(lens_k):
        nblk = (L + 255) // 256
        for i in range(nblk):
            blk = int(bt[b, i])
            used[blk] = True
            if i == nblk - 1 and L % 256:
                kc[blk, L % 256:] = float("nan")
                vc[blk, L % 256:] = float("nan")
        bt[b, nblk:] = FAR_BLOCK
    kc[~used] = float("nan")
    vc[~used] = float("nan")
    return kc.cuda(), vc.cuda(), bt.cuda()


def test_scatter_writes_only_the_mapped_slots(ops):
    Hkv, D, nblocks = 4, 128, 3
    g = torch.Generator().manual_seed(3)
    N = 300
    perm = torch.randperm(nblocks * 256, generator=g)
    perm = perm[(perm != 0) & (perm != nblocks * 256 - 1)]
    slots = torch.cat([torch.tensor([0, nblocks * 256 - 1]), perm[:N - 2]]).to(torch.int32)     # first and last slot of the cache, no duplicates
    slots[5:9] = -1                                                 # padding rows: skipped
    parent = torch.full((N + 2, 3 * Hkv * D), float("nan")).bfloat16().cuda()
    qkv = parent[1:N + 1]
    qkv.copy_(torch.randn(N, 3 * Hkv * D, generator=g).bfloat16())
    key = qkv[:, Hkv * D:2 * Hkv * D].view(N, Hkv, D)
    val = qkv[:, 2 * Hkv * D:].view(N, Hkv, D)
    kc, vc = Guarded((nblocks, 256, Hkv, D)), Guarded((nblocks, 256, Hkv, D))
    ops.store_kvcache(key, val, kc.t, vc.t, slots.cuda())
    torch.cuda.synchronize()
    hit = torch.zeros(nblocks * 256, dtype=torch.bool, device="cuda")
    hit[slots[slots >= 0].long().cuda()] = True
    for cache, src in ((kc, key), (vc, val)):
        flat = cache.t.view(nblocks * 256, Hkv, D)
        assert cache.guards_intact()
        assert bool((flat[~hit].view(torch.int16) == CANARY).all().item())           # unmapped slots untouched
        live = slots >= 0
        assert torch.equal(flat[slots[live].long().cuda()], src[live.cuda()])


@pytest.mark.parametrize("Hq,Hkv,splits", [(32, 8, None), (16, 1, 5), (5, 1, 3), (8, 8, 1)])
def test_decode_reads_nothing_outside_contexts_and_writes_only_out(ops, Hq, Hkv, splits):
    import b200attn
    lib = b200attn.load_library()
    lens = [700, 1, 0, 256, 257, 1023]
    B = len(lens)
    c = make_paged_case(40 + Hq, [1] * B, lens, Hq, Hkv, pad=0, table_width=6)
    clean = ops.attn_decode(c["q"].cuda(), c["k_cache"].cuda(), c["v_cache"].cuda(), c["context_lens"].cuda(),
                            c["block_tables"].cuda(), SCALE, num_splits=splits)
    kc, vc, bt = hostile_paged(c, lens)
    qparent = torch.full((B + 2, Hq + 1, 128), float("nan")).bfloat16().cuda()
    q = qparent[1:B + 1, :Hq]                                       # strided rows inside a NaN parent
    q.copy_(c["q"])
    out = Guarded((B, Hq, 128))
    ns = splits if splits is not None else lib.b200_decode_plan(B, Hkv, 6 * 256)
    need = lib.b200_decode_workspace_bytes(B, Hq, ns)
    ws = torch.zeros(need + 2 * GUARD, dtype=torch.uint8, device="cuda")
    ws[:GUARD] = 0x5A
    ws[GUARD + need:] = 0x5A
    wsp = ws[GUARD:GUARD + need]
    from b200attn._lib import check
    check(lib.b200_attn_decode(q.data_ptr(), q.stride(0), kc.data_ptr(), vc.data_ptr(), out.t.data_ptr(), out.t.stride(0),
                               c["context_lens"].cuda().data_ptr(), bt.data_ptr(), bt.stride(0), B, Hq, Hkv, 128, 256,
                               float(SCALE), ns, wsp.data_ptr() if ns > 1 else None, need if ns > 1 else 0,
                               torch.cuda.current_stream().cuda_stream), "b200_attn_decode")
    torch.cuda.synchronize()
    assert out.guards_intact()
    assert bool((ws[:GUARD] == 0x5A).all().item() and (ws[GUARD + need:] == 0x5A).all().item())
    assert torch.isfinite(out.t.float()).all()
    assert torch.equal(out.t, clean)
    want = oracle.attn_decode(c["q"], c["k_cache"], c["v_cache"], c["context_lens"], c["block_tables"], SCALE)
    assert (out.t.float().cpu() - want.float()).abs().max().item() <= 2e-2


@pytest.mark.parametrize("Hq,Hkv", [(16, 4), (5, 1), (8, 1)])
def test_planned_prefill_paged_hostile_cache_and_guarded_out(ops, Hq, Hkv):
    """Mixed step (chunked-prefill rows with a prefix, fresh prompts, decode rows): partial query tiles (rows past
    a sequence are computed by the tile and must be dropped), partial KV tiles (NaN past the context)."""
    lens_q = [1, 127, 129, 300, 1, 64]
    lens_k = [513, 127, 400, 300, 1, 1000]
    c = make_paged_case(70 + Hq, lens_q, lens_k, Hq, Hkv, pad=0, table_width=5)
    T = sum(lens_q)
    args = (c["cu_seqlens_q"].cuda(), c["cu_seqlens_k"].cuda(), max(lens_q), max(lens_k), SCALE)
    clean = ops.attn_varlen_planned(c["q"].cuda(), c["k_cache"].cuda(), c["v_cache"].cuda(), *args,
                                    block_table=c["block_tables"].cuda())
    kc, vc, bt = hostile_paged(c, lens_k)
    qparent = torch.full((T + 2, Hq + 1, 128), float("nan")).bfloat16().cuda()
    q = qparent[1:T + 1, :Hq]
    q.copy_(c["q"])
    out = Guarded((T, Hq, 128))
    ops.attn_varlen_planned(q, kc, vc, *args, block_table=bt, out=out.t)
    torch.cuda.synchronize()
    assert out.guards_intact() and torch.isfinite(out.t.float()).all()
    assert torch.equal(out.t, clean)
    want = oracle.attn_varlen(c["q"], c["k_cache"], c["v_cache"], c["cu_seqlens_q"], c["cu_seqlens_k"], SCALE,
                              block_tables=c["block_tables"])
    assert (out.t.float().cpu() - want.float()).abs().max().item() <= 2e-2


@pytest.mark.parametrize("Hq,Hkv", [(16, 4), (5, 1)])
def test_planned_prefill_unpaged_views_inside_nan_parent(ops, Hq, Hkv):
    """q/k/v are column slices of one qkv row (qwen3.py:185-190): the neighbours of every row are hostile."""
    lens = [1, 130, 255, 256, 77]
    T = sum(lens)
    g = torch.Generator().manual_seed(11 + Hq)
    W = (Hq + 2 * Hkv) * 128
    parent = torch.full((T + 2, W + 128), float("nan")).bfloat16().cuda()
    qkv = parent[1:T + 1, 64:64 + W]
    qkv.copy_(torch.randn(T, W, generator=g).bfloat16())
    q = qkv[:, :Hq * 128].view(T, Hq, 128)
    k = qkv[:, Hq * 128:(Hq + Hkv) * 128].view(T, Hkv, 128)
    v = qkv[:, (Hq + Hkv) * 128:].view(T, Hkv, 128)
    cu = torch.tensor([0] + [sum(lens[:i + 1]) for i in range(len(lens))], dtype=torch.int32)
    clean = ops.attn_varlen_planned(q.contiguous(), k.contiguous(), v.contiguous(), cu.cuda(), cu.cuda(), max(lens), max(lens), SCALE)
    out = Guarded((T, Hq, 128))
    ops.attn_varlen_planned(q, k, v, cu.cuda(), cu.cuda(), max(lens), max(lens), SCALE, out=out.t)
    torch.cuda.synchronize()
    assert out.guards_intact() and torch.isfinite(out.t.float()).all()
    assert torch.equal(out.t, clean)
    want = oracle.attn_varlen(q.cpu().contiguous(), k.cpu().contiguous(), v.cpu().contiguous(), cu, cu, SCALE)
    assert (out.t.float().cpu() - want.float()).abs().max().item() <= 2e-2


def _stream():
    return torch.cuda.current_stream().cuda_stream


@pytest.mark.parametrize("T", [1, 37, 1000])
def test_pointwise_kernels_write_only_their_outputs(ops, T):
    import b200attn
    from b200attn._lib import check
    lib = b200attn.load_library()
    g = torch.Generator().manual_seed(T)
    Hq, Hkv, D, hidden, inter = 16, 4, 128, 1024, 1536
    # ---- rmsnorm / add+rmsnorm: out and the in-place residual
    x = torch.randn(T, hidden, generator=g).bfloat16().cuda()
    w = (1 + 0.1 * torch.randn(hidden, generator=g)).bfloat16().cuda()
    r0 = torch.randn(T, hidden, generator=g).bfloat16().cuda()
    want_y, want_r = ops.rms_norm(x, w, 1e-6, r0.clone())
    out, res = Guarded((T, hidden)), Guarded((T, hidden), fill=r0)
    check(lib.b200_rmsnorm(x.data_ptr(), res.t.data_ptr(), w.data_ptr(), 1e-6, out.t.data_ptr(), T, hidden, _stream()), "b200_rmsnorm")
    assert out.guards_intact() and res.guards_intact() and torch.equal(out.t, want_y) and torch.equal(res.t, want_r)
    # ---- silu_and_mul
    gu = torch.randn(T, 2 * inter, generator=g).bfloat16().cuda()
    want = ops.silu_and_mul(gu)
    out = Guarded((T, inter))
    check(lib.b200_silu_and_mul(gu.data_ptr(), out.t.data_ptr(), T, inter, _stream()), "b200_silu_and_mul")
    assert out.guards_intact() and torch.equal(out.t, want)
    # ---- q/k norm + rope (+ store): q', k', and only the mapped cache slots
    W = (Hq + 2 * Hkv) * D
    parent = torch.full((T + 2, W + 128), float("nan")).bfloat16().cuda()
    qkv = parent[1:T + 1, 64:64 + W]
    qkv.copy_(torch.randn(T, W, generator=g).bfloat16())
    q = qkv[:, :Hq * D].view(T, Hq, D)
    k = qkv[:, Hq * D:(Hq + Hkv) * D].view(T, Hkv, D)
    v = qkv[:, (Hq + Hkv) * D:].view(T, Hkv, D)
    pos = torch.randint(0, 4096, (T,), generator=g).cuda()
    inv = 1.0 / (1e6 ** (torch.arange(0, D, 2).float() / D))
    fr = torch.arange(4096).float()[:, None] * inv[None]
    cache = torch.cat([fr.cos(), fr.sin()], -1).contiguous().cuda()
    qw = (1 + 0.1 * torch.randn(D, generator=g)).bfloat16().cuda()
    kw = (1 + 0.1 * torch.randn(D, generator=g)).bfloat16().cuda()
    want_q, want_k = ops.qk_norm_rope(q, k, pos, qw, kw, cache, 1e-6)
    qo, ko = Guarded((T, Hq, D)), Guarded((T, Hkv, D))
    check(lib.b200_qk_norm_rope(q.data_ptr(), q.stride(0), k.data_ptr(), k.stride(0), pos.data_ptr(), cache.data_ptr(),
                                qw.data_ptr(), kw.data_ptr(), 1e-6, qo.t.data_ptr(), ko.t.data_ptr(), T, Hq, Hkv, D, _stream()),
          "b200_qk_norm_rope")
    assert qo.guards_intact() and ko.guards_intact() and torch.equal(qo.t, want_q) and torch.equal(ko.t, want_k)
    nblocks = (T + 255) // 256 + 1
    slots = torch.randperm(nblocks * 256, generator=g)[:T].to(torch.int32)
    if T > 4:
        slots[2] = -1
    kc, vc = Guarded((nblocks, 256, Hkv, D)), Guarded((nblocks, 256, Hkv, D))
    qo, ko = Guarded((T, Hq, D)), Guarded((T, Hkv, D))
    check(lib.b200_qk_norm_rope_store(q.data_ptr(), q.stride(0), k.data_ptr(), k.stride(0), v.data_ptr(), v.stride(0),
                                      pos.data_ptr(), cache.data_ptr(), qw.data_ptr(), kw.data_ptr(), 1e-6,
                                      qo.t.data_ptr(), ko.t.data_ptr(), kc.t.data_ptr(), vc.t.data_ptr(),
                                      slots.cuda().data_ptr(), T, Hq, Hkv, D, _stream()), "b200_qk_norm_rope_store")
    torch.cuda.synchronize()
    hit = torch.zeros(nblocks * 256, dtype=torch.bool, device="cuda")
    live = (slots >= 0).cuda()
    hit[slots.cuda()[live].long()] = True
    for cch, src in ((kc, want_k), (vc, v)):
        flat = cch.t.view(nblocks * 256, Hkv, D)
        assert cch.guards_intact() and bool((flat[~hit].view(torch.int16) == CANARY).all().item())
        assert torch.equal(flat[slots.cuda()[live].long()], src[live])
    assert qo.guards_intact() and ko.guards_intact() and torch.equal(qo.t, want_q) and torch.equal(ko.t, want_k)


def test_single_rank_collectives_write_only_their_outputs(ops):
    nv = ops.NvlsAllReduce(None, 1 << 20)
    g = torch.Generator().manual_seed(2)
    T, H = 33, 2048
    x = torch.randn(T, H, generator=g).bfloat16().cuda()
    w = (1 + 0.1 * torch.randn(H, generator=g)).bfloat16().cuda()
    r0 = torch.randn(T, H, generator=g).bfloat16().cuda()
    want_y, want_r = ops.rms_norm(x, w, 1e-6, r0.clone())
    for form in (0, 1, 2, 3):
        out, res = Guarded((T, H)), Guarded((T, H), fill=r0)
        nv.all_reduce_add_rms_norm(x, res.t, w, 1e-6, form=form, out=out.t)
        torch.cuda.synchronize()
        assert out.guards_intact() and res.guards_intact() and torch.equal(res.t, want_r), form
        assert (out.t == want_y).float().mean().item() > 0.99, form
    assert nv.status() == 0
    nv.destroy()
