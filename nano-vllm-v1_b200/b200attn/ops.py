"""Tensor-level wrappers: validate like the reference does, pass raw pointers to the C ABI.

PyTorch is plumbing only here (device memory, the current stream, output allocation).
Error conventions follow the reference: layout problems are `AssertionError`s
(nanovllm/layers/attention.py:121-124), wrong dtype/device `RuntimeError`s (flash-attn's
TORCH_CHECKs).
"""
from typing import Optional, Tuple

import torch

from ._lib import check, load_library

PAGE = 256
HEAD_DIM = 128


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _need_cuda_bf16(t: torch.Tensor, name: str) -> None:
    if not t.is_cuda:
        raise RuntimeError(f"b200attn: {name} must be a CUDA tensor (no CPU fallback on this path)")
    if t.dtype != torch.bfloat16:
        raise RuntimeError(f"b200attn: {name} must be bfloat16, got {t.dtype}")


def _need_i32(t: torch.Tensor, name: str) -> None:
    if not t.is_cuda or t.dtype != torch.int32:
        raise RuntimeError(f"b200attn: {name} must be a CUDA int32 tensor")
    assert t.is_contiguous() or t.dim() == 2 and t.stride(1) == 1, f"{name} must be contiguous"


# ------------------------------------------------------------------------------------ (1)
def store_kvcache(key: torch.Tensor, value: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor,
                  slot_mapping: torch.Tensor) -> None:
    """Same contract as nanovllm.layers.attention.store_kvcache (attention.py:93-127)."""
    N, num_heads, head_dim = key.shape
    D = num_heads * head_dim
    assert key.stride(-1) == 1 and value.stride(-1) == 1
    assert key.stride(1) == head_dim and value.stride(1) == head_dim
    assert k_cache.stride(1) == D and v_cache.stride(1) == D
    assert slot_mapping.numel() == N
    for t, n in ((key, "key"), (value, "value"), (k_cache, "k_cache"), (v_cache, "v_cache")):
        _need_cuda_bf16(t, n)
    _need_i32(slot_mapping, "slot_mapping")
    lib = load_library()
    check(lib.b200_kv_scatter(key.data_ptr(), value.data_ptr(), key.stride(0), value.stride(0),
                              k_cache.data_ptr(), v_cache.data_ptr(), slot_mapping.data_ptr(),
                              N, num_heads, head_dim, _stream()), "b200_kv_scatter")


# ------------------------------------------------------------------------------------ (3)
_WS_CACHE = {}
_WS_RETIRED = []        # superseded workspaces: CUDA graphs captured earlier still hold their raw pointers


def _device_key(device) -> int:
    device = torch.device(device)
    return device.index if device.index is not None else torch.cuda.current_device()


def _grow_workspace(key: int, nbytes: int, device) -> torch.Tensor:
    """Replace the per-device workspace by a larger one.  The old buffer is RETIRED, never freed: the
    reference's capture_cudagraph (model_runner.py:691-705) warms every bucket eagerly from the largest
    batch downward and the split count is not monotone in the batch, so a later warm-up may need more
    bytes than a graph captured before it -- that graph keeps replaying into the old pointer."""
    old = _WS_CACHE.get(key)
    if old is not None:
        _WS_RETIRED.append(old)
    # zeros: the head of the workspace holds the arrival tickets of the fused split merge, which every
    # launch expects to find at zero and leaves at zero (include/b200attn.h, b200_attn_decode)
    ws = torch.zeros(nbytes, dtype=torch.uint8, device=device)
    _WS_CACHE[key] = ws
    return ws


def _workspace_for(batch: int, num_heads: int, splits: int, device) -> Optional[torch.Tensor]:
    if splits <= 1:
        return None
    key = _device_key(device)
    ws = _WS_CACHE.get(key)
    need = load_library().b200_decode_workspace_bytes(batch, num_heads, splits)
    if ws is None or ws.numel() < need:
        if torch.cuda.is_current_stream_capturing():
            raise RuntimeError("b200attn: decode workspace must be sized before CUDA-graph capture "
                               "(call ops.reserve_decode_workspace first)")
        ws = _grow_workspace(key, int(need * 1.25) + 256, device)
    return ws


def reserve_decode_workspace(max_batch: int, num_heads: int, num_kv_heads: int, max_context: int, device) -> None:
    """Size the per-device split-K workspace for the worst graph bucket (call before capture)."""
    lib = load_library()
    worst = 0
    cands = sorted({1, 2, 4, 8, *range(16, max_batch + 1, 16), max_batch})
    for bs in cands:
        s = lib.b200_decode_plan(bs, num_kv_heads, max_context)
        worst = max(worst, lib.b200_decode_workspace_bytes(bs, num_heads, s))
    key = _device_key(device)
    cur = _WS_CACHE.get(key)
    if cur is None or cur.numel() < worst:
        _grow_workspace(key, worst + 256, device)


def attn_decode(q: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor, context_lens: torch.Tensor,
                block_tables: torch.Tensor, scale: float, num_splits: Optional[int] = None,
                out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """q [B, Hq, D] -> [B, Hq, D].  Semantics of flash_attn_with_kvcache at attention.py:229-232."""
    B, Hq, D = q.shape
    Hkv = k_cache.shape[2]
    for t, n in ((q, "q"), (k_cache, "k_cache"), (v_cache, "v_cache")):
        _need_cuda_bf16(t, n)
    _need_i32(context_lens, "context_lens")
    _need_i32(block_tables, "block_tables")
    assert q.stride(-1) == 1 and q.stride(1) == D
    assert k_cache.is_contiguous() and v_cache.is_contiguous() and k_cache.shape[1] == PAGE
    assert block_tables.shape[0] == B and context_lens.numel() == B
    lib = load_library()
    if num_splits is None:
        num_splits = lib.b200_decode_plan(B, Hkv, block_tables.shape[1] * PAGE)
    ws = _workspace_for(B, Hq, num_splits, q.device)
    if out is None:
        out = torch.empty((B, Hq, D), dtype=q.dtype, device=q.device)
    check(lib.b200_attn_decode(q.data_ptr(), q.stride(0), k_cache.data_ptr(), v_cache.data_ptr(),
                               out.data_ptr(), out.stride(0), context_lens.data_ptr(),
                               block_tables.data_ptr(), block_tables.stride(0), B, Hq, Hkv, D, PAGE,
                               float(scale), num_splits, ws.data_ptr() if ws is not None else None,
                               ws.numel() if ws is not None else 0, _stream()), "b200_attn_decode")
    return out


# ------------------------------------------------------------------------------------ (2)
def attn_varlen(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, cu_seqlens_q: torch.Tensor,
                cu_seqlens_k: torch.Tensor, max_seqlen_q: int, max_seqlen_k: int, scale: float,
                block_table: Optional[torch.Tensor] = None, variant: int = 0,
                out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Semantics of flash_attn_varlen_func(..., causal=True, block_table=...) at attention.py:242-248."""
    T, Hq, D = q.shape
    for t, n in ((q, "q"), (k, "k"), (v, "v")):
        _need_cuda_bf16(t, n)
    _need_i32(cu_seqlens_q, "cu_seqlens_q")
    _need_i32(cu_seqlens_k, "cu_seqlens_k")
    assert q.stride(-1) == 1 and q.stride(1) == D
    B = cu_seqlens_q.numel() - 1
    assert cu_seqlens_k.numel() == B + 1
    if block_table is not None:
        _need_i32(block_table, "block_table")
        assert k.dim() == 4 and k.is_contiguous() and v.is_contiguous() and k.shape[1] == PAGE
        Hkv = k.shape[2]
        ks = vs = 0
        nblocks = k.shape[0]
        bt_ptr, bt_stride = block_table.data_ptr(), block_table.stride(0)
        assert block_table.shape[0] == B
    else:
        assert k.dim() == 3 and k.stride(-1) == 1 and v.stride(-1) == 1
        assert k.stride(1) == D and v.stride(1) == D
        Hkv = k.shape[1]
        ks, vs = k.stride(0), v.stride(0)
        nblocks = 0
        bt_ptr, bt_stride = None, 0
    if out is None:
        out = torch.empty((T, Hq, D), dtype=q.dtype, device=q.device)
    lib = load_library()
    check(lib.b200_attn_prefill(q.data_ptr(), q.stride(0), k.data_ptr(), v.data_ptr(), ks, vs,
                                out.data_ptr(), out.stride(0), cu_seqlens_q.data_ptr(), cu_seqlens_k.data_ptr(),
                                bt_ptr, bt_stride, B, T, int(max_seqlen_q), int(max_seqlen_k), Hq, Hkv, D, PAGE,
                                nblocks, float(scale), variant, _stream()), "b200_attn_prefill")
    return out


# ------------------------------------------------------------------------------------ (2'/3')
class StepPlan:
    """Device-built routing of one varlen step (b200_attn_plan): work list of the persistent
    prefill kernel + list of single-token rows for the decode kernel.  Built once per engine step
    and shared by every layer (the Context it is derived from is the same for all layers)."""

    def __init__(self, buf: torch.Tensor, batch: int, total_q: int, route_decode: bool, heads) -> None:
        self.buf, self.batch, self.total_q, self.route_decode, self.heads = buf, batch, total_q, route_decode, heads

    def counts(self):
        """(prefill work items = query tiles of multi-token rows, decode rows) -- synchronises; for tests and tools only.
        plan words: [0] work units, [1] decode rows, [2] tile rows, [3] items."""
        head = self.buf[:16].view(torch.int32).cpu()
        return int(head[3]), int(head[1])


_PLAN_RING = {}          # device index -> [buffers, next slot]: the engine's per-step plans alternate between two buffers


def _plan_buffer(nbytes: int, device: torch.device) -> torch.Tensor:
    """A plan buffer that lives until the plan after next is built (the engine builds one plan per step, uses it for
    that step's layers only, and everything is ordered on one stream): no allocation on the step path once the largest
    step has been seen (the ABI side never allocates; this keeps the host side from doing it per step)."""
    ring = _PLAN_RING.setdefault(device.index, [[None, None], 0])
    slot = ring[1]
    ring[1] = slot ^ 1
    buf = ring[0][slot]
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(max(nbytes, 1 << 20), dtype=torch.uint8, device=device)
        ring[0][slot] = buf
    return buf[:nbytes]


def attn_plan(cu_seqlens_q: torch.Tensor, cu_seqlens_k: torch.Tensor, total_q: int, num_heads: int, num_kv_heads: int,
              route_decode: bool, reuse: bool = False) -> StepPlan:
    """reuse=True: the buffer comes from a two-deep ring (valid until the plan after next is built) -- what the
    Attention layer asks for; the default allocates, so a caller may keep any number of plans alive."""
    _need_i32(cu_seqlens_q, "cu_seqlens_q")
    _need_i32(cu_seqlens_k, "cu_seqlens_k")
    B = cu_seqlens_q.numel() - 1
    lib = load_library()
    nbytes = lib.b200_attn_plan_bytes(max(B, 1), total_q)
    buf = _plan_buffer(nbytes, cu_seqlens_q.device) if reuse else torch.empty(nbytes, dtype=torch.uint8, device=cu_seqlens_q.device)
    check(lib.b200_attn_plan(cu_seqlens_q.data_ptr(), cu_seqlens_k.data_ptr(), B, total_q, num_heads, num_kv_heads,
                             1 if route_decode else 0, buf.data_ptr(), nbytes, _stream()), "b200_attn_plan")
    return StepPlan(buf, B, total_q, route_decode, (num_heads, num_kv_heads))


def attn_varlen_planned(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, cu_seqlens_q: torch.Tensor,
                        cu_seqlens_k: torch.Tensor, max_seqlen_q: int, max_seqlen_k: int, scale: float,
                        block_table: Optional[torch.Tensor] = None, plan: Optional[StepPlan] = None,
                        out: Optional[torch.Tensor] = None, num_splits: Optional[int] = None) -> torch.Tensor:
    """flash_attn_varlen_func(..., causal=True, block_table=...) semantics (attention.py:242-248) through
    the step plan: multi-token rows on the persistent tcgen05 kernel, single-token rows (paged calls
    only) on the split-K decode kernel."""
    T, Hq, D = q.shape
    for t, n in ((q, "q"), (k, "k"), (v, "v")):
        _need_cuda_bf16(t, n)
    assert q.stride(-1) == 1 and q.stride(1) == D
    B = cu_seqlens_q.numel() - 1
    paged = block_table is not None
    if paged:
        _need_i32(block_table, "block_table")
        assert k.dim() == 4 and k.is_contiguous() and v.is_contiguous() and k.shape[1] == PAGE
        assert block_table.shape[0] == B
        Hkv, ks, vs, nblocks = k.shape[2], 0, 0, k.shape[0]
        bt_ptr, bt_stride = block_table.data_ptr(), block_table.stride(0)
    else:
        assert k.dim() == 3 and k.stride(-1) == 1 and v.stride(-1) == 1 and k.stride(1) == D and v.stride(1) == D
        Hkv, ks, vs, nblocks = k.shape[1], k.stride(0), v.stride(0), 0
        bt_ptr, bt_stride = None, 0
    if plan is None:
        plan = attn_plan(cu_seqlens_q, cu_seqlens_k, T, Hq, Hkv, route_decode=paged)
    assert plan.batch == B and plan.total_q == T and (not plan.route_decode or paged) and plan.heads == (Hq, Hkv)
    if out is None:
        out = torch.empty((T, Hq, D), dtype=q.dtype, device=q.device)
    lib = load_library()
    # The prefill kernel runs whenever the plan may hold work for it: always for an unpaged call (its
    # 1-token rows are NOT routed to decode -- a step of 1-token prompts without a prefix hit must still
    # produce out = v, as flash_attn_varlen_func does), and for a paged call with any multi-token row.
    if max_seqlen_q > 1 or not plan.route_decode:
        check(lib.b200_attn_prefill_planned(q.data_ptr(), q.stride(0), k.data_ptr(), v.data_ptr(), ks, vs,
                                            out.data_ptr(), out.stride(0), cu_seqlens_q.data_ptr(), cu_seqlens_k.data_ptr(),
                                            bt_ptr, bt_stride, B, T, int(max_seqlen_q), int(max_seqlen_k), Hq, Hkv, D, PAGE,
                                            nblocks, float(scale), plan.buf.data_ptr(), plan.buf.numel(), _stream()),
              "b200_attn_prefill_planned")
    if plan.route_decode:
        if num_splits is None:
            num_splits = lib.b200_decode_plan(B, Hkv, int(max_seqlen_k))
        ws = _workspace_for(B, Hq, num_splits, q.device)
        check(lib.b200_attn_decode_varlen(q.data_ptr(), q.stride(0), k.data_ptr(), v.data_ptr(), out.data_ptr(), out.stride(0),
                                          cu_seqlens_q.data_ptr(), cu_seqlens_k.data_ptr(), bt_ptr, bt_stride, B, Hq, Hkv, D,
                                          PAGE, float(scale), num_splits, ws.data_ptr() if ws is not None else None,
                                          ws.numel() if ws is not None else 0, plan.buf.data_ptr(), plan.buf.numel(),
                                          _stream()), "b200_attn_decode_varlen")
    return out


# ------------------------------------------------------------------------------------ (4)
def qk_norm_rope(q: torch.Tensor, k: torch.Tensor, positions: torch.Tensor, q_weight: Optional[torch.Tensor],
                 k_weight: Optional[torch.Tensor], cos_sin_cache: torch.Tensor, eps: float = 1e-6) -> Tuple[torch.Tensor, torch.Tensor]:
    """q [T,Hq,128], k [T,Hkv,128] (views of qkv) -> fresh contiguous (q', k'); qwen3.py:191-195."""
    T, Hq, D = q.shape
    Hkv = k.shape[1]
    _need_cuda_bf16(q, "q")
    _need_cuda_bf16(k, "k")
    assert q.stride(-1) == 1 and q.stride(1) == D and k.stride(-1) == 1 and k.stride(1) == D
    assert positions.dtype == torch.int64 and positions.is_cuda and positions.is_contiguous()
    assert cos_sin_cache.dtype == torch.float32 and cos_sin_cache.is_contiguous() and cos_sin_cache.shape[-1] == D
    q_out = torch.empty((T, Hq, D), dtype=q.dtype, device=q.device)
    k_out = torch.empty((T, Hkv, D), dtype=k.dtype, device=k.device)
    check(load_library().b200_qk_norm_rope(q.data_ptr(), q.stride(0), k.data_ptr(), k.stride(0),
                                           positions.data_ptr(), cos_sin_cache.data_ptr(),
                                           q_weight.data_ptr() if q_weight is not None else None,
                                           k_weight.data_ptr() if k_weight is not None else None, float(eps), q_out.data_ptr(), k_out.data_ptr(),
                                           T, Hq, Hkv, D, _stream()), "b200_qk_norm_rope")
    return q_out, k_out


def qk_norm_rope_store(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, positions: torch.Tensor,
                       q_weight: Optional[torch.Tensor], k_weight: Optional[torch.Tensor], cos_sin_cache: torch.Tensor,
                       eps: float, k_cache: torch.Tensor, v_cache: torch.Tensor, slot_mapping: torch.Tensor,
                       need_k_out: bool = True) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
    """qk_norm_rope + store_kvcache in one launch.  Returns (q', k' or None); k' (rotated)
    and v are in the caches at `slot_mapping` when this returns (stream order)."""
    T, Hq, D = q.shape
    Hkv = k.shape[1]
    for t, n in ((q, "q"), (k, "k"), (v, "v"), (k_cache, "k_cache"), (v_cache, "v_cache")):
        _need_cuda_bf16(t, n)
    _need_i32(slot_mapping, "slot_mapping")
    assert q.stride(-1) == 1 and q.stride(1) == D and k.stride(-1) == 1 and k.stride(1) == D
    assert v.stride(-1) == 1 and v.stride(1) == D and v.shape == k.shape
    assert k_cache.stride(1) == Hkv * D and v_cache.stride(1) == Hkv * D and k_cache.is_contiguous() and v_cache.is_contiguous()
    assert slot_mapping.numel() == T
    assert positions.dtype == torch.int64 and positions.is_cuda and positions.is_contiguous()
    assert cos_sin_cache.dtype == torch.float32 and cos_sin_cache.is_contiguous() and cos_sin_cache.shape[-1] == D
    q_out = torch.empty((T, Hq, D), dtype=q.dtype, device=q.device)
    k_out = torch.empty((T, Hkv, D), dtype=k.dtype, device=k.device) if need_k_out else None
    check(load_library().b200_qk_norm_rope_store(
        q.data_ptr(), q.stride(0), k.data_ptr(), k.stride(0), v.data_ptr(), v.stride(0),
        positions.data_ptr(), cos_sin_cache.data_ptr(),
        q_weight.data_ptr() if q_weight is not None else None, k_weight.data_ptr() if k_weight is not None else None,
        float(eps), q_out.data_ptr(), k_out.data_ptr() if k_out is not None else None,
        k_cache.data_ptr(), v_cache.data_ptr(), slot_mapping.data_ptr(), T, Hq, Hkv, D, _stream()), "b200_qk_norm_rope_store")
    return q_out, k_out


def rms_norm(x: torch.Tensor, weight: torch.Tensor, eps: float, residual: Optional[torch.Tensor] = None):
    """layernorm.py:143-165.  With `residual`, returns (normed, residual_out) like the reference;
    residual is updated in place."""
    _need_cuda_bf16(x, "x")
    assert x.is_contiguous()
    hidden = x.shape[-1]
    T = x.numel() // hidden
    out = torch.empty_like(x)
    if residual is not None:
        _need_cuda_bf16(residual, "residual")
        assert residual.is_contiguous() and residual.shape == x.shape
    check(load_library().b200_rmsnorm(x.data_ptr(), residual.data_ptr() if residual is not None else None,
                                      weight.data_ptr(), float(eps), out.data_ptr(), T, hidden, _stream()),
          "b200_rmsnorm")
    return out if residual is None else (out, residual)


def silu_and_mul(x: torch.Tensor) -> torch.Tensor:
    """activation.py:36-58: [.., 2I] -> [.., I]."""
    _need_cuda_bf16(x, "x")
    assert x.is_contiguous() and x.shape[-1] % 2 == 0
    inter = x.shape[-1] // 2
    T = x.numel() // (2 * inter)
    out = torch.empty(x.shape[:-1] + (inter,), dtype=x.dtype, device=x.device)
    check(load_library().b200_silu_and_mul(x.data_ptr(), out.data_ptr(), T, inter, _stream()), "b200_silu_and_mul")
    return out


# ---------------------------------------------------------------------------- (f3) TP all-reduce
def ar_choose_form(world: int, nbytes: int, max_bytes: int, two_shot_min_bytes: int, two_shot_form: int = 2) -> int:
    """Which form of csrc/allreduce.cu carries a message (include/b200attn.h: 0 one-shot/pull, 1 two-shot/pull,
    2 two-shot/push, 3 one-shot/push).  Pure function of sizes, so every rank picks the same."""
    if world > 2 and nbytes >= two_shot_min_bytes:
        return two_shot_form
    if nbytes * world <= 2 * max_bytes:                         # every rank's copy fits one slot
        return 3
    return two_shot_form if world > 2 else 0


class TpAllReduce:
    """Small-message all-reduce over NVLink peer memory (csrc/allreduce.cu) for the partial sums of
    RowParallelLinear (nanovllm/layers/linear.py:407-410).  One instance per process; the IPC handles
    travel through the reference's own torch.distributed group.  Messages larger than `max_bytes`
    (prefill-sized) stay on dist.all_reduce, exactly the reference's call."""

    def __init__(self, rank: int, world: int, max_bytes: int, gather_handles):
        import ctypes
        lib = load_library()
        self.rank, self.world, self.max_bytes = rank, world, max_bytes
        self.two_shot_form = int(__import__("os").environ.get("B200ATTN_AR_TWO_SHOT_FORM", 2))     # 1 = pull, 2 = push
        self.two_shot_min_bytes = int(__import__("os").environ.get("B200ATTN_AR_TWO_SHOT_MIN", _ar_two_shot_min_bytes(world)))
        nbytes = lib.b200_ar_buffer_bytes(max_bytes)
        ptr, handle = ctypes.c_void_p(), ctypes.create_string_buffer(64)
        check(lib.b200_ar_alloc(nbytes, ctypes.byref(ptr), handle), "b200_ar_alloc")
        handles = gather_handles(bytes(handle.raw))             # list of `world` 64-byte handles, rank order
        assert len(handles) == world and handles[rank] == bytes(handle.raw)
        peers = (ctypes.c_void_p * world)()
        for r in range(world):
            if r == rank:
                peers[r] = ptr.value
            else:
                p = ctypes.c_void_p()
                check(lib.b200_ar_open(ctypes.create_string_buffer(handles[r], 64), ctypes.byref(p)), "b200_ar_open")
                peers[r] = p.value
        comm = ctypes.c_void_p()
        check(lib.b200_ar_create(rank, world, peers, max_bytes, ctypes.byref(comm)), "b200_ar_create")
        self._comm, self._lib = comm, lib

    def choose_form(self, nbytes: int) -> int:
        return ar_choose_form(self.world, nbytes, self.max_bytes, self.two_shot_min_bytes, self.two_shot_form)

    def fits(self, t: torch.Tensor) -> bool:
        """Carrier choice depends on dtype, shape and layout only -- never on the address, which may differ
        between ranks: a rank that went to NCCL alone would leave its peers spinning."""
        return (t.is_cuda and t.dtype == torch.bfloat16 and t.is_contiguous() and t.numel() % 8 == 0
                and t.numel() * 2 <= self.max_bytes)

    def all_reduce_(self, t: torch.Tensor, two_shot: Optional[bool] = None) -> torch.Tensor:
        if not self.fits(t):
            raise RuntimeError("b200attn: tensor does not fit the peer-memory all-reduce (bf16, contiguous, "
                               f"numel % 8 == 0, <= {self.max_bytes} bytes)")
        if two_shot is None:
            two_shot = self.choose_form(t.numel() * 2)
        work = t if t.data_ptr() % 16 == 0 else t.clone()        # a misaligned view is reduced through a copy
        check(self._lib.b200_allreduce_bf16(self._comm, work.data_ptr(), work.numel(), int(two_shot), _stream()),
              "b200_allreduce_bf16")
        if work is not t:
            t.copy_(work)
        return t

    def status(self) -> int:
        """0 = healthy, 1 = a wait inside a launch timed out since creation.  Synchronises the device."""
        import ctypes
        out = ctypes.c_int32(0)
        check(self._lib.b200_ar_status(self._comm, ctypes.byref(out)), "b200_ar_status")
        return out.value

    def destroy(self) -> None:
        """Close the peers' IPC mappings and free the buffers (the caller synchronised device and group)."""
        if self._comm is not None:
            check(self._lib.b200_ar_destroy(self._comm), "b200_ar_destroy")
            self._comm = None


_TP_AR = None
# Measured on B200 NVSwitch boxes against NCCL 2.28 (tools/ar_check.py, profiles/r1h_allreduce_tp{2,4,8}.json,
# per CUDA-graph-replayed call, us, the form this module picks vs NCCL):
#   TP2  8 KB 10.0/14.4   64 KB 13.2/17.1   512 KB 13.9/18.4   2 MB 15.7/25.3   4 MB 19.0/43.7   (one-shot push)
#   TP4  8 KB 10.6/19.3   64 KB 15.2/21.2   512 KB 16.2/23.4   2 MB 23.1/30.4   4 MB 29.7/42.4   (one-shot push; 4 MB two-shot push)
#   TP8  8 KB 12.4/30.2   64 KB 19.1/32.1   512 KB 20.4/34.2   2 MB 27.3/38.1   4 MB 34.0/58.6   (one-shot push; two-shot push from 1 MB)
# Peer stores are posted, peer loads are round trips, so the push forms beat the pull forms by 2-6 us everywhere;
# the two-shot form pays a second cross-GPU barrier (~8 us) and wins once (world-1) x message no longer fits.
AR_DEFAULT_MAX_BYTES = 4 << 20


def _ar_default_max_bytes(world: int) -> int:
    return AR_DEFAULT_MAX_BYTES


def _ar_two_shot_min_bytes(world: int) -> int:
    return (4 << 20) if world <= 4 else (1 << 20)


def tp_all_reduce_(t: torch.Tensor) -> torch.Tensor:
    """In-place sum over the tensor-parallel group: peer-memory kernel when the message fits, the
    reference's dist.all_reduce otherwise.  The communicator is built on first use, which every rank
    reaches at the same point (the first forward of the warm-up, model_runner.py:311-316)."""
    import os
    import torch.distributed as dist
    global _TP_AR
    if not t.is_cuda:
        raise RuntimeError("b200attn: tp_all_reduce_ needs a CUDA tensor (no CPU fallback on this path)")
    if _TP_AR is None:
        world = dist.get_world_size()
        if world > 8:
            _TP_AR = False
        else:
            def gather(mine: bytes):
                out = [None] * world
                dist.all_gather_object(out, mine)
                return out
            _TP_AR = TpAllReduce(dist.get_rank(), world, int(os.environ.get("B200ATTN_AR_MAX_BYTES", _ar_default_max_bytes(world))), gather)
    if _TP_AR is not False and _TP_AR.fits(t):
        return _TP_AR.all_reduce_(t)
    dist.all_reduce(t)
    return t


# ------------------------------------------------- (f3) all-reduce fused with add + RMSNorm on NVLS multicast
class NvlsAllReduce:
    """csrc/allreduce_nvls.cu: ONE launch per rank = all-reduce of the RowParallelLinear partial sums (in-switch
    multimem.ld_reduce) + residual add + RMSNorm (linear.py:407-410 followed by layernorm.py:95-141).  The
    symmetric buffer and its multicast mapping come from torch.distributed._symmetric_memory on the reference's
    own process group (plumbing only: cuMemCreate / cuMulticast* and the handle exchange)."""

    def __init__(self, group, max_bytes: int, device=None, group_barrier: bool = True):
        """group_barrier=False: the caller runs the group-wide synchronisation that must follow b200_nvls_create on
        every rank itself (tp_nvls folds it into its agreement all-reduce, so a rank whose set-up threw is not left
        facing its peers' barrier with a different collective)."""
        import ctypes
        import torch.distributed as dist
        lib = load_library()
        self._lib = lib
        self.world = dist.get_world_size(group) if group is not None else 1
        self.rank = dist.get_rank(group) if group is not None else 0
        self.max_bytes = int(max_bytes)
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        nbytes = lib.b200_nvls_buffer_bytes(self.max_bytes)
        mc_ptr, peers = None, None
        if self.world > 1:
            import torch.distributed._symmetric_memory as symm_mem
            self.buf = symm_mem.empty(nbytes, dtype=torch.uint8, device=dev)
            self.hdl = symm_mem.rendezvous(self.buf, group)
            mc_ptr = int(self.hdl.multicast_ptr)
            if not mc_ptr:
                raise RuntimeError("b200attn: this group has no NVLS multicast mapping (multicast_ptr == 0)")
            peers = (ctypes.c_void_p * self.world)(*[int(p) for p in self.hdl.buffer_ptrs])
            assert int(self.hdl.buffer_ptrs[self.rank]) == self.buf.data_ptr()
        else:
            self.buf = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        comm = ctypes.c_void_p()
        check(lib.b200_nvls_create(self.rank, self.world, self.buf.data_ptr(), mc_ptr, peers, self.max_bytes,
                                   ctypes.byref(comm)), "b200_nvls_create")
        self._comm = comm
        self.ll_max_bytes = min(self.max_bytes, 4 << 20)
        env = __import__("os").environ
        # thresholds measured with tools/nvls_check.py (profiles/r2_nvls_tp*.json): bytes a rank RECEIVES in the LL one-shot
        # form; largest message of the LL two-shot form; extra bytes the one-shot barrier form may move to save a barrier
        self.ll1_budget = int(env.get("B200ATTN_NVLS_LL1_BUDGET", 1 << 20))
        self.ll2_max = min(int(env.get("B200ATTN_NVLS_LL2_MAX", 1 << 20)), self.ll_max_bytes)
        self.one_shot_budget = int(env.get("B200ATTN_NVLS_ONE_SHOT_BUDGET", 3 << 19))
        self.fused_max_pair = int(env.get("B200ATTN_NVLS_FUSED_MAX_TP2", 1 << 20))
        self.forced_form = int(env["B200ATTN_NVLS_FORM"]) if "B200ATTN_NVLS_FORM" in env else None
        if self.world > 1 and group_barrier:
            dist.barrier(group)            # every rank zeroed its buffer before anyone can write into it (b200attn.h)

    def choose_form(self, nbytes: int) -> int:
        """Form policy (include/b200attn.h), a pure function of (group size, message bytes) so every rank picks the
        same: the flag-in-data forms save every barrier (~6 us each) and pay with twice the bytes on the wire --
        one-shot (2) while the 2 x world x message bytes a rank receives stay small, two-shot (3) up to ~1 MB;
        beyond that the in-switch reduction, every rank for itself (1) or reduce-scatter + all-gather (0)."""
        if self.forced_form is not None:
            return self.forced_form
        if 2 * self.world * nbytes <= min(self.ll1_budget, 4 * self.ll_max_bytes):
            return 2
        if nbytes <= self.ll2_max:
            return 3
        return 1 if nbytes * max(self.world - 2, 0) <= self.one_shot_budget else 0

    def form_fits(self, form: int, nbytes: int) -> bool:
        """Whether `form` can carry a message of nbytes (the range checks of b200_nvls_allreduce_add_rmsnorm)."""
        if form == 2:
            return 2 * self.world * nbytes <= 4 * self.ll_max_bytes + 1024
        if form == 3:
            return nbytes <= self.ll_max_bytes
        return form in (0, 1) and nbytes <= self.max_bytes

    def pays(self, x: torch.Tensor) -> bool:
        """Whether deferring RowParallelLinear's all-reduce of `x` into the RMSNorm launch beats the peer-memory
        all-reduce followed by the RMSNorm kernel.  Measured (profiles/r2_nvls_tp2.json, r3_nvls_tp4.json, r3_nvls_tp8.json; us per
        graph-replayed call, chosen fused form vs peer-store all-reduce + norm kernel): TP8 11.3/17.0 at 8 KB, 16.9/26.6 at
        512 KB, 28.0/35.8 at 2 MB, 36.4/41.9 at 4 MB (1.15-1.9x); TP4 10.8/16.1 at 8 KB, 13.8/21.1 at 512 KB, a tie at 2 MB
        (28.1/27.9); with two ranks only up to ~1 MB -- beyond, one posted peer store of the whole message beats the
        round trips of multimem.ld_reduce."""
        return self.fits(x) and (self.world > 2 or x.numel() * 2 <= self.fused_max_pair)

    def fits(self, x: torch.Tensor) -> bool:
        return (x.is_cuda and x.dtype == torch.bfloat16 and x.dim() == 2 and x.is_contiguous()
                and x.shape[1] % 8 == 0 and x.shape[1] <= 16384 and 0 < x.numel() * 2 <= self.max_bytes)

    def all_reduce_add_rms_norm(self, x: torch.Tensor, residual: Optional[torch.Tensor], weight: Optional[torch.Tensor],
                                eps: float, form: Optional[int] = None, out: Optional[torch.Tensor] = None):
        """(normed, residual) like RMSNorm.forward(x_allreduced, residual); weight None => plain all-reduce -> out."""
        if not self.fits(x):
            raise RuntimeError("b200attn: tensor does not fit the NVLS fused all-reduce")
        T, hidden = x.shape
        if residual is not None:
            _need_cuda_bf16(residual, "residual")
            assert residual.is_contiguous() and residual.shape == x.shape
        if weight is not None:
            _need_cuda_bf16(weight, "weight")
            assert weight.is_contiguous() and weight.numel() == hidden
        if form is None:
            form = self.choose_form(x.numel() * 2)
        if out is None:
            out = torch.empty_like(x)
        xin = x if x.data_ptr() % 16 == 0 else x.clone()
        check(self._lib.b200_nvls_allreduce_add_rmsnorm(
            self._comm, xin.data_ptr(), residual.data_ptr() if residual is not None else None,
            weight.data_ptr() if weight is not None else None, float(eps), out.data_ptr(), T, hidden,
            int(form), _stream()), "b200_nvls_allreduce_add_rmsnorm")
        return out if residual is None else (out, residual)

    def status(self) -> int:
        import ctypes
        out = ctypes.c_int32(0)
        check(self._lib.b200_nvls_status(self._comm, ctypes.byref(out)), "b200_nvls_status")
        return out.value

    def destroy(self) -> None:
        if self._comm is not None:
            check(self._lib.b200_nvls_destroy(self._comm), "b200_nvls_destroy")
            self._comm = None


_TP_NVLS = None          # None = not tried yet, False = unavailable (multicast setup failed; the peer-store forms carry on)
_DEFER_DEPTH = 0         # > 0 while a patched model forward is running: RowParallelLinear may hand over partial sums
_PENDING = None          # partial sums of the last RowParallelLinear whose all-reduce has been deferred to its RMSNorm
NVLS_DEFAULT_MAX_BYTES = 8 << 20


def tp_nvls():
    """The process-wide NVLS communicator of the tensor-parallel group, built on first use (every rank reaches the
    first RowParallelLinear of the warm-up forward together, model_runner.py:311-316).  B200ATTN_NVLS=0, a group
    without multicast support or a failed rendezvous leave the (already measured) peer-store all-reduce + separate
    RMSNorm kernel in charge -- never a silent NCCL-only path for messages the peer-memory kernels can carry."""
    import os
    import torch.distributed as dist
    global _TP_NVLS
    if _TP_NVLS is None:
        _TP_NVLS = False
        if os.environ.get("B200ATTN_NVLS", "1") != "0" and dist.is_initialized() and 1 < dist.get_world_size() <= 8:
            ok = 1
            try:
                comm = NvlsAllReduce(dist.group.WORLD, int(os.environ.get("B200ATTN_NVLS_MAX_BYTES", NVLS_DEFAULT_MAX_BYTES)),
                                     group_barrier=False)
            except Exception as exc:                            # noqa: BLE001 -- any set-up failure means "unavailable"
                comm, ok = None, 0
                print(f"b200attn[rank {dist.get_rank()}]: NVLS fused all-reduce unavailable ({type(exc).__name__}: {exc}); "
                      "using the peer-store all-reduce + RMSNorm kernels", flush=True)
            # The FIRST collective after the try block, on success and on failure alike (a rank that threw
            # must not meet its peers' barrier with an all-reduce).  It does two jobs: all ranks must agree, or some
            # would defer their all-reduce and others not; and it is the group-wide synchronisation after
            # b200_nvls_create -- the buffer was zeroed on the legacy default stream this all-reduce is ordered behind,
            # and .item() returns only once every rank has contributed.
            flag = torch.tensor([ok], device="cuda", dtype=torch.int32)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
            if int(flag.item()) == 1:
                _TP_NVLS = comm
            elif comm is not None:
                comm.destroy()
    return _TP_NVLS


class deferred_all_reduce:
    """Context manager around a model forward: inside it `tp_row_parallel_output` may return UN-reduced partial
    sums and remember them; the RMSNorm that consumes them (`tp_norm_input`) runs the fused kernel.  Anything
    still pending at exit is reduced in place, so no caller ever sees partial sums outside."""

    def __enter__(self):
        global _DEFER_DEPTH
        _DEFER_DEPTH += 1
        return self

    def __exit__(self, *exc):
        global _DEFER_DEPTH
        _DEFER_DEPTH -= 1
        if _DEFER_DEPTH == 0:
            flush_pending_all_reduce()
        return False


def flush_pending_all_reduce() -> None:
    global _PENDING
    if _PENDING is not None:
        t, _PENDING = _PENDING, None
        tp_all_reduce_(t)


def tp_row_parallel_output(y: torch.Tensor) -> torch.Tensor:
    """RowParallelLinear's `dist.all_reduce(y)` (linear.py:410): immediately, or -- inside `deferred_all_reduce`,
    for messages the NVLS kernel carries -- left to the RMSNorm that follows (qwen3.py:370,374,440)."""
    global _PENDING
    flush_pending_all_reduce()                                  # an earlier deferred sum nobody normalised
    if _DEFER_DEPTH > 0:
        nv = tp_nvls()
        if nv and nv.pays(y):
            _PENDING = y
            return y
    return tp_all_reduce_(y)


def tp_add_rms_norm(x: torch.Tensor, residual: Optional[torch.Tensor], weight: torch.Tensor, eps: float):
    """RMSNorm.forward(x, residual) (layernorm.py:143-165); when `x` is the pending partial-sum tensor the
    all-reduce happens inside the same launch."""
    global _PENDING
    if _PENDING is not None:
        if _PENDING is x and residual is not None and x.is_contiguous() and residual.is_contiguous():
            _PENDING = None
            return _TP_NVLS.all_reduce_add_rms_norm(x, residual, weight, eps)
        # x is something else -- possibly a VIEW of the pending tensor: reduce the original in place first, so
        # that whatever copy is made below reads reduced sums
        flush_pending_all_reduce()
    if not x.is_contiguous():
        x = x.contiguous()
    if residual is not None and not residual.is_contiguous():
        residual = residual.contiguous()
    return rms_norm(x, weight, eps, residual)


def collectives_status() -> int:
    """Worst status word of the live peer-memory communicators: 0 = healthy, 1 = some wait inside a launch gave up
    after B200ATTN_AR_TIMEOUT_S (b200_ar_status / b200_nvls_status).  Costs one 4-byte device->host copy per
    communicator, so call it where the host has just synchronised anyway."""
    worst = 0
    for c in (_TP_AR, _TP_NVLS):
        if c not in (None, False):
            worst = max(worst, int(c.status()))
    return worst


def raise_if_collectives_failed() -> None:
    """After a timeout the barrier and packet epochs of the ranks no longer agree, and every later all-reduce or
    fused norm returns garbage: turn that into an exception instead of wrong tokens.  The reference's
    NCCL path would abort the process in the same situation (watchdog, linear.py:407-410's dist.all_reduce)."""
    if collectives_status() != 0:
        raise RuntimeError("b200attn: a tensor-parallel all-reduce kernel timed out waiting for a peer rank "
                           "(B200ATTN_AR_TIMEOUT_S); every result since then is invalid -- restart the engine "
                           "(B200ATTN_ALLREDUCE=0 B200ATTN_NVLS=0 keeps the collectives on NCCL)")


def shutdown_collectives() -> None:
    """Tear the peer-memory communicators down (b200_ar_destroy / b200_nvls_destroy) before the reference
    destroys its process group (ModelRunner.exit, model_runner.py:187-193).  Collective: every rank calls it."""
    import torch.distributed as dist
    global _TP_AR, _TP_NVLS, _PENDING
    _PENDING = None
    live = [c for c in (_TP_AR, _TP_NVLS) if c not in (None, False)]
    if not live:
        return
    torch.cuda.synchronize()
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.barrier()                              # no rank is still inside a call that reads this rank's buffer
    failed = any(int(c.status()) != 0 for c in live)
    for c in live:
        c.destroy()
    _TP_AR = _TP_NVLS = None
    if failed:                                      # workers have no per-step check: this is where they report
        raise RuntimeError("b200attn: a tensor-parallel all-reduce kernel had timed out before shutdown "
                           "(B200ATTN_AR_TIMEOUT_S); results of this process after that point were invalid")
