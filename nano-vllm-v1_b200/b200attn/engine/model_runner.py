"""The in-scope methods of nanovllm/engine/model_runner.py, re-done for the B200 path
(SURVEY.md rows a11-a13).  They are bound onto the reference's own `ModelRunner` class by
`b200attn.install.install()`; everything else of the engine stays the reference's code.

  allocate_kv_cache   model_runner.py:319-375  same tensor, same num_blocks arithmetic (the
                      reference method itself runs), plus the split-K workspace of the decode
                      kernel sized once for the largest CUDA-graph bucket so that capture and
                      replay never allocate.
  prepare_model_input model_runner.py:400-495  same eight Context fields with the same values
                      and dtypes; built with numpy instead of per-token Python list appends,
                      staged in ONE pinned buffer and moved with ONE H2D copy (the reference
                      makes 7-8 pinned allocations + copies per step).
  run                 model_runner.py:602-638  same pipeline and return shape; the indices of the
                      sequences that produced a token are returned as a host list (they were built on
                      the host in prepare_model_input) instead of the CUDA int32 tensor, so that the
                      untouched `Scheduler.postprocess` (scheduler.py:248) no longer pays one
                      device->host sync per sequence when it iterates them (3.6 ms of an
                      11.7 ms Qwen3-8B decode step at 256 sequences, profiles/r1f_step_phases.json).
  capture_cudagraph   model_runner.py:640-715  unchanged logic (the reference method runs);
                      the decode kernel it captures reads every length from device memory and
                      its launch geometry depends on host integers only.
"""
import itertools

import numpy as np
import torch

from .. import ops

_ALIGN = 16


def allocate_kv_cache(self, _reference_impl):
    _reference_impl(self)
    hf = self.config.hf_config
    tp = self.world_size
    num_heads = hf.num_attention_heads // tp
    num_kv_heads = hf.num_key_value_heads // tp
    max_bs = max(min(self.config.max_num_seqs, 512), 1)
    max_blocks = (self.config.max_model_len + self.block_size - 1) // self.block_size
    ops.reserve_decode_workspace(max_bs, num_heads, num_kv_heads, max_blocks * self.block_size,
                                 torch.device("cuda", torch.cuda.current_device()))


class _Staging:
    """Two pinned arenas used alternately; an event per arena guards reuse."""

    def __init__(self):
        self.bufs = [None, None]
        self.events = [None, None]
        self.turn = 0

    def get(self, nbytes: int) -> torch.Tensor:
        self.turn ^= 1
        i = self.turn
        if self.events[i] is not None:
            self.events[i].synchronize()
        if self.bufs[i] is None or self.bufs[i].numel() < nbytes:
            self.bufs[i] = torch.empty(max(nbytes * 2, 1 << 16), dtype=torch.uint8, device="cpu",
                                       pin_memory=torch.cuda.is_available())
        return self.bufs[i]

    def sent(self):
        if not torch.cuda.is_available():
            return
        ev = torch.cuda.Event()
        ev.record()
        self.events[self.turn] = ev


def build_model_input(seqs, block_size: int):
    """Pure host part of prepare_model_input: numpy arrays with the reference's values.

    Returns a dict of arrays (block_tables None when sum(k) == sum(q), model_runner.py:483) and
    the two host ints.  Kept separate from the H2D staging so it is testable on CPU against
    fixtures generated from the reference's own method (tests/golden/prepare_*.pt)."""
    n = len(seqs)
    cached = np.fromiter((s.num_cached_tokens for s in seqs), dtype=np.int64, count=n)
    new = np.fromiter((s.num_new_tokens for s in seqs), dtype=np.int64, count=n)
    ctx = cached + new
    cu_q = np.zeros(n + 1, dtype=np.int64)
    cu_k = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(new, out=cu_q[1:])
    np.cumsum(ctx, out=cu_k[1:])
    total = int(cu_q[-1])
    tok_seq = np.repeat(np.arange(n, dtype=np.int64), new)               # owning sequence of every new token
    positions = np.arange(total, dtype=np.int64) - (cu_q[:-1] - cached)[tok_seq]
    input_ids = np.fromiter(itertools.chain.from_iterable(s[int(c):int(x)] for s, c, x in zip(seqs, cached, ctx)),
                            dtype=np.int64, count=total)
    has_table = np.fromiter((1 if s.block_table else 0 for s in seqs), dtype=np.int64, count=n)
    lens = np.fromiter((len(s) for s in seqs), dtype=np.int64, count=n)
    need_logits = np.nonzero((lens == ctx) & (has_table == 1))[0].astype(np.int32)
    width = max((len(s.block_table) for s in seqs), default=0)
    table = None
    if width:
        table = np.full((n, width), -1, dtype=np.int32)
        for i, s in enumerate(seqs):
            bt = s.block_table
            if bt:
                table[i, :len(bt)] = bt
    if has_table.all() and total:
        slot = table[tok_seq, positions // block_size].astype(np.int64) * block_size + positions % block_size
        slot_mapping = slot.astype(np.int32)
    else:                                                                 # warmup: sequences without a block table
        keep = has_table[tok_seq] == 1
        if keep.any():
            p, sq = positions[keep], tok_seq[keep]
            slot_mapping = (table[sq, p // block_size].astype(np.int64) * block_size + p % block_size).astype(np.int32)
        else:
            slot_mapping = np.zeros(0, dtype=np.int32)
    arrays = dict(input_ids=input_ids, positions=positions, cu_seqlens_q=cu_q.astype(np.int32),
                  cu_seqlens_k=cu_k.astype(np.int32), slot_mapping=slot_mapping, context_lens=ctx.astype(np.int32),
                  seq_need_compute_logits=need_logits,
                  block_tables=table if (int(cu_k[-1]) > total and table is not None) else None)
    return arrays, int(new.max()) if n else 0, int(ctx.max()) if n else 0


_ORDER = ("input_ids", "positions", "cu_seqlens_q", "cu_seqlens_k", "slot_mapping", "context_lens",
          "seq_need_compute_logits", "block_tables")


def _layout_of(arrays):
    """[(name, offset, nbytes, dtype_char, shape)] for the non-None arrays, 16-byte aligned offsets."""
    layout, nbytes = [], 0
    for name in _ORDER:
        a = arrays[name]
        if a is None:
            continue
        layout.append((name, nbytes, a.nbytes, "q" if a.dtype == np.int64 else "i", tuple(a.shape)))
        nbytes += (a.nbytes + _ALIGN - 1) // _ALIGN * _ALIGN
    return layout, max(nbytes, _ALIGN)


def _staging_of(self):
    staging = getattr(self, "_b200_staging", None)
    if staging is None:
        staging = self._b200_staging = _Staging()
    return staging


def _pack(self, arrays):
    """Arrays -> one pinned host arena (returns its layout and the pinned uint8 tensor)."""
    layout, nbytes = _layout_of(arrays)
    host = _staging_of(self).get(nbytes)
    host_np = host.numpy()
    for name, off, nb, _, _ in layout:
        host_np[off:off + nb] = arrays[name].reshape(-1).view(np.uint8)
    return layout, nbytes, host


def _upload_and_set_context(self, layout, nbytes, host, max_q, max_k, _set_context):
    """ONE H2D copy of the packed arena, then the eight Context fields as views of it."""
    dev = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    dev.copy_(host[:nbytes], non_blocking=True)
    _staging_of(self).sent()
    out = dict.fromkeys(_ORDER)
    for name, off, nb, ch, shape in layout:
        out[name] = dev[off:off + nb].view(torch.int64 if ch == "q" else torch.int32).view(shape)
    _set_context(out["cu_seqlens_q"], out["cu_seqlens_k"], max_q, max_k, out["slot_mapping"], out["context_lens"],
                 out["block_tables"], out["seq_need_compute_logits"])
    return out["input_ids"], out["positions"]


def prepare_model_input(self, seqs, _set_context):
    pre = getattr(self, "_b200_prebuilt", None)
    if pre is not None and pre[0] is seqs:                 # rank 0 under TP: `call` already built and packed this step
        _, layout, nbytes, host, max_q, max_k, need = pre
        self._b200_prebuilt = None
    else:
        arrays, max_q, max_k = build_model_input(seqs, self.block_size)
        layout, nbytes, host = _pack(self, arrays)
        need = arrays["seq_need_compute_logits"].tolist()
    self._b200_need_logits_host = need
    return _upload_and_set_context(self, layout, nbytes, host, max_q, max_k, _set_context)

def run(self, seqs, _get_context, _reset_context):
    """ModelRunner.run (model_runner.py:602-638): prepare -> forward -> sample; see module docstring."""
    self._b200_need_logits_host = None
    input_ids, positions = self.prepare_model_input(seqs)
    temperatures = self.prepare_sample(seqs) if self.rank == 0 else None
    logits = self.run_model(input_ids, positions)
    token_ids = self.sampler(logits, temperatures).tolist() if self.rank == 0 else None
    if self.rank == 0 and self.world_size > 1:
        # .tolist() has just synchronised the device: the cheapest place to learn that an all-reduce kernel of this
        # step gave up on a stalled peer, before its garbage becomes tokens (workers report at shutdown)
        ops.raise_if_collectives_failed()
    need = self._b200_need_logits_host
    if need is None:                                       # prepare_model_input was not ours
        need = _get_context().seq_need_compute_logits
    _reset_context()
    return token_ids, need


# ---- tensor-parallel control plane (model_runner.py:241-288) ---------------------------
# The reference pickles the scheduled `Sequence` objects -- every token id of every sequence -- into the
# shared-memory mailbox on every step, and each worker unpickles them and re-runs prepare_model_input.
# All a worker needs is the packed metadata arena rank 0 builds anyway, so rank 0 sends THAT (one bytes
# object + a tiny layout) as a "run_packed" call; workers upload it and run the forward.  Message
# format, mailbox, events and the `loop`/`read_shm`/`write_shm` methods are the reference's, untouched.
# An arena larger than the reference's 1 MiB mailbox travels through a side shared-memory segment
# (the reference's own message would overflow there too): the cap is lifted, not inherited.
_MAILBOX_HEADROOM = 4096
_ARENA_NAME = "nanovllm_b200_arena"


def _shared_memory(name, create=False, size=0):
    from multiprocessing.shared_memory import SharedMemory
    if create:
        try:                                              # a crashed run may have left the segment behind (the
            stale = SharedMemory(name=name)               # reference cleans up its own the same way, model_runner.py:148-157)
            stale.close()
            stale.unlink()
        except FileNotFoundError:
            pass
    return SharedMemory(name=name, create=create, size=size)


def _side_arena(self, nbytes):
    """Rank 0: a shared-memory segment of at least `nbytes` for arenas the reference's 1 MiB mailbox
    (model_runner.py:160) cannot carry (the reference itself overflows there).  Grown by
    doubling under a new name; workers attach by the name that travels in the message."""
    cur = getattr(self, "_b200_arena", None)
    if cur is None or cur.size < nbytes:
        gen = getattr(self, "_b200_arena_gen", 0) + 1
        self._b200_arena_gen = gen
        size = max(1 << 22, 1 << (int(nbytes) - 1).bit_length() + 1)
        new = _shared_memory(f"{_ARENA_NAME}_{gen}", create=True, size=size)
        if cur is not None:
            cur.close()
            cur.unlink()
        cur = self._b200_arena = new
    return cur


def release_side_arena(self):
    """Close (and, on rank 0, unlink) the side arena; called from the patched ModelRunner.exit."""
    cur = getattr(self, "_b200_arena", None)
    if cur is not None:
        self._b200_arena = None
        cur.close()
        if getattr(self, "rank", 0) == 0:
            try:
                cur.unlink()
            except FileNotFoundError:
                pass


def call(self, method_name, *args, _reference_call=None):
    if method_name == "run" and self.world_size > 1 and self.rank == 0 and len(args) == 1:
        seqs = args[0]
        arrays, max_q, max_k = build_model_input(seqs, self.block_size)
        layout, nbytes, host = _pack(self, arrays)
        if nbytes + _MAILBOX_HEADROOM <= self.shm.size:
            self.write_shm("run_packed", layout, nbytes, max_q, max_k, host.numpy()[:nbytes].tobytes())
        else:
            # larger than the reference's mailbox: the arena travels through a side segment, only its name
            # and layout through the mailbox.  Lock-step is the reference's own: rank 0 cannot finish this
            # step (its forward holds collectives with every worker) before each worker copied the arena out.
            arena = _side_arena(self, nbytes)
            arena.buf[:nbytes] = host.numpy()[:nbytes].data
            self.write_shm("run_packed", layout, nbytes, max_q, max_k, (arena.name, arena.size))
        self._b200_prebuilt = (seqs, layout, nbytes, host, max_q, max_k, arrays["seq_need_compute_logits"].tolist())
        return self.run(seqs)
    if method_name == "exit":
        if self.rank == 0 and self.world_size > 1:
            self.write_shm("exit")                         # workers first: the tear-down below is collective
        try:
            late_error = None
            try:
                ops.shutdown_collectives()                 # b200_ar_destroy / b200_nvls_destroy, before the group goes
            except RuntimeError as exc:                    # a collective had timed out: finish the tear-down, then say so
                late_error = exc
            result = self.exit()
            if late_error is not None:
                raise late_error
            return result
        finally:
            release_side_arena(self)
    return _reference_call(self, method_name, *args)


def run_packed(self, layout, nbytes, max_q, max_k, payload, _set_context, _reset_context):
    """Worker side of `call("run", seqs)`: same forward as ModelRunner.run on rank > 0 (no sampling there,
    model_runner.py:631-633), fed from rank 0's packed arena instead of re-deriving it from Sequences.
    `payload` is the arena itself (bytes) or (name, size) of the side segment that holds it."""
    host = _staging_of(self).get(nbytes)
    if isinstance(payload, (bytes, bytearray, memoryview)):
        host.numpy()[:nbytes] = np.frombuffer(payload, dtype=np.uint8, count=nbytes)
    else:
        name, _size = payload
        cur = getattr(self, "_b200_arena", None)
        if cur is None or cur.name != name:
            if cur is not None:
                cur.close()
            cur = self._b200_arena = _shared_memory(name)
        host.numpy()[:nbytes] = np.frombuffer(cur.buf, dtype=np.uint8, count=nbytes)
    input_ids, positions = _upload_and_set_context(self, layout, nbytes, host, max_q, max_k, _set_context)
    self.run_model(input_ids, positions)
    _reset_context()
    return None
