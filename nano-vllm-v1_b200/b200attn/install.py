"""Apply the drop-in to an importable `nanovllm` (the reference, unmodified on disk).

Methods are replaced ON the reference's own classes, so module objects, parameters, attribute
names (`k_cache`, `v_cache`, `weight`, `cos_sin_cache`, ...) and every `from x import y` binding
stay valid whatever was imported first:

    nanovllm.layers.attention.Attention.forward        -> scatter + decode / prefill kernels
    nanovllm.layers.attention.store_kvcache            -> b200_kv_scatter
    nanovllm.models.qwen3.Qwen3Attention.forward       -> qkv_proj, ONE fused q/k-norm+RoPE+KV-scatter launch,
                                                          attention, o_proj (qwen3.py:183-199 order kept;
                                                          B200ATTN_FUSE_STORE=0 keeps the scatter separate)
    nanovllm.layers.layernorm.RMSNorm.forward          -> b200_rmsnorm (fused residual add), or -- when its input is the
                                                          pending output of a RowParallelLinear -- the NVLS launch that
                                                          all-reduces, adds the residual and normalises in one go
    nanovllm.layers.linear.RowParallelLinear.forward   -> peer-memory all-reduce / deferred to that RMSNorm launch
    nanovllm.models.qwen3.Qwen3Model.forward           -> same code, wrapped in ops.deferred_all_reduce()
    nanovllm.layers.rotary_embedding.RotaryEmbedding.forward -> rotation-only kernel
    nanovllm.layers.activation.SiluAndMul.forward      -> b200_silu_and_mul
    nanovllm.engine.model_runner.ModelRunner.{allocate_kv_cache, prepare_model_input, run, call, run_packed}

`LLM.generate`, `Scheduler`, `BlockManager`, `Sequence`, the linears, the sampler and NCCL are untouched.
`install(pointwise=False)` limits the patch to attention + scatter; `install(runner=False)`
leaves ModelRunner alone.
"""
import importlib
import os

import torch

from . import ops
from .layers import attention as _attention

_INSTALLED = False
_FUSE_STORE = os.environ.get("B200ATTN_FUSE_STORE", "1") != "0"


def _attention_forward(self, q, k, v):
    return _attention.Attention.forward(self, q, k, v)


def _qwen3_attention_forward(self, positions, hidden_states):
    qkv = self.qkv_proj(hidden_states)
    q, k, v = qkv.split([self.q_size, self.kv_size, self.kv_size], dim=-1)
    q = q.view(-1, self.num_heads, self.head_dim)
    k = k.view(-1, self.num_kv_heads, self.head_dim)
    v = v.view(-1, self.num_kv_heads, self.head_dim)
    rope = self.rotary_emb
    wq, wk, eps = (self.q_norm.weight, self.k_norm.weight, self.q_norm.eps) if not self.qkv_bias else (None, None, 1e-6)
    attn = self.attn
    if _FUSE_STORE and attn.k_cache.numel() and attn.v_cache.numel():
        # one launch: q/k norm + RoPE + KV-cache scatter; fresh k is only read by unpaged prefill
        ctx = _attention._context()
        need_k = ctx.cu_seqlens_q is not None and ctx.block_tables is None
        q, k = ops.qk_norm_rope_store(q, k, v, positions, wq, wk, rope.cos_sin_cache, eps,
                                      attn.k_cache, attn.v_cache, ctx.slot_mapping, need_k_out=need_k)
        o = _attention.Attention.forward(attn, q, k, v, kv_stored=True)
    else:
        q, k = ops.qk_norm_rope(q, k, positions, wq, wk, rope.cos_sin_cache, eps)
        o = attn(q, k, v)
    return self.o_proj(o.flatten(1, -1))


def _rmsnorm_forward(self, x, residual=None):
    if residual is None:
        ops.flush_pending_all_reduce()
        return ops.rms_norm(x if x.is_contiguous() else x.contiguous(), self.weight, self.eps)
    # x may be the un-reduced output of the RowParallelLinear in front (ops.tp_row_parallel_output): the
    # all-reduce then runs inside this launch
    # No .contiguous() here: a copy taken before a pending all-reduce has run would hold un-reduced partial sums
    # (ADVICE r2).  tp_add_rms_norm settles the pending sum first and only then makes its inputs contiguous.
    return ops.tp_add_rms_norm(x, residual, self.weight, self.eps)


def _qwen3_model_forward(self, input_ids, positions, _reference_forward=None):
    with ops.deferred_all_reduce():
        return _reference_forward(self, input_ids, positions)


def _rotary_forward(self, positions, query, key):
    return ops.qk_norm_rope(query, key, positions, None, None, self.cos_sin_cache)


def _silu_forward(self, x):
    return ops.silu_and_mul(x if x.is_contiguous() else x.contiguous())


def install(pointwise: bool = True, runner: bool = True) -> None:
    global _INSTALLED
    if _INSTALLED:
        return
    ops.load_library()                                     # fail loudly before touching anything
    att = importlib.import_module("nanovllm.layers.attention")
    ctx = importlib.import_module("nanovllm.utils.context")
    _attention.use_context_getter(ctx.get_context)
    att.Attention.forward = _attention_forward
    att.store_kvcache = ops.store_kvcache
    if pointwise:
        qwen3 = importlib.import_module("nanovllm.models.qwen3")
        qwen3.Qwen3Attention.forward = _qwen3_attention_forward
        importlib.import_module("nanovllm.layers.layernorm").RMSNorm.forward = _rmsnorm_forward
        importlib.import_module("nanovllm.layers.rotary_embedding").RotaryEmbedding.forward = _rotary_forward
        importlib.import_module("nanovllm.layers.activation").SiluAndMul.forward = _silu_forward
    if os.environ.get("B200ATTN_ALLREDUCE", "1") != "0":      # B200ATTN_ALLREDUCE=0 keeps every all-reduce on NCCL
        from .layers.linear import row_parallel_forward
        importlib.import_module("nanovllm.layers.linear").RowParallelLinear.forward = row_parallel_forward
        if pointwise:                                          # the fused form needs our RMSNorm.forward on the other end
            qwen3 = importlib.import_module("nanovllm.models.qwen3")
            ref_model_forward = qwen3.Qwen3Model.forward
            qwen3.Qwen3Model.forward = lambda self, input_ids, positions: _qwen3_model_forward(
                self, input_ids, positions, _reference_forward=ref_model_forward)
    if runner:
        from .engine import model_runner as ours
        mr = importlib.import_module("nanovllm.engine.model_runner")
        ref_alloc = mr.ModelRunner.allocate_kv_cache
        mr.ModelRunner.allocate_kv_cache = lambda self: ours.allocate_kv_cache(self, ref_alloc)
        mr.ModelRunner.prepare_model_input = lambda self, seqs: ours.prepare_model_input(self, seqs, ctx.set_context)
        mr.ModelRunner.run = lambda self, seqs: ours.run(self, seqs, ctx.get_context, ctx.reset_context)
        ref_call = mr.ModelRunner.call
        mr.ModelRunner.call = lambda self, method_name, *args: ours.call(self, method_name, *args, _reference_call=ref_call)
        mr.ModelRunner.run_packed = lambda self, layout, nbytes, max_q, max_k, payload: ours.run_packed(
            self, layout, nbytes, max_q, max_k, payload, ctx.set_context, ctx.reset_context)
    _INSTALLED = True


def installed() -> bool:
    return _INSTALLED
