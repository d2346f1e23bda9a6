"""b200attn -- host-side mirror of nano-vllm-v1's attention hot path on B200 (sm_100a).

Thin Python over the C ABI in include/b200attn.h (libb200attn.so, hand-written CUDA).
Module layout mirrors the reference so parity tests read like the reference's call sites:

    b200attn.layers.attention          <- nanovllm/layers/attention.py   (Attention, store_kvcache)
    b200attn.layers.linear             <- nanovllm/layers/linear.py      (RowParallelLinear.forward, f3)
    b200attn.utils.context             <- nanovllm/utils/context.py      (Context, get/set/reset_context)
    b200attn.engine.model_runner       <- nanovllm/engine/model_runner.py (the in-scope ModelRunner methods)
    b200attn.install                   patches an importable `nanovllm` in place (drop-in): RMSNorm,
                                       RotaryEmbedding and SiluAndMul forward are bound onto the
                                       reference's own classes there, straight to b200attn.ops

There is NO CPU fallback, no Triton and no flash-attn on this path: if the CUDA library is
missing or a tensor is not on a CUDA device the ops raise.
"""
from . import ops  # noqa: F401
from ._lib import lib_path, load_library  # noqa: F401

__version__ = "0.1.0"
