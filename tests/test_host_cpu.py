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
