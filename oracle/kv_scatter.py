"""Oracle: KV-cache scatter driven by slot_mapping (test infrastructure only).

Restates /root/reference/nanovllm/layers/attention.py:34-90 (`store_kvcache_kernel`)
and its wrapper :93-127.  Pure 16-bit copies => the CUDA kernel must be BIT-EXACT.
Pinned against the reference's Triton kernel under TRITON_INTERPRET=1
(tests/golden/scatter_*.pt).
"""
import torch


def kv_scatter(key: torch.Tensor, value: torch.Tensor, k_cache: torch.Tensor,
               v_cache: torch.Tensor, slot_mapping: torch.Tensor) -> None:
    """In place: cache.view(-1, Hkv*D)[slot_mapping[i]] = x[i] for every i with slot != -1.

    key/value: [T, Hkv, D] (value may be a strided view, attention.py:121-122);
    caches: [num_blocks, block_size, Hkv, D] contiguous (attention.py:123);
    slot_mapping: int32 [T] (attention.py:124); slot -1 is skipped (attention.py:79-80).
    Later rows win on duplicate slots only by program order in the reference; the
    reference never produces duplicates (model_runner.py:466-481), so tests do not either.
    """
    T, H, D = key.shape
    row = H * D
    assert key.stride(-1) == 1 and value.stride(-1) == 1
    assert key.stride(1) == D and value.stride(1) == D
    assert k_cache.stride(1) == row and v_cache.stride(1) == row
    assert slot_mapping.numel() == T
    kc = k_cache.view(-1, row)
    vc = v_cache.view(-1, row)
    slots = slot_mapping.to(torch.int64)
    keep = slots != -1
    idx = slots[keep]
    kc[idx] = key.reshape(T, row)[keep]
    vc[idx] = value.reshape(T, row)[keep]
