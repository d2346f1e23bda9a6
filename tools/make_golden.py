#!/usr/bin/env python
"""Generate tests/golden/{scatter,pointwise}_*.pt by running the REFERENCE's own code on CPU.

Run in the build container only (needs /root/reference):
    python tools/make_golden.py

  * scatter_*.pt ..... nanovllm.layers.attention.store_kvcache (the Triton kernel,
                       attention.py:34-127) executed with TRITON_INTERPRET=1 on CPU tensors.
  * pointwise_*.pt ... nanovllm.layers.{layernorm.RMSNorm, rotary_embedding.RotaryEmbedding,
                       activation.SiluAndMul} executed eagerly (TORCHDYNAMO_DISABLE=1; the CPU
                       inductor backend cannot compile in this image).
Only inputs and the reference's outputs are stored -- no reference source.
"""
import os
import sys

os.environ["TRITON_INTERPRET"] = "1"
os.environ["TORCHDYNAMO_DISABLE"] = "1"
REF = os.environ.get("NANOVLLM_REFERENCE", "/root/reference")
sys.path.insert(0, REF)

import torch  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def scatter_case(name, seed, T, Hq, Hkv, D, num_blocks, slots):
    from nanovllm.layers.attention import store_kvcache
    g = torch.Generator().manual_seed(seed)
    qkv = torch.randn(T, (Hq + 2 * Hkv) * D, generator=g).bfloat16()
    q, k, v = qkv.split([Hq * D, Hkv * D, Hkv * D], dim=-1)
    k = k.reshape(T, Hkv, D).contiguous()        # rotary output is a fresh tensor
    v = v.view(T, Hkv, D)                         # strided view of qkv
    k_cache = torch.randn(num_blocks, 256, Hkv, D, generator=g).bfloat16()
    v_cache = torch.randn(num_blocks, 256, Hkv, D, generator=g).bfloat16()
    slot_mapping = torch.tensor(slots, dtype=torch.int32)
    k0, v0 = k_cache.clone(), v_cache.clone()
    store_kvcache(k, v, k_cache, v_cache, slot_mapping)
    # store only the touched rows + a checksum of the whole cache to keep fixtures small
    touched = sorted({s for s in slots if s >= 0})
    torch.save({
        "seed": seed, "T": T, "Hq": Hq, "Hkv": Hkv, "D": D, "num_blocks": num_blocks,
        "slot_mapping": slot_mapping,
        "touched": torch.tensor(touched, dtype=torch.int64),
        "k_rows": k_cache.view(-1, Hkv * D)[touched].clone(),
        "v_rows": v_cache.view(-1, Hkv * D)[touched].clone(),
        "k_untouched_equal": bool(_untouched_equal(k_cache, k0, touched)),
        "v_untouched_equal": bool(_untouched_equal(v_cache, v0, touched)),
    }, os.path.join(OUT, f"scatter_{name}.pt"))
    print("wrote scatter", name)


def _untouched_equal(after, before, touched):
    mask = torch.ones(after.shape[0] * after.shape[1], dtype=torch.bool)
    mask[touched] = False
    D = after.shape[2] * after.shape[3]
    return torch.equal(after.view(-1, D)[mask], before.view(-1, D)[mask])


def pointwise_case(name, seed, T, Hq, Hkv, D, inter, max_pos, base=1e6, eps=1e-6):
    from nanovllm.layers.layernorm import RMSNorm
    from nanovllm.layers.rotary_embedding import RotaryEmbedding
    from nanovllm.layers.activation import SiluAndMul
    g = torch.Generator().manual_seed(seed)
    torch.set_default_dtype(torch.bfloat16)      # model_runner.py:126
    qn, kn = RMSNorm(D, eps), RMSNorm(D, eps)
    hn = RMSNorm(Hq * D // 2, eps)
    rot = RotaryEmbedding(D, D, max_pos, base)
    torch.set_default_dtype(torch.float32)
    with torch.no_grad():
        qn.weight.copy_((1 + 0.1 * torch.randn(D, generator=g)).bfloat16())
        kn.weight.copy_((1 + 0.1 * torch.randn(D, generator=g)).bfloat16())
        hn.weight.copy_((1 + 0.1 * torch.randn(Hq * D // 2, generator=g)).bfloat16())
    qkv = (3 * torch.randn(T, (Hq + 2 * Hkv) * D, generator=g)).bfloat16()
    positions = torch.randint(0, max_pos, (T,), generator=g, dtype=torch.int64)
    q, k, v = qkv.split([Hq * D, Hkv * D, Hkv * D], dim=-1)
    with torch.no_grad():
        q1 = qn(q.view(T, Hq, D))
        k1 = kn(k.view(T, Hkv, D))
        q2, k2 = rot(positions, q1, k1)
        x = (2 * torch.randn(T, 2 * inter, generator=g)).bfloat16()
        act = SiluAndMul()(x)
        h = torch.randn(T, Hq * D // 2, generator=g).bfloat16()
        r = torch.randn(T, Hq * D // 2, generator=g).bfloat16()
        h_norm = hn(h.clone())
        hr_norm, hr_res = hn(h.clone(), r.clone())
    torch.save({
        "seed": seed, "T": T, "Hq": Hq, "Hkv": Hkv, "D": D, "inter": inter, "max_pos": max_pos,
        "base": base, "eps": eps,
        "qkv": qkv, "positions": positions, "q_weight": qn.weight.data.clone(),
        "k_weight": kn.weight.data.clone(), "cos_sin_cache_sample_pos": positions[:4].clone(),
        "cos_sin_cache_sample": rot.cos_sin_cache[positions[:4]].clone(),
        "q_normed": q1, "k_normed": k1, "q_out": q2, "k_out": k2,
        "act_in": x, "act_out_eager": act,
        "h": h, "r": r, "h_weight": hn.weight.data.clone(), "h_norm": h_norm,
        "hr_norm": hr_norm, "hr_res": hr_res,
    }, os.path.join(OUT, f"pointwise_{name}.pt"))
    print("wrote pointwise", name)


def prepare_case(name, seed, specs):
    """Run the reference's ModelRunner.prepare_model_input (model_runner.py:400-495) on CPU: the only
    CUDA-specific bits are `pin_memory=True` and `.cuda()`, shimmed to no-ops here.
    specs: list of (num_tokens, num_cached_tokens, num_new_tokens, has_block_table)."""
    import random
    from types import SimpleNamespace
    from nanovllm.engine.model_runner import ModelRunner
    from nanovllm.engine.sequence import Sequence
    from nanovllm.utils.context import get_context, reset_context
    rnd = random.Random(seed)
    real_tensor, real_cuda = torch.tensor, torch.Tensor.cuda
    torch.tensor = lambda *a, pin_memory=False, **k: real_tensor(*a, **k)
    torch.Tensor.cuda = lambda self, *a, **k: self
    try:
        seqs, states = [], []
        free = list(range(4096))
        rnd.shuffle(free)
        for (ntok, cached, new, has_bt) in specs:
            s = Sequence([rnd.randrange(0, 150000) for _ in range(ntok)])
            s.num_cached_tokens, s.num_new_tokens = cached, new
            nblk = (cached + new + 255) // 256
            s.block_table = [free.pop() for _ in range(nblk)] if has_bt else []
            seqs.append(s)
            states.append(dict(token_ids=list(s.token_ids), num_cached_tokens=cached, num_new_tokens=new,
                               block_table=list(s.block_table)))
        fake = SimpleNamespace(block_size=256)
        fake.prepare_block_tables = lambda sq: ModelRunner.prepare_block_tables(fake, sq)
        input_ids, positions = ModelRunner.prepare_model_input(fake, seqs)
        c = get_context()
        rec = dict(states=states, input_ids=input_ids, positions=positions, cu_seqlens_q=c.cu_seqlens_q,
                   cu_seqlens_k=c.cu_seqlens_k, max_seqlen_q=c.max_seqlen_q, max_seqlen_k=c.max_seqlen_k,
                   slot_mapping=c.slot_mapping, context_lens=c.context_lens, block_tables=c.block_tables,
                   seq_need_compute_logits=c.seq_need_compute_logits)
        reset_context()
    finally:
        torch.tensor, torch.Tensor.cuda = real_tensor, real_cuda
    torch.save(rec, os.path.join(OUT, f"prepare_{name}.pt"))
    print("wrote prepare", name)


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    # prefill-like contiguous run, decode-like single rows, graph padding (-1), page crossing
    scatter_case("tp1_prefill", 1, 40, 16, 8, 128, 4, list(range(500, 540)))
    scatter_case("tp8_mixed", 2, 9, 2, 1, 128, 3, [767, -1, 0, 255, 256, -1, 511, 300, 1])
    scatter_case("tp2_decode", 3, 6, 8, 4, 128, 2, [5, 261, 17, -1, 400, 129])
    # (len(seq), num_cached_tokens, num_new_tokens, has_block_table)
    prepare_case("pure_prefill", 21, [(128, 0, 128, True), (1, 0, 1, True), (300, 0, 300, True), (257, 0, 257, True)])
    prepare_case("mixed_chunked", 22, [(700, 512, 188, True), (130, 129, 1, True), (2000, 0, 128, True), (1, 0, 1, True),
                                       (258, 257, 1, True), (900, 256, 257, True), (64, 0, 64, True), (1040, 1000, 40, True),
                                       (513, 512, 1, True), (256, 0, 256, True)])
    prepare_case("decode", 23, [(4096, 4095, 1, True), (2, 1, 1, True), (257, 256, 1, True), (1301, 1300, 1, True)])
    prepare_case("warmup", 24, [(600, 0, 600, False), (600, 0, 600, False)])
    pointwise_case("q0p6b", 11, 7, 16, 8, 128, 96, 4096)
    pointwise_case("tp8_14b", 12, 5, 5, 1, 128, 64, 40960)
