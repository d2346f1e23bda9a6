#!/usr/bin/env python
"""Generate tests/golden/attn_ref_*.pt ON THE GPU BOX by running the UNMODIFIED reference:
`nanovllm.layers.attention.Attention.forward` (Triton store_kvcache + flash-attn 2.8.3) from the
offline install under baseline/_ref (made in the build container with
`pip install --no-deps --target baseline/_ref /root/reference`; git-ignored, travels with gpurun).

    gpurun -- python tools/make_golden_gpu.py        # writes gpurun_out/golden/attn_ref_*.pt
    cp gpurun_out/golden/attn_ref_*.pt tests/golden/ # in the build container, then commit

Only the case description (seed, shapes), the reference's output and a digest of the cache after
the step are stored -- inputs are re-created from the seed by tests/helpers.make_forward_case.
This pins oracle/paged_attention.py to the reference (tests/test_oracle_cpu.py) without a GPU.
"""
import hashlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "baseline", "_ref"))
sys.path.insert(0, os.path.join(ROOT, "tests"))

import torch  # noqa: E402

from helpers import FORWARD_CASES, make_forward_case  # noqa: E402


def digest(t: torch.Tensor) -> str:
    return hashlib.sha256(t.contiguous().view(torch.int16).cpu().numpy().tobytes()).hexdigest()


def main():
    from nanovllm.layers.attention import Attention            # the reference's own module
    from nanovllm.utils.context import reset_context, set_context
    import flash_attn
    out_dir = os.path.join(ROOT, "gpurun_out", "golden")
    os.makedirs(out_dir, exist_ok=True)
    dev = torch.device("cuda:0")
    for name, meta in FORWARD_CASES.items():
        c = make_forward_case(meta)
        Hq, Hkv = meta["Hq"], meta["Hkv"]
        attn = Attention(Hq, 128, 128 ** -0.5, Hkv)
        attn.k_cache, attn.v_cache = c["k_cache"].to(dev), c["v_cache"].to(dev)
        qkv = c["qkv"].to(dev)
        n = qkv.shape[0]
        v = qkv[:, (Hq + Hkv) * 128:].view(n, Hkv, 128)         # strided view, as qwen3.py hands it over
        set_context(*[t.to(dev) if isinstance(t, torch.Tensor) else t for t in c["ctx"]])
        with torch.inference_mode():
            o = attn(c["q"].to(dev), c["k"].to(dev), v)
        reset_context()
        torch.cuda.synchronize()
        rec = dict(meta, name=name, out=o.view(n, Hq, 128).cpu().clone(),
                   k_cache_sha256=digest(attn.k_cache), v_cache_sha256=digest(attn.v_cache),
                   flash_attn=str(flash_attn.__version__), torch=str(torch.__version__), gpu=torch.cuda.get_device_name(0))
        torch.save(rec, os.path.join(out_dir, f"attn_ref_{name}.pt"))
        print("wrote", name, tuple(o.shape), flush=True)


if __name__ == "__main__":
    main()
