"""CPU tests: the C-ABI library builds, loads, and exports exactly what include/b200attn.h
declares (no compute calls -- there is no GPU here)."""
import os
import re
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    sys.path.insert(0, ROOT)
    import __graft_entry__ as g
    g.build()
    import b200attn
    return b200attn.load_library()


def _header_symbols():
    src = open(os.path.join(ROOT, "include", "b200attn.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_all_exported_and_bound(lib):
    from b200attn._lib import SIGNATURES, lib_path
    syms = _header_symbols()
    assert len(syms) >= 11
    out = subprocess.run(["nm", "-D", "--defined-only", lib_path()], capture_output=True, text=True, check=True).stdout
    exported = set(re.findall(r" T (b200_[a-z0-9_]+)", out))
    for s in syms:
        assert s in exported, f"{s} declared in the header but not exported"
        assert s in SIGNATURES, f"{s} has no ctypes signature"
        assert getattr(lib, s) is not None
    assert set(SIGNATURES) <= set(syms)


def test_version_and_status_strings(lib):
    assert lib.b200_version() == 100
    assert lib.b200_status_string(0) == b"ok"
    for code in range(-7, 0):
        assert lib.b200_status_string(code).startswith(b"b200attn:")


def test_argument_errors_without_gpu(lib):
    # validation happens before any CUDA call
    assert lib.b200_kv_scatter(None, None, 0, 0, None, None, None, 4, 8, 128, None) == -1
    assert lib.b200_kv_scatter(None, None, 0, 0, None, None, None, 0, 8, 128, None) == 0     # empty batch ok
    assert lib.b200_kv_scatter(16, 16, 1024, 1024, 16, 16, 16, 4, 8, 64, None) == -2         # head_dim != 128
    assert lib.b200_kv_scatter(16, 8, 1024, 1024, 16, 16, 16, 4, 8, 128, None) == -3         # misaligned value
    assert lib.b200_attn_decode(16, 128, 16, 16, 16, 128, 16, 16, 4, 2, 6, 4, 128, 256, 1.0, 1, None, 0, None) == -2
    assert lib.b200_attn_decode(16, 128, 16, 16, 16, 128, 16, 16, 4, 2, 8, 4, 128, 256, 1.0, 4, None, 0, None) == -4
    assert lib.b200_silu_and_mul(16, 16, 3, 100, None) == -2
    # the no-allocation decode refuses, before touching CUDA, what its built-in 8 MiB scratch cannot hold
    assert lib.b200_decode_workspace_bytes(16, 32, 64) > 8 * 1024 * 1024
    assert lib.b200_attn_decode_static_scratch(16, 128, 16, 16, 16, 128, 16, 16, 4, 16, 32, 8, 128, 256, 1.0, 64, None) == -7
    assert lib.b200_attn_decode_static_scratch(16, 128, 16, 16, 16, 128, 16, 16, 4, 2, 6, 4, 128, 256, 1.0, 1, None) == -2


def test_decode_plan_is_static_and_bounded(lib):
    assert lib.b200_decode_plan(512, 8, 4096) == 1          # cfg3: enough (seq, kv-head) items
    s = lib.b200_decode_plan(64, 1, 32832)                   # cfg5 decode: few items, long context
    assert 2 <= s <= 64
    assert lib.b200_decode_plan(1, 8, 300) <= 2              # never finer than one page per split
    # round 2 (profiles/r2_sweep_decode_tp.txt): an HBM-bound launch with at least one full wave of items is NOT split to
    # "fill the last wave" -- cfg3 at the TP2/TP4/TP8 heads runs unsplit (TP8: 158 us unsplit vs 172 us with 4 splits)
    for hkv in (4, 2, 1):
        assert lib.b200_decode_plan(512, hkv, 4096) == 1
    assert lib.b200_decode_plan(256, 2, 1280) == 1
    assert lib.b200_decode_plan(1, 8, 4096) >= 8             # a lone sequence is spread over the machine
    assert lib.b200_decode_plan(512, 1, 4096) == lib.b200_decode_plan(512, 1, 4096)   # pure function of its arguments
    assert lib.b200_decode_workspace_bytes(4, 32, 1) == 0
    assert lib.b200_decode_workspace_bytes(4, 32, 3) == 64 * 1024 + 4 * 3 * 32 * 129 * 4   # tickets + partial O + LSE


def test_ops_refuse_cpu_tensors(lib):
    import torch
    from b200attn import ops
    k = torch.zeros(2, 1, 128, dtype=torch.bfloat16)
    kc = torch.zeros(1, 256, 1, 128, dtype=torch.bfloat16)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.store_kvcache(k, k, kc, kc, torch.zeros(2, dtype=torch.int32))
    with pytest.raises(AssertionError):
        ops.store_kvcache(k, k, kc, kc, torch.zeros(3, dtype=torch.int32))


def test_missing_library_fails_loudly(monkeypatch):
    import b200attn._lib as L
    monkeypatch.setattr(L, "_LIB", None)
    monkeypatch.setenv("B200ATTN_LIB", "/nonexistent/libb200attn.so")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        L.load_library()


def test_context_mirror_matches_reference_field_order():
    from b200attn.utils.context import Context, get_context, reset_context, set_context
    import dataclasses
    names = [f.name for f in dataclasses.fields(Context)]
    assert names == ["cu_seqlens_q", "cu_seqlens_k", "max_seqlen_q", "max_seqlen_k", "slot_mapping",
                     "context_lens", "block_tables", "seq_need_compute_logits"]
    set_context(1, 2, 3, 4, 5, 6, 7, 8)
    c = get_context()
    assert (c.cu_seqlens_q, c.max_seqlen_k, c.seq_need_compute_logits) == (1, 4, 8)
    reset_context()
    assert get_context().cu_seqlens_q is None
