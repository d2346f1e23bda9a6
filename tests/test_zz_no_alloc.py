"""GPU test of b200_attn_decode_static_scratch, the decode entry point for callers that cannot allocate.

Some users of libb200attn.so have no way to obtain device memory at call time: an engine that gave all
free HBM to its KV cache, a cgo/JNI/ctypes host without a device allocator, a loop that must not stall on
cudaMalloc.  For them the library carries its split-K scratch as a fixed 8 MiB device array and this entry
point uses it instead of a caller workspace.  The test calls it straight through the C ABI (ctypes, raw
pointers) and compares with b200_attn_decode -- same kernel, caller-provided workspace -- on the same
inputs, at the attention tolerance of tests/test_gpu_kernels.py.

Two calls in a row with DIFFERENT shapes: the array is shared and keeps the first call's partial sums and
tickets, so the second call is the one that shows stale contents are handled.  The file sorts last on
purpose (pytest -x would hide everything after a failure here); it needs nothing from the other tests and
passes alone in a fresh process as well as at the end of the whole suite.
"""
import pytest
import torch

from helpers import make_paged_case

pytestmark = pytest.mark.gpu
ATOL = 2e-2                      # attention bar of tests/test_gpu_kernels.py (BASELINE.json north_star)
SCALE = 128 ** -0.5
PAGE = 256
SCRATCH_LIMIT = 8 * 1024 * 1024  # B200_DECODE_STATIC_SCRATCH_BYTES in include/b200attn.h
ERR_SCRATCH_LIMIT = -7


@pytest.fixture(scope="module")
def lib():
    assert torch.cuda.is_available()
    import b200attn
    return b200attn.load_library()


def _case(seed, lens_k, Hq, Hkv):
    c = make_paged_case(seed, [1] * len(lens_k), lens_k, Hq, Hkv, pad=-1)
    return {k: (v.cuda() if isinstance(v, torch.Tensor) else v) for k, v in c.items()}


def _call(lib, fn_name, c, Hq, Hkv, splits, out, workspace=None):
    """One raw C-ABI call; `out` (and `workspace`) are preallocated by the test, nothing is allocated here."""
    q, bt = c["q"], c["block_tables"]
    B = q.shape[0]
    stream = torch.cuda.current_stream().cuda_stream
    head = (q.data_ptr(), q.stride(0), c["k_cache"].data_ptr(), c["v_cache"].data_ptr(), out.data_ptr(), out.stride(0),
            c["context_lens"].data_ptr(), bt.data_ptr(), bt.stride(0), B, Hq, Hkv, 128, PAGE, SCALE, splits)
    if fn_name == "b200_attn_decode":
        return lib.b200_attn_decode(*head, workspace.data_ptr(), workspace.numel(), stream)
    return lib.b200_attn_decode_static_scratch(*head, stream)


def _reference(lib, c, Hq, Hkv, splits):
    """The existing entry point with a zeroed caller workspace (its documented contract)."""
    B = c["q"].shape[0]
    ws = torch.zeros(lib.b200_decode_workspace_bytes(B, Hq, splits) + 256, dtype=torch.uint8, device="cuda")
    out = torch.empty_like(c["q"])
    assert _call(lib, "b200_attn_decode", c, Hq, Hkv, splits, out, ws) == 0
    torch.cuda.synchronize()
    return out


def test_static_scratch_decode_matches_workspace_decode_twice_in_a_row(lib):
    # call 1: 6 sequences, ragged contexts (one empty, one a single token), 32q/8kv, 4 splits
    # call 2: other batch, other heads, other split count -> every region of the shared array is laid out
    #         differently and still holds call 1's data when call 2 starts
    shapes = [(11, [1000, 257, 0, 4096, 1, 777], 32, 8, 4),
              (12, [3000, 512, 2049], 16, 2, 7)]
    cases = [(_case(seed, lens, Hq, Hkv), Hq, Hkv, s) for seed, lens, Hq, Hkv, s in shapes]
    want = [_reference(lib, c, Hq, Hkv, s) for c, Hq, Hkv, s in cases]
    outs = [torch.full_like(c["q"], float("nan")) for c, _, _, _ in cases]

    c, Hq, Hkv, s = cases[0]
    assert lib.b200_decode_workspace_bytes(c["q"].shape[0], Hq, s) <= SCRATCH_LIMIT
    assert _call(lib, "b200_attn_decode_static_scratch", c, Hq, Hkv, s, outs[0]) == 0     # first call: per-device set-up
    torch.cuda.synchronize()

    # the second call must not touch any allocator: torch's stays put and the device's free memory does not shrink
    torch_before = torch.cuda.memory_allocated()
    free_before, _ = torch.cuda.mem_get_info()
    c, Hq, Hkv, s = cases[1]
    assert _call(lib, "b200_attn_decode_static_scratch", c, Hq, Hkv, s, outs[1]) == 0
    torch.cuda.synchronize()
    free_after, _ = torch.cuda.mem_get_info()
    assert torch.cuda.memory_allocated() == torch_before
    assert free_after >= free_before

    for got, ref in zip(outs, want):
        assert not torch.isnan(got.float()).any()
        assert (got.float() - ref.float()).abs().max().item() <= ATOL

    # and once more with the first shape: call 2's leftovers must not leak into it either
    c, Hq, Hkv, s = cases[0]
    again = torch.full_like(c["q"], float("nan"))
    assert _call(lib, "b200_attn_decode_static_scratch", c, Hq, Hkv, s, again) == 0
    torch.cuda.synchronize()
    assert (again.float() - want[0].float()).abs().max().item() <= ATOL


def test_static_scratch_decode_refuses_what_does_not_fit(lib):
    """Beyond the 8 MiB array the call is refused with its own status code and launches nothing."""
    Hq, Hkv, splits = 32, 8, 64
    c = _case(13, [300] * 16, Hq, Hkv)
    assert lib.b200_decode_workspace_bytes(16, Hq, splits) > SCRATCH_LIMIT
    out = torch.full_like(c["q"], 7.0)
    assert _call(lib, "b200_attn_decode_static_scratch", c, Hq, Hkv, splits, out) == ERR_SCRATCH_LIMIT
    torch.cuda.synchronize()
    assert (out == 7.0).all()
    msg = lib.b200_status_string(ERR_SCRATCH_LIMIT).decode()
    assert "built-in scratch" in msg and "b200_attn_decode" in msg
    # one split needs no scratch at all and is always accepted
    assert _call(lib, "b200_attn_decode_static_scratch", c, Hq, Hkv, 1, out) == 0
    torch.cuda.synchronize()
    ref = _reference(lib, c, Hq, Hkv, 2)
    assert (out.float() - ref.float()).abs().max().item() <= ATOL
