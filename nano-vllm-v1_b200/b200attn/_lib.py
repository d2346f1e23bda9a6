"""ctypes binding of include/b200attn.h.  Fails loudly when the library is absent."""
import ctypes
import os
from ctypes import POINTER, c_char_p, c_float, c_int32, c_int64, c_size_t, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

# name -> (restype, argtypes).  Kept in sync with include/b200attn.h; tests/test_abi.py checks
# that every symbol the header declares is exported and listed here.
SIGNATURES = {
    "b200_version": (c_int32, []),
    "b200_status_string": (c_char_p, [c_int32]),
    "b200_sm_count": (c_int32, []),
    "b200_debug_prefill_trace": (c_int32, [c_void_p, c_int32]),
    "b200_ar_buffer_bytes": (c_size_t, [c_size_t]),
    "b200_ar_alloc": (c_int32, [c_size_t, POINTER(c_void_p), c_void_p]),
    "b200_ar_open": (c_int32, [c_void_p, POINTER(c_void_p)]),
    "b200_ar_create": (c_int32, [c_int32, c_int32, POINTER(c_void_p), c_size_t, POINTER(c_void_p)]),
    "b200_allreduce_bf16": (c_int32, [c_void_p, c_void_p, c_int64, c_int32, c_void_p]),
    "b200_ar_destroy": (c_int32, [c_void_p]),
    "b200_ar_status": (c_int32, [c_void_p, POINTER(c_int32)]),
    "b200_nvls_buffer_bytes": (c_size_t, [c_size_t]),
    "b200_nvls_create": (c_int32, [c_int32, c_int32, c_void_p, c_void_p, POINTER(c_void_p), c_size_t, POINTER(c_void_p)]),
    "b200_nvls_allreduce_add_rmsnorm": (c_int32, [c_void_p, c_void_p, c_void_p, c_void_p, c_float, c_void_p,
                                                  c_int32, c_int32, c_int32, c_void_p]),
    "b200_nvls_status": (c_int32, [c_void_p, POINTER(c_int32)]),
    "b200_nvls_destroy": (c_int32, [c_void_p]),
    "b200_kv_scatter": (c_int32, [c_void_p, c_void_p, c_int64, c_int64, c_void_p, c_void_p, c_void_p,
                                  c_int32, c_int32, c_int32, c_void_p]),
    "b200_decode_plan": (c_int32, [c_int32, c_int32, c_int32]),
    "b200_decode_workspace_bytes": (c_size_t, [c_int32, c_int32, c_int32]),
    "b200_attn_decode": (c_int32, [c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_int64, c_void_p,
                                   c_void_p, c_int32, c_int32, c_int32, c_int32, c_int32, c_int32, c_float,
                                   c_int32, c_void_p, c_size_t, c_void_p]),
    "b200_attn_decode_static_scratch": (c_int32, [c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_int64, c_void_p,
                                                  c_void_p, c_int32, c_int32, c_int32, c_int32, c_int32, c_int32,
                                                  c_float, c_int32, c_void_p]),
    "b200_attn_prefill": (c_int32, [c_void_p, c_int64, c_void_p, c_void_p, c_int64, c_int64, c_void_p, c_int64,
                                    c_void_p, c_void_p, c_void_p, c_int32, c_int32, c_int32, c_int32, c_int32,
                                    c_int32, c_int32, c_int32, c_int32, c_int64, c_float, c_int32, c_void_p]),
    "b200_attn_plan_bytes": (c_size_t, [c_int32, c_int32]),
    "b200_attn_plan": (c_int32, [c_void_p, c_void_p, c_int32, c_int32, c_int32, c_int32, c_int32, c_void_p, c_size_t, c_void_p]),
    "b200_attn_prefill_planned": (c_int32, [c_void_p, c_int64, c_void_p, c_void_p, c_int64, c_int64, c_void_p, c_int64,
                                            c_void_p, c_void_p, c_void_p, c_int32, c_int32, c_int32, c_int32, c_int32,
                                            c_int32, c_int32, c_int32, c_int32, c_int64, c_float, c_void_p, c_size_t, c_void_p]),
    "b200_attn_decode_varlen": (c_int32, [c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_int64, c_void_p, c_void_p,
                                          c_void_p, c_int32, c_int32, c_int32, c_int32, c_int32, c_int32, c_float,
                                          c_int32, c_void_p, c_size_t, c_void_p, c_size_t, c_void_p]),
    "b200_qk_norm_rope": (c_int32, [c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_void_p,
                                    c_float, c_void_p, c_void_p, c_int32, c_int32, c_int32, c_int32, c_void_p]),
    "b200_qk_norm_rope_store": (c_int32, [c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_void_p,
                                          c_void_p, c_void_p, c_float, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                          c_int32, c_int32, c_int32, c_int32, c_void_p]),
    "b200_rmsnorm": (c_int32, [c_void_p, c_void_p, c_void_p, c_float, c_void_p, c_int32, c_int32, c_void_p]),
    "b200_silu_and_mul": (c_int32, [c_void_p, c_void_p, c_int32, c_int32, c_void_p]),
}


def lib_path() -> str:
    return os.environ.get("B200ATTN_LIB", os.path.join(_HERE, "libb200attn.so"))


def load_library():
    """dlopen libb200attn.so (built in-tree by `__graft_entry__.build()` / csrc/Makefile)."""
    global _LIB
    if _LIB is not None:
        return _LIB
    path = lib_path()
    if not os.path.exists(path):
        raise RuntimeError(
            f"b200attn: CUDA library not found at {path}. Build it with "
            f"`python -c 'import __graft_entry__ as g; g.build()'` or `make -C nano-vllm-v1_b200/csrc`. "
            f"There is no CPU fallback for this path.")
    lib = ctypes.CDLL(path)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if the symbol is missing: loud by design
        fn.restype = res
        fn.argtypes = args
    _LIB = lib
    return lib


def check(status: int, what: str) -> None:
    if status != 0:
        msg = load_library().b200_status_string(status).decode()
        raise RuntimeError(f"{what} failed with status {status}: {msg}")
