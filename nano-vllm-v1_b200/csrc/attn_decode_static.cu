// b200_attn_decode_static_scratch: split-K paged decode for callers that cannot allocate device memory.
//
// Same kernel and same arithmetic as b200_attn_decode (attn_decode.cu); the only difference is where the
// split-K scratch (arrival tickets + fp32 partial outputs + fp32 log-sum-exps) lives.  b200_attn_decode
// takes it from the caller; here it is a fixed array that is part of the library image, so the call needs
// no cudaMalloc -- not by the caller, not by us.
//
// Who needs that: a serving engine that has already handed every free byte of HBM to the KV cache and so
// cannot count on a later cudaMalloc succeeding; a host binding (cgo / JNI / bare ctypes) that owns no device
// allocator and is not allowed to create one inside someone else's memory pool; a request loop in which an
// allocation would stall the whole device (cudaMalloc / cudaFree synchronise it).
//
// The price: every GPU on which this library is loaded carries B200_DECODE_STATIC_SCRATCH_BYTES of device
// memory from module load until the process exits, used or not, and all calls on one device share it.
#include "../../include/b200attn.h"

#include <cuda_runtime.h>

#include <mutex>

#include "launch.h"

namespace {

// The scratch itself.  In HOST code this name is only a handle for the runtime's symbol table -- it is NOT
// the array's device address and must never be handed to a kernel as a pointer.  The per-device address
// comes from cudaGetSymbolAddress (static_scratch_base below).
static __device__ __align__(256) unsigned char g_decode_static_scratch[B200_DECODE_STATIC_SCRATCH_BYTES];

std::mutex g_setup_mutex;
unsigned char* g_scratch_base[64] = {};          // per device ordinal; nullptr = not looked up yet

// Device address of g_decode_static_scratch on the CURRENT device.  Looked up once per device, with that
// device current (each device has its own copy of the array at its own address).  No allocation happens:
// the array was placed when the module was loaded onto the device.
int static_scratch_base(unsigned char** base_out) {
    const int dev = b200::current_device();
    if (dev < 0 || dev >= 64) return B200_ERR_RANGE;
    std::lock_guard<std::mutex> lock(g_setup_mutex);
    if (g_scratch_base[dev] == nullptr) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_decode_static_scratch);
        if (e != cudaSuccess) { (void)cudaGetLastError(); return (int)e; }
        if (p == nullptr) return (int)cudaErrorInvalidSymbol;
        g_scratch_base[dev] = static_cast<unsigned char*>(p);
    }
    *base_out = g_scratch_base[dev];
    return B200_OK;
}

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }
inline bool stride_ok(int64_t elems) { return ((elems * 2) & 15) == 0; }

}  // namespace

extern "C" int b200_attn_decode_static_scratch(const void* q, int64_t q_token_stride, const void* k_cache,
                                               const void* v_cache, void* out, int64_t out_token_stride,
                                               const int32_t* context_lens, const int32_t* block_tables,
                                               int32_t bt_row_stride, int32_t batch, int32_t num_heads,
                                               int32_t num_kv_heads, int32_t head_dim, int32_t page_size,
                                               float softmax_scale, int32_t num_splits, void* stream) {
    using namespace b200;
    // Same argument checks, in the same order, as b200_attn_decode.
    if (batch < 0 || num_splits < 1 || num_splits > 256 || bt_row_stride < 0) return B200_ERR_RANGE;
    if (batch == 0) return B200_OK;
    if (!q || !k_cache || !v_cache || !out || !context_lens || !block_tables) return B200_ERR_NULL;
    if (head_dim != B200_HD_CONST || page_size != 256 || num_kv_heads <= 0 || num_heads % num_kv_heads != 0 ||
        num_heads / num_kv_heads > 16)
        return B200_ERR_SHAPE;
    if (!aligned16(q) || !aligned16(k_cache) || !aligned16(v_cache) || !aligned16(out) ||
        !stride_ok(q_token_stride) || !stride_ok(out_token_stride))
        return B200_ERR_ALIGN;
    // The fixed array is all there is: a problem that does not fit is refused, never truncated.
    const size_t need = b200_decode_workspace_bytes(batch, num_heads, num_splits);   // 0 when num_splits == 1
    if (need > (size_t)B200_DECODE_STATIC_SCRATCH_BYTES) return B200_ERR_SCRATCH_LIMIT;

    DecodeArgs a;
    a.q = q; a.q_token_stride = q_token_stride; a.k_cache = k_cache; a.v_cache = v_cache;
    a.out = out; a.out_token_stride = out_token_stride; a.context_lens = context_lens;
    a.block_tables = block_tables; a.bt_row_stride = bt_row_stride;
    a.batch = batch; a.Hq = num_heads; a.Hkv = num_kv_heads; a.scale = softmax_scale; a.num_splits = num_splits;
    a.ws_o = nullptr; a.ws_lse = nullptr; a.tickets = nullptr;
    if (num_splits > 1) {
        unsigned char* base = nullptr;
        const int rc = static_scratch_base(&base);
        if (rc != B200_OK) return rc;
        // Layout of b200_attn_decode's workspace: [ tickets (DEC_TICKET_BYTES) | partial O fp32 | LSE fp32 ].
        a.ws_o = reinterpret_cast<float*>(base + DEC_TICKET_BYTES);
        a.ws_lse = a.ws_o + (size_t)batch * num_splits * num_heads * B200_HD_CONST;
        const size_t ticket_bytes = (size_t)batch * num_kv_heads * sizeof(int32_t);
        if (ticket_bytes <= DEC_TICKET_BYTES) {
            // The array outlives every call and keeps whatever the previous one left.  The partial O / LSE
            // regions are fully rewritten before they are read; the arrival tickets are the one part that
            // must start at zero.  A launch that ran to completion leaves them zero, one that was cut short
            // by an error need not, so clear the words this call uses.
            const cudaError_t e = cudaMemsetAsync(base, 0, ticket_bytes, (cudaStream_t)stream);
            if (e != cudaSuccess) return (int)e;
            a.tickets = reinterpret_cast<int32_t*>(base);
        }                                            // else: no tickets => launch_attn_decode adds the separate merge launch
    }
    return launch_attn_decode(a, (cudaStream_t)stream);
}
