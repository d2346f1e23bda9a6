// extern "C" surface declared in include/b200attn.h: argument validation + dispatch.
// No torch headers, no exceptions across the boundary, no allocation, no host sync.
#include "../../include/b200attn.h"

#include <cuda_runtime.h>
#include <stdlib.h>
#include <string.h>

#include <math.h>
#include "launch.h"

namespace b200 {
int current_device() {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) { (void)cudaGetLastError(); return -1; }
    return dev;
}

int sm_count() {
    static int cached[64] = {0};                                  // per device: a process may touch several GPUs
    const int dev = current_device();
    if (dev >= 0 && dev < 64 && cached[dev] > 0) return cached[dev];
    int n = 0;
    if (dev < 0 || cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
        n = 148;                                                  // B200
    if (dev >= 0 && dev < 64) cached[dev] = n;
    return n;
}

// B200ATTN_DECODE_FUSED_REDUCE=0 restores the separate decode_reduce_kernel launch (A/B knob).
static bool decode_fused_reduce() {
    static int v = -1;
    if (v < 0) {
        const char* env = getenv("B200ATTN_DECODE_FUSED_REDUCE");
        v = (env && env[0] == '0') ? 0 : 1;
    }
    return v == 1;
}

// Carve the caller's split-K workspace: [ tickets (DEC_TICKET_BYTES, zero) | partial O fp32 | LSE fp32 ].
static void carve_decode_workspace(DecodeArgs& a, void* workspace) {
    a.ws_o = nullptr; a.ws_lse = nullptr; a.tickets = nullptr;
    if (!workspace) return;
    a.ws_o = (float*)((char*)workspace + DEC_TICKET_BYTES);
    a.ws_lse = a.ws_o + (size_t)a.batch * a.num_splits * a.Hq * B200_HD_CONST;
    if (decode_fused_reduce() && (size_t)a.batch * a.Hkv * sizeof(int32_t) <= DEC_TICKET_BYTES)
        a.tickets = (int32_t*)workspace;
}
}  // namespace b200

using namespace b200;

static inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }
static inline bool stride_ok(int64_t elems) { return ((elems * 2) & 15) == 0; }

extern "C" {

int b200_version(void) { return B200ATTN_VERSION; }

const char* b200_status_string(int status) {
    switch (status) {
        case B200_OK: return "ok";
        case B200_ERR_NULL: return "b200attn: required pointer is NULL";
        case B200_ERR_SHAPE: return "b200attn: unsupported shape (need head_dim 128, page 256, Hq % Hkv == 0, Hq/Hkv <= 16)";
        case B200_ERR_ALIGN: return "b200attn: pointer or stride not 16-byte aligned";
        case B200_ERR_WORKSPACE: return "b200attn: split-K workspace too small";
        case B200_ERR_RANGE: return "b200attn: negative or out-of-range count";
        case B200_ERR_UNAVAILABLE: return "b200attn: requested kernel variant unavailable";
        case B200_ERR_SCRATCH_LIMIT:
            return "b200attn: batch x heads x splits needs more than the 8 MiB built-in scratch "
                   "(B200_DECODE_STATIC_SCRATCH_BYTES); lower num_splits or pass your own workspace to b200_attn_decode";
        default: break;
    }
    if (status > 0) return cudaGetErrorString(static_cast<cudaError_t>(status));
    return "b200attn: unknown status";
}

int b200_sm_count(void) { return sm_count(); }

int b200_kv_scatter(const void* key, const void* value, int64_t key_row_stride, int64_t value_row_stride,
                    void* k_cache, void* v_cache, const int32_t* slot_mapping,
                    int32_t num_tokens, int32_t num_kv_heads, int32_t head_dim, void* stream) {
    if (num_tokens < 0 || num_kv_heads <= 0) return B200_ERR_RANGE;
    if (num_tokens == 0) return B200_OK;
    if (!key || !value || !k_cache || !v_cache || !slot_mapping) return B200_ERR_NULL;
    if (head_dim != B200_HD_CONST) return B200_ERR_SHAPE;
    if (!aligned16(key) || !aligned16(value) || !aligned16(k_cache) || !aligned16(v_cache) ||
        !stride_ok(key_row_stride) || !stride_ok(value_row_stride))
        return B200_ERR_ALIGN;
    return launch_kv_scatter(key, value, key_row_stride, value_row_stride, k_cache, v_cache, slot_mapping,
                             num_tokens, num_kv_heads, head_dim, (cudaStream_t)stream);
}

// Host-only heuristic (static under CUDA graphs, like flash-attn's num_splits which is fixed at
// capture time): the split count that minimises a small cost model of the launch, never finer
// than one page per split.
int b200_decode_plan(int32_t batch, int32_t num_kv_heads, int32_t max_context_len) {
    if (batch <= 0 || num_kv_heads <= 0) return 1;
    const char* env = getenv("B200ATTN_DECODE_SPLITS");
    if (env && atoi(env) > 0) return atoi(env);
    // Cost model fitted to the (CTA shape x split count) sweep of round 2 (profiles/r2_sweep_decode_tp.txt), in bytes
    // moved at HBM speed.  The kernel is HBM-bound, so a partly filled last wave costs nothing while it still holds
    // enough CTAs to saturate HBM: cfg3 at the TP8 heads (512 items on 296 slots) runs 9 % faster unsplit than with the
    // 4 splits that "fill the last wave".  What does cost: (a) a CTA on its own streams at ~1/100 of HBM bandwidth
    // (64 GB/s measured), so a sparse wave lasts at least 100 x the bytes of one CTA; (b) ~3 us per wave of fixed work
    // (prologue latency, cross-warp merge, partial write + ticket), i.e. ~20 MB at HBM speed; (c) below.
    const double slots = (double)sm_count() * 2;                 // 2 resident CTAs per SM
    const double items = (double)batch * num_kv_heads;
    const double row_bytes = (double)(max_context_len > 0 ? max_context_len : 1) * 512.0;   // K + V of one kv head
    int cap = (max_context_len + 255) / 256;                     // never finer than one page per split
    if (cap < 1) cap = 1;
    if (cap > 64) cap = 64;
    int best = 1;
    double best_cost = 0.0;
    for (int s = 1; s <= cap; ++s) {
        const double ctas = items * s, per = row_bytes / s;
        const double full = floor(ctas / slots), tail = ctas - full * slots;
        double cost = full * slots * per + (full + (tail > 0 ? 1 : 0)) * 20.0e6;
        if (tail > 0) cost += per * (tail > 100.0 ? tail : 100.0);
        if (s > 1) cost += ctas * 4096.0 + 5.0e6;                // partial O/LSE round trip + the last arriver's merge
        if (ctas < slots) cost *= 1.0 + 0.04 * (1.0 - ctas / slots);   // (c) a machine that is not full streams a few % slower
        if (s == 1 || cost < best_cost * 0.995) { best_cost = cost; best = s; }
    }
    return best;
}

size_t b200_decode_workspace_bytes(int32_t batch, int32_t num_heads, int32_t num_splits) {
    if (num_splits <= 1 || batch <= 0 || num_heads <= 0) return 0;
    return DEC_TICKET_BYTES + (size_t)batch * num_splits * num_heads * (B200_HD_CONST + 1) * sizeof(float);
}

int b200_attn_decode(const void* q, int64_t q_token_stride, const void* k_cache, const void* v_cache,
                     void* out, int64_t out_token_stride, const int32_t* context_lens,
                     const int32_t* block_tables, int32_t bt_row_stride,
                     int32_t batch, int32_t num_heads, int32_t num_kv_heads, int32_t head_dim, int32_t page_size,
                     float softmax_scale, int32_t num_splits, void* workspace, size_t workspace_bytes, void* stream) {
    if (batch < 0 || num_splits < 1 || num_splits > 256 || bt_row_stride < 0) return B200_ERR_RANGE;
    if (batch == 0) return B200_OK;
    if (!q || !k_cache || !v_cache || !out || !context_lens || !block_tables) return B200_ERR_NULL;
    if (head_dim != B200_HD_CONST || page_size != 256 || num_kv_heads <= 0 || num_heads % num_kv_heads != 0 ||
        num_heads / num_kv_heads > 16)
        return B200_ERR_SHAPE;
    if (!aligned16(q) || !aligned16(k_cache) || !aligned16(v_cache) || !aligned16(out) ||
        !stride_ok(q_token_stride) || !stride_ok(out_token_stride))
        return B200_ERR_ALIGN;
    const size_t need = b200_decode_workspace_bytes(batch, num_heads, num_splits);
    if (num_splits > 1 && (!workspace || workspace_bytes < need)) return B200_ERR_WORKSPACE;
    DecodeArgs a;
    a.q = q; a.q_token_stride = q_token_stride; a.k_cache = k_cache; a.v_cache = v_cache;
    a.out = out; a.out_token_stride = out_token_stride; a.context_lens = context_lens;
    a.block_tables = block_tables; a.bt_row_stride = bt_row_stride;
    a.batch = batch; a.Hq = num_heads; a.Hkv = num_kv_heads; a.scale = softmax_scale; a.num_splits = num_splits;
    carve_decode_workspace(a, workspace);
    return launch_attn_decode(a, (cudaStream_t)stream);
}

int b200_attn_prefill(const void* q, int64_t q_token_stride, const void* k, const void* v,
                      int64_t k_token_stride, int64_t v_token_stride, void* out, int64_t out_token_stride,
                      const int32_t* cu_seqlens_q, const int32_t* cu_seqlens_k,
                      const int32_t* block_tables, int32_t bt_row_stride,
                      int32_t batch, int32_t total_q, int32_t max_seqlen_q, int32_t max_seqlen_k,
                      int32_t num_heads, int32_t num_kv_heads, int32_t head_dim, int32_t page_size,
                      int64_t num_cache_blocks, float softmax_scale, int32_t variant, void* stream) {
    if (batch < 0 || total_q < 0 || max_seqlen_q < 0 || max_seqlen_k < 0) return B200_ERR_RANGE;
    if (batch == 0 || total_q == 0) return B200_OK;
    if (!q || !k || !v || !out || !cu_seqlens_q || !cu_seqlens_k) return B200_ERR_NULL;
    if (head_dim != B200_HD_CONST || (block_tables && page_size != 256) || num_kv_heads <= 0 ||
        num_heads % num_kv_heads != 0)
        return B200_ERR_SHAPE;
    if (!aligned16(q) || !aligned16(k) || !aligned16(v) || !aligned16(out) || !stride_ok(q_token_stride) ||
        !stride_ok(out_token_stride) || (!block_tables && (!stride_ok(k_token_stride) || !stride_ok(v_token_stride))))
        return B200_ERR_ALIGN;
    PrefillArgs a;
    a.q = q; a.q_token_stride = q_token_stride; a.k = k; a.v = v;
    a.k_token_stride = k_token_stride; a.v_token_stride = v_token_stride;
    a.out = out; a.out_token_stride = out_token_stride; a.cu_q = cu_seqlens_q; a.cu_k = cu_seqlens_k;
    a.block_tables = block_tables; a.bt_row_stride = bt_row_stride;
    a.batch = batch; a.total_q = total_q; a.max_q = max_seqlen_q; a.max_k = max_seqlen_k;
    a.Hq = num_heads; a.Hkv = num_kv_heads; a.num_cache_blocks = num_cache_blocks; a.scale = softmax_scale;
    if (variant == 0) {
        const char* env = getenv("B200ATTN_PREFILL");
        if (env && !strcmp(env, "mma")) variant = 1;
        else if (env && !strcmp(env, "tc")) variant = 2;
        else variant = prefill_tc_available() ? 2 : 1;
    }
    if (variant == 2) {
        if (!prefill_tc_available()) return B200_ERR_UNAVAILABLE;
        return launch_attn_prefill_tc(a, (cudaStream_t)stream);
    }
    if (variant == 3) {
        if (!prefill_tc_available()) return B200_ERR_UNAVAILABLE;
        return launch_attn_prefill_tc1(a, (cudaStream_t)stream);
    }
    if (variant != 1) return B200_ERR_RANGE;
    return launch_attn_prefill_mma(a, (cudaStream_t)stream);
}

size_t b200_attn_plan_bytes(int32_t batch, int32_t total_q) {
    if (batch <= 0 || total_q < 0) return 0;
    return (plan_bytes(batch, total_q) + 15) & ~(size_t)15;
}

int b200_attn_plan(const int32_t* cu_seqlens_q, const int32_t* cu_seqlens_k, int32_t batch, int32_t total_q,
                   int32_t num_heads, int32_t num_kv_heads, int32_t route_decode,
                   void* plan, size_t plan_bytes_given, void* stream) {
    if (batch < 0 || total_q < 0) return B200_ERR_RANGE;
    if (num_kv_heads <= 0 || num_heads <= 0 || num_heads % num_kv_heads != 0) return B200_ERR_SHAPE;
    if (!cu_seqlens_q || !cu_seqlens_k || !plan) return B200_ERR_NULL;
    if (!aligned16(plan)) return B200_ERR_ALIGN;
    if (plan_bytes_given < plan_bytes(batch > 0 ? batch : 1, total_q)) return B200_ERR_WORKSPACE;
    if (plan_units_per_item(num_heads, num_kv_heads) > PLAN_MAX_UNITS_PER_ITEM) return B200_ERR_SHAPE;
    return launch_attn_plan(cu_seqlens_q, cu_seqlens_k, batch, route_decode ? 1 : 0,
                            plan_tile_rows(num_heads, num_kv_heads), plan_units_per_item(num_heads, num_kv_heads),
                            prefill_pack_ok(num_heads, num_kv_heads) ? 1 : 0, (int32_t*)plan,
                            plan_max_items(batch, total_q), (cudaStream_t)stream);
}

int b200_attn_prefill_planned(const void* q, int64_t q_token_stride, const void* k, const void* v,
                              int64_t k_token_stride, int64_t v_token_stride, void* out, int64_t out_token_stride,
                              const int32_t* cu_seqlens_q, const int32_t* cu_seqlens_k,
                              const int32_t* block_tables, int32_t bt_row_stride,
                              int32_t batch, int32_t total_q, int32_t max_seqlen_q, int32_t max_seqlen_k,
                              int32_t num_heads, int32_t num_kv_heads, int32_t head_dim, int32_t page_size,
                              int64_t num_cache_blocks, float softmax_scale,
                              const void* plan, size_t plan_bytes_given, void* stream) {
    if (batch < 0 || total_q < 0 || max_seqlen_q < 0 || max_seqlen_k < 0) return B200_ERR_RANGE;
    if (batch == 0 || total_q == 0) return B200_OK;
    if (!q || !k || !v || !out || !cu_seqlens_q || !cu_seqlens_k || !plan) return B200_ERR_NULL;
    if (head_dim != B200_HD_CONST || (block_tables && page_size != 256) || num_kv_heads <= 0 ||
        num_heads % num_kv_heads != 0)
        return B200_ERR_SHAPE;
    if (!aligned16(q) || !aligned16(k) || !aligned16(v) || !aligned16(out) || !stride_ok(q_token_stride) ||
        !stride_ok(out_token_stride) || (!block_tables && (!stride_ok(k_token_stride) || !stride_ok(v_token_stride))))
        return B200_ERR_ALIGN;
    if (plan_bytes_given < plan_bytes(batch, total_q)) return B200_ERR_WORKSPACE;
    if (!prefill_tc_available()) return B200_ERR_UNAVAILABLE;
    PrefillArgs a;
    a.q = q; a.q_token_stride = q_token_stride; a.k = k; a.v = v;
    a.k_token_stride = k_token_stride; a.v_token_stride = v_token_stride;
    a.out = out; a.out_token_stride = out_token_stride; a.cu_q = cu_seqlens_q; a.cu_k = cu_seqlens_k;
    a.block_tables = block_tables; a.bt_row_stride = bt_row_stride;
    a.batch = batch; a.total_q = total_q; a.max_q = max_seqlen_q; a.max_k = max_seqlen_k;
    a.Hq = num_heads; a.Hkv = num_kv_heads; a.num_cache_blocks = num_cache_blocks; a.scale = softmax_scale;
    return launch_attn_prefill_persist(a, (const int32_t*)plan, plan_max_items(batch, total_q), (cudaStream_t)stream);
}

int b200_attn_decode_varlen(const void* q, int64_t q_token_stride, const void* k_cache, const void* v_cache,
                            void* out, int64_t out_token_stride, const int32_t* cu_seqlens_q, const int32_t* cu_seqlens_k,
                            const int32_t* block_tables, int32_t bt_row_stride,
                            int32_t batch, int32_t num_heads, int32_t num_kv_heads, int32_t head_dim, int32_t page_size,
                            float softmax_scale, int32_t num_splits, void* workspace, size_t workspace_bytes,
                            const void* plan, size_t plan_bytes_given, void* stream) {
    if (batch < 0 || num_splits < 1 || num_splits > 256 || bt_row_stride < 0) return B200_ERR_RANGE;
    if (batch == 0) return B200_OK;
    if (!q || !k_cache || !v_cache || !out || !cu_seqlens_q || !cu_seqlens_k || !block_tables || !plan) return B200_ERR_NULL;
    if (head_dim != B200_HD_CONST || page_size != 256 || num_kv_heads <= 0 || num_heads % num_kv_heads != 0 ||
        num_heads / num_kv_heads > 16)
        return B200_ERR_SHAPE;
    if (!aligned16(q) || !aligned16(k_cache) || !aligned16(v_cache) || !aligned16(out) ||
        !stride_ok(q_token_stride) || !stride_ok(out_token_stride))
        return B200_ERR_ALIGN;
    const size_t need = b200_decode_workspace_bytes(batch, num_heads, num_splits);
    if (num_splits > 1 && (!workspace || workspace_bytes < need)) return B200_ERR_WORKSPACE;
    if (plan_bytes_given < (size_t)PLAN_HEADER_WORDS * 4 + (size_t)batch * 4) return B200_ERR_WORKSPACE;
    DecodeArgs a;
    a.q = q; a.q_token_stride = q_token_stride; a.k_cache = k_cache; a.v_cache = v_cache;
    a.out = out; a.out_token_stride = out_token_stride; a.context_lens = nullptr;
    a.block_tables = block_tables; a.bt_row_stride = bt_row_stride;
    a.batch = batch; a.Hq = num_heads; a.Hkv = num_kv_heads; a.scale = softmax_scale; a.num_splits = num_splits;
    carve_decode_workspace(a, workspace);
    const int32_t* pl = (const int32_t*)plan;
    a.cu_q = cu_seqlens_q; a.cu_k = cu_seqlens_k;
    a.n_rows = pl + 1;
    a.row_list = pl + PLAN_HEADER_WORDS;
    return launch_attn_decode(a, (cudaStream_t)stream);
}

int b200_qk_norm_rope(const void* q, int64_t q_token_stride, const void* k, int64_t k_token_stride,
                      const int64_t* positions, const float* cos_sin_cache, const void* q_weight,
                      const void* k_weight, float eps, void* q_out, void* k_out,
                      int32_t num_tokens, int32_t num_heads, int32_t num_kv_heads, int32_t head_dim, void* stream) {
    if (num_tokens < 0 || num_heads <= 0 || num_kv_heads <= 0) return B200_ERR_RANGE;
    if (num_tokens == 0) return B200_OK;
    if (!q || !k || !positions || !cos_sin_cache || !q_out || !k_out) return B200_ERR_NULL;   /* NULL weights => rotation only */
    if ((q_weight == nullptr) != (k_weight == nullptr)) return B200_ERR_NULL;                 /* both or neither */
    if (head_dim != B200_HD_CONST) return B200_ERR_SHAPE;
    if (!aligned16(q) || !aligned16(k) || !aligned16(q_out) || !aligned16(k_out) || !stride_ok(q_token_stride) ||
        !stride_ok(k_token_stride) || (reinterpret_cast<uintptr_t>(cos_sin_cache) & 15))
        return B200_ERR_ALIGN;                                      /* the kernel uses 16-byte accesses throughout */
    return launch_qk_norm_rope(q, q_token_stride, k, k_token_stride, positions, cos_sin_cache, q_weight, k_weight,
                               eps, q_out, k_out, num_tokens, num_heads, num_kv_heads, (cudaStream_t)stream);
}

int b200_qk_norm_rope_store(const void* q, int64_t q_token_stride, const void* k, int64_t k_token_stride,
                            const void* v, int64_t v_token_stride,
                            const int64_t* positions, const float* cos_sin_cache, const void* q_weight,
                            const void* k_weight, float eps, void* q_out, void* k_out,
                            void* k_cache, void* v_cache, const int32_t* slot_mapping,
                            int32_t num_tokens, int32_t num_heads, int32_t num_kv_heads, int32_t head_dim, void* stream) {
    if (num_tokens < 0 || num_heads <= 0 || num_kv_heads <= 0) return B200_ERR_RANGE;
    if (num_tokens == 0) return B200_OK;
    if (!q || !k || !v || !positions || !cos_sin_cache || !q_out || !k_cache || !v_cache || !slot_mapping)
        return B200_ERR_NULL;                                       /* k_out and the weights may be NULL */
    if ((q_weight == nullptr) != (k_weight == nullptr)) return B200_ERR_NULL;   /* both weights or neither */
    if (head_dim != B200_HD_CONST) return B200_ERR_SHAPE;
    if (!aligned16(q) || !aligned16(k) || !aligned16(v) || !aligned16(q_out) || (k_out && !aligned16(k_out)) ||
        !aligned16(k_cache) || !aligned16(v_cache) || !stride_ok(q_token_stride) || !stride_ok(k_token_stride) ||
        !stride_ok(v_token_stride) || (reinterpret_cast<uintptr_t>(cos_sin_cache) & 15))
        return B200_ERR_ALIGN;
    return launch_qk_norm_rope_store(q, q_token_stride, k, k_token_stride, v, v_token_stride, positions, cos_sin_cache,
                                     q_weight, k_weight, eps, q_out, k_out, k_cache, v_cache, slot_mapping,
                                     num_tokens, num_heads, num_kv_heads, (cudaStream_t)stream);
}

int b200_rmsnorm(const void* x, void* residual, const void* weight, float eps, void* out,
                 int32_t num_tokens, int32_t hidden, void* stream) {
    if (num_tokens < 0 || hidden <= 0) return B200_ERR_RANGE;
    if (num_tokens == 0) return B200_OK;
    if (!x || !weight || !out) return B200_ERR_NULL;
    if (hidden % 8 != 0 || hidden > 16384) return B200_ERR_SHAPE;
    if (!aligned16(x) || !aligned16(weight) || !aligned16(out) || (residual && !aligned16(residual))) return B200_ERR_ALIGN;
    return launch_rmsnorm(x, residual, weight, eps, out, num_tokens, hidden, (cudaStream_t)stream);
}

int b200_silu_and_mul(const void* x, void* out, int32_t num_tokens, int32_t inter, void* stream) {
    if (num_tokens < 0 || inter <= 0) return B200_ERR_RANGE;
    if (num_tokens == 0) return B200_OK;
    if (!x || !out) return B200_ERR_NULL;
    if (inter % 8 != 0) return B200_ERR_SHAPE;
    if (!aligned16(x) || !aligned16(out)) return B200_ERR_ALIGN;
    return launch_silu_and_mul(x, out, num_tokens, inter, (cudaStream_t)stream);
}

}  // extern "C"
