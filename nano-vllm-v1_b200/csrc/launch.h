// Internal host-side launcher declarations (C++), one per kernel family.  The extern "C"
// surface in abi.cu validates arguments and forwards here.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>
#include <stdlib.h>

#define B200_HD_CONST 128

namespace b200 {

int sm_count();                          // of the CURRENT device (cached per device)
int current_device();

// One-time per-device setup guard (cudaFuncSetAttribute opt-ins are per device; a process may touch
// several GPUs, e.g. tests on cuda:1).
struct PerDeviceOnce {
    uint64_t mask = 0;
    int dev = 0;
    bool need() { dev = current_device(); return dev < 0 || dev >= 64 || !(mask >> dev & 1ull); }
    void done() { if (dev >= 0 && dev < 64) mask |= 1ull << dev; }
};

int launch_kv_scatter(const void* key, const void* value, int64_t key_row_stride, int64_t value_row_stride,
                      void* k_cache, void* v_cache, const int32_t* slot_mapping,
                      int32_t T, int32_t Hkv, int32_t D, cudaStream_t stream);
int launch_kv_scatter_generic(const void* key, const void* value, int64_t key_row_stride, int64_t value_row_stride,
                              void* k_cache, void* v_cache, const int32_t* slot_mapping,
                              int32_t T, int32_t Hkv, int32_t D, cudaStream_t stream);
int launch_qk_norm_rope(const void* q, int64_t q_token_stride, const void* k, int64_t k_token_stride,
                        const int64_t* positions, const float* cos_sin, const void* qw, const void* kw, float eps,
                        void* q_out, void* k_out, int32_t T, int32_t Hq, int32_t Hkv, cudaStream_t stream);
int launch_qk_norm_rope_store(const void* q, int64_t q_token_stride, const void* k, int64_t k_token_stride,
                              const void* v, int64_t v_token_stride,
                              const int64_t* positions, const float* cos_sin, const void* qw, const void* kw, float eps,
                              void* q_out, void* k_out, void* k_cache, void* v_cache, const int32_t* slot_mapping,
                              int32_t T, int32_t Hq, int32_t Hkv, cudaStream_t stream);
int launch_rmsnorm(const void* x, void* residual, const void* weight, float eps, void* out,
                   int32_t T, int32_t hidden, cudaStream_t stream);
int launch_silu_and_mul(const void* x, void* out, int32_t T, int32_t inter, cudaStream_t stream);

struct DecodeArgs {
    const void* q; int64_t q_token_stride;
    const void* k_cache; const void* v_cache;
    void* out; int64_t out_token_stride;
    const int32_t* context_lens; const int32_t* block_tables; int32_t bt_row_stride;
    int32_t batch, Hq, Hkv; float scale; int32_t num_splits;
    float* ws_o; float* ws_lse;
    int32_t* tickets = nullptr;           // zeroed int32[batch * Hkv]: fused split merge (NULL => separate reduce launch)
    // varlen form: non-null => rows are row_list[0 .. *n_rows), lengths/tokens from cu_k / cu_q
    const int32_t* cu_q = nullptr; const int32_t* cu_k = nullptr;
    const int32_t* row_list = nullptr; const int32_t* n_rows = nullptr;
};
int launch_attn_decode(const DecodeArgs& a, cudaStream_t stream);
// Fixed ticket region at the head of the split-K workspace (one int32 per (row, kv head)); must be zero
// before the first launch, every launch leaves it zero.
constexpr size_t DEC_TICKET_BYTES = 64 * 1024;

struct PrefillArgs {
    const void* q; int64_t q_token_stride;
    const void* k; const void* v; int64_t k_token_stride, v_token_stride;
    void* out; int64_t out_token_stride;
    const int32_t* cu_q; const int32_t* cu_k; const int32_t* block_tables; int32_t bt_row_stride;
    int32_t batch, total_q, max_q, max_k, Hq, Hkv; int64_t num_cache_blocks; float scale;
};
int launch_attn_prefill_mma(const PrefillArgs& a, cudaStream_t stream);
int launch_attn_prefill_tc(const PrefillArgs& a, cudaStream_t stream);   // tcgen05/TMEM/TMA, 2 query tiles in flight
int launch_attn_prefill_tc1(const PrefillArgs& a, cudaStream_t stream);  // same, 1 query tile (bring-up / small shapes)
bool prefill_tc_available();

// ---- step plan (attn_plan.cu) + persistent tcgen05 prefill (attn_prefill_persist.cu) ----------
// plan buffer, int32 words: [0] n_units  [1] n_decode_rows  [2] tile_rows  [3] n_items  [4..15] reserved
//                           [16 .. 16 + batch)        rows with exactly one query token (region padded to 4 words)
//                           [.. + 4*max_items)        scratch of the plan kernel: {sequence, tile, first unit, units} per item
//                           [.. + 8*max_units)        one 32-byte record per WORK UNIT of the persistent prefill kernel,
//                                                     heaviest first: {sequence, tile, q0, Lq} {k0, Lk, sub, pack} -- all a
//                                                     role needs about its next unit in ONE load level.  An item (query tile
//                                                     of a sequence) has Hq/2 units (two q heads each), Hq/4 when its tile is
//                                                     packed (<= 64 valid rows, two heads per tile), Hq in row-pair mode.
constexpr int PLAN_HEADER_WORDS = 16;
constexpr int PLAN_ITEM_WORDS = 4;
constexpr int PLAN_UNIT_WORDS = 8;
constexpr int PLAN_MAX_UNITS_PER_ITEM = 64;      // Hq <= 128 in pair mode, <= 64 in row-pair mode
constexpr int PLAN_TILE_ROWS_MIN = 128;
// tile rows of a launch: 128 when two q heads of one kv head share a CTA (even GQA group), else 256
inline int32_t plan_tile_rows(int32_t Hq, int32_t Hkv) { return ((Hq / Hkv) % 2 == 0) ? 128 : 256; }
inline int32_t plan_units_per_item(int32_t Hq, int32_t Hkv) { return plan_tile_rows(Hq, Hkv) == 128 ? Hq / 2 : Hq; }
inline int32_t plan_max_items(int32_t batch, int32_t total_q) { return total_q / PLAN_TILE_ROWS_MIN + batch; }
constexpr size_t plan_items_offset(int32_t batch) { return (size_t)PLAN_HEADER_WORDS + (((size_t)batch + 3) & ~(size_t)3); }
constexpr size_t plan_units_offset(int32_t batch, int32_t max_items) { return plan_items_offset(batch) + (size_t)PLAN_ITEM_WORDS * max_items; }
inline size_t plan_bytes(int32_t batch, int32_t total_q) {
    const int32_t mi = plan_max_items(batch, total_q);
    return (plan_units_offset(batch, mi) + (size_t)PLAN_UNIT_WORDS * PLAN_MAX_UNITS_PER_ITEM * mi) * sizeof(int32_t);
}
bool prefill_pack_ok(int32_t Hq, int32_t Hkv);     // attn_prefill_persist.cu: may short query tiles pack two heads per tile?
int launch_attn_plan(const int32_t* cu_q, const int32_t* cu_k, int32_t batch, int32_t route_decode, int32_t tile_rows,
                     int32_t units_per_item, int32_t pack_ok, int32_t* plan, int32_t max_items, cudaStream_t stream);
int launch_attn_prefill_persist(const PrefillArgs& a, const int32_t* plan, int32_t max_items, cudaStream_t stream);


// ---- peer-memory collectives (allreduce.cu, allreduce_nvls.cu) ----------------------------------
// Launch with the cooperative attribute: the driver either places ALL blocks at once or fails the launch, so
// the blocks that spin on each other are co-resident whatever else is running on the device (another stream's
// kernel holding SMs can delay the launch, never deadlock it).  Capturable in CUDA graphs.
template <typename... KArgs, typename... Args>
inline int launch_coop(void (*kernel)(KArgs...), int grid, int threads, cudaStream_t stream, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid); cfg.blockDim = dim3((unsigned)threads); cfg.dynamicSmemBytes = 0; cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeCooperative;
    attr[0].val.cooperative = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    return (int)cudaLaunchKernelEx(&cfg, kernel, args...);
}

uint64_t ar_timeout_ns();   // B200ATTN_AR_TIMEOUT_S (default 30 s) in ns

}  // namespace b200
