// Internal host-side launcher declarations (C++), one per kernel family.  The extern "C"
// surface in abi.cu validates arguments and forwards here.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

#define B200_HD_CONST 128

namespace b200 {

int sm_count();

int launch_kv_scatter(const void* key, const void* value, int64_t key_row_stride, int64_t value_row_stride,
                      void* k_cache, void* v_cache, const int32_t* slot_mapping,
                      int32_t T, int32_t Hkv, int32_t D, cudaStream_t stream);
int launch_kv_scatter_generic(const void* key, const void* value, int64_t key_row_stride, int64_t value_row_stride,
                              void* k_cache, void* v_cache, const int32_t* slot_mapping,
                              int32_t T, int32_t Hkv, int32_t D, cudaStream_t stream);
int launch_qk_norm_rope(const void* q, int64_t q_token_stride, const void* k, int64_t k_token_stride,
                        const int64_t* positions, const float* cos_sin, const void* qw, const void* kw, float eps,
                        void* q_out, void* k_out, int32_t T, int32_t Hq, int32_t Hkv, cudaStream_t stream);
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
};
int launch_attn_decode(const DecodeArgs& a, cudaStream_t stream);

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

}  // namespace b200
