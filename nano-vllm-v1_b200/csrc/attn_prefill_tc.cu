// (2, variant 2) Varlen causal prefill / chunked prefill on Blackwell tensor cores.
//
//   TMA  (cp.async.bulk.tensor.3d, 128B swizzle) stages Q once and K/V tiles of 128 tokens
//        straight out of the paged cache [num_blocks*256, Hkv, 128] (or the packed k/v of an
//        unpaged prefill) into shared memory; a 128-token tile never straddles a 256-token page.
//   tcgen05.mma (kind::f16, M=128, N=128, cta_group::1), issued by ONE thread:
//        S = Q K^T   : A, B from shared memory (K-major, SWIZZLE_128B descriptors)
//        O += P V    : A = P from TENSOR MEMORY (bf16, written by the softmax warps over S),
//                      B = V from shared memory (MN-major descriptor -- V is never transposed)
//   TMEM (512 columns) holds, per 128-row query tile, S/P (128 cols) and O (128 cols); two query
//        tiles are in flight per CTA so the tensor pipe works on one tile while the softmax
//        warpgroup of the other is busy (ping-pong).
//   Softmax: one thread per query row (tcgen05.ld 32x32b), no shuffles; base-2 with the scale
//        folded in; P rounded to bf16 before PV like flash-attn; O is rescaled in TMEM lazily
//        (only when a row maximum grows by more than 2^8), the final division happens once.
//
// Warp roles (NQ = 2: 384 threads; NQ = 1: 256 threads)
//   warp 0   TMA producer (one elected lane)          warp 2  TMEM allocator
//   warp 1   MMA issuer   (one elected lane)          warp 3  idle
//   warps 4..7 / 8..11   softmax + correction + epilogue warpgroup of query tile 0 / 1
//
// Replaces flash_attn_varlen_func (attention.py:242-248), paged and unpaged.
#include "launch.h"
#include "tc_ptx.cuh"

namespace b200 {

struct TcParams {
    const int32_t* cu_q; const int32_t* cu_k; const int32_t* block_tables; int32_t bt_row_stride;
    __nv_bfloat16* out; int64_t out_token_stride;
    int32_t Hq, G; float scale_log2; int32_t num_m_tiles;
};

template <int NQ>
struct TcSmem {
    static constexpr int Q_OFF = 0;
    static constexpr int K_OFF = Q_OFF + NQ * TC_TILE_BYTES;
    static constexpr int V_OFF = K_OFF + TC_KV_STAGES * TC_TILE_BYTES;
    static constexpr int BAR_OFF = V_OFF + TC_KV_STAGES * TC_TILE_BYTES;
    // barriers: q_full[NQ], k_full[S], k_empty[S], v_full[S], v_empty[S], s_full[NQ], p_full[NQ], o_final[NQ]
    static constexpr int NUM_BARS = 4 * NQ + 4 * TC_KV_STAGES;
    static constexpr int TMEM_PTR_OFF = BAR_OFF + NUM_BARS * 8;
    static constexpr int TOTAL = TMEM_PTR_OFF + 16;
    static constexpr int ALLOC = TOTAL + 1024;           // slack for manual 1024-byte alignment
};

template <int NQ>
__global__ void __launch_bounds__(128 + NQ * 128, 1) prefill_tc_kernel(
    const __grid_constant__ CUtensorMap map_q, const __grid_constant__ CUtensorMap map_k,
    const __grid_constant__ CUtensorMap map_v, const TcParams p) {
    using L = TcSmem<NQ>;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int b = blockIdx.z, head = blockIdx.y, kvh = head / p.G;
    const int mt = p.num_m_tiles - 1 - (int)blockIdx.x;            // heavy (late) tiles first
    const int q0 = p.cu_q[b], Lq = p.cu_q[b + 1] - q0;
    const int k0 = p.cu_k[b], Lk = p.cu_k[b + 1] - k0;
    const int m0 = mt * (NQ * TC_BM);
    if (m0 >= Lq) return;
    const int offset = Lk - Lq;                                     // bottom-right causal alignment

    // number of KV tiles each query tile attends to
    int n_t[NQ];
    int n_max = 0;
#pragma unroll
    for (int t = 0; t < NQ; ++t) {
        const int r_begin = m0 + t * TC_BM;
        int kv_end = min(Lk, r_begin + TC_BM + offset);
        if (r_begin >= Lq || kv_end < 0) kv_end = 0;
        n_t[t] = (kv_end + TC_BN - 1) / TC_BN;
        n_max = max(n_max, n_t[t]);
    }

    const uint32_t sQ = smem_base + L::Q_OFF, sK = smem_base + L::K_OFF, sV = smem_base + L::V_OFF;
    const uint32_t bars = smem_base + L::BAR_OFF;
    auto q_full = [&](int t) { return bars + 8 * t; };
    auto k_full = [&](int s) { return bars + 8 * (NQ + s); };
    auto k_empty = [&](int s) { return bars + 8 * (NQ + TC_KV_STAGES + s); };
    auto v_full = [&](int s) { return bars + 8 * (NQ + 2 * TC_KV_STAGES + s); };
    auto v_empty = [&](int s) { return bars + 8 * (NQ + 3 * TC_KV_STAGES + s); };
    auto s_full = [&](int t) { return bars + 8 * (NQ + 4 * TC_KV_STAGES + t); };
    auto p_full = [&](int t) { return bars + 8 * (2 * NQ + 4 * TC_KV_STAGES + t); };
    auto o_final = [&](int t) { return bars + 8 * (3 * NQ + 4 * TC_KV_STAGES + t); };
    volatile uint32_t* tmem_ptr_smem = reinterpret_cast<volatile uint32_t*>(smem_gen + L::TMEM_PTR_OFF);
    constexpr uint32_t TMEM_COLS = NQ * 256;

    if (warp == 1 && lane == 0) {
#pragma unroll
        for (int t = 0; t < NQ; ++t) {
            mbar_init(q_full(t), 1); mbar_init(s_full(t), 1); mbar_init(p_full(t), 128); mbar_init(o_final(t), 1);
        }
#pragma unroll
        for (int s = 0; s < TC_KV_STAGES; ++s) {
            mbar_init(k_full(s), 1); mbar_init(k_empty(s), 1); mbar_init(v_full(s), 1); mbar_init(v_empty(s), 1);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    } else if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_base + L::TMEM_PTR_OFF), "r"(TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    } else if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_q) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_k) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_v) : "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr_smem;

    if (warp == 0) {
        // ===================================== TMA producer =====================================
        if (lane == 0) {
#pragma unroll
            for (int t = 0; t < NQ; ++t) {
                if (n_t[t] > 0) {
                    mbar_expect_tx(q_full(t), TC_TILE_BYTES);
                    tma_load_3d(sQ + t * TC_TILE_BYTES, &map_q, q_full(t), 0, head, q0 + m0 + t * TC_BM);
                    tma_load_3d(sQ + t * TC_TILE_BYTES + TC_HALF_BYTES, &map_q, q_full(t), 64, head, q0 + m0 + t * TC_BM);
                }
            }
            const bool paged = p.block_tables != nullptr;
            const int32_t* bt = paged ? p.block_tables + (int64_t)b * p.bt_row_stride : nullptr;
            int nxt_blk = (paged && n_max > 0) ? __ldg(bt) : 0;
            for (int j = 0; j < n_max; ++j) {
                const int n0 = j * TC_BN;
                int row;
                if (paged) {
                    const int blk = nxt_blk;
                    if (j + 1 < n_max) nxt_blk = __ldg(bt + (n0 + TC_BN) / B200_PAGE);
                    row = blk * B200_PAGE + (n0 & (B200_PAGE - 1));
                } else {
                    row = k0 + n0;
                }
                const int st = j % TC_KV_STAGES;
                const uint32_t par = (uint32_t)((j / TC_KV_STAGES) & 1);
                mbar_wait(k_empty(st), par ^ 1u);
                mbar_expect_tx(k_full(st), TC_TILE_BYTES);
                tma_load_3d(sK + st * TC_TILE_BYTES, &map_k, k_full(st), 0, kvh, row);
                tma_load_3d(sK + st * TC_TILE_BYTES + TC_HALF_BYTES, &map_k, k_full(st), 64, kvh, row);
                mbar_wait(v_empty(st), par ^ 1u);
                mbar_expect_tx(v_full(st), TC_TILE_BYTES);
                tma_load_3d(sV + st * TC_TILE_BYTES, &map_v, v_full(st), 0, kvh, row);
                tma_load_3d(sV + st * TC_TILE_BYTES + TC_HALF_BYTES, &map_v, v_full(st), 64, kvh, row);
            }
        }
    } else if (warp == 1) {
        // ===================================== MMA issuer =======================================
        if (lane == 0) {
            constexpr uint32_t IDESC_QK = umma_idesc(0);           // B = K tile, K-major
            constexpr uint32_t IDESC_PV = umma_idesc(1);           // B = V tile, MN-major
            auto issue_qk = [&](int t, int j) {
                const uint32_t kst = sK + (j % TC_KV_STAGES) * TC_TILE_BYTES;
                const uint32_t qst = sQ + t * TC_TILE_BYTES;
                const uint32_t d = tmem_base + t * 256;
#pragma unroll
                for (int ks = 0; ks < 8; ++ks) {                   // 8 x K=16 over head_dim
                    const uint32_t off = (ks >> 2) * TC_HALF_BYTES + (ks & 3) * 32;
                    tc_mma_ss(d, umma_desc(qst + off, 0, 1024), umma_desc(kst + off, 0, 1024), IDESC_QK, ks > 0);
                }
            };
            auto issue_pv = [&](int t, int j) {
                const uint32_t vst = sV + (j % TC_KV_STAGES) * TC_TILE_BYTES;
                const uint32_t d = tmem_base + t * 256 + 128;
                const uint32_t a = tmem_base + t * 256;            // P (bf16) aliases S
#pragma unroll
                for (int kk = 0; kk < 8; ++kk) {                   // 8 x K=16 over kv tokens
                    // MN-major B: 64-dim atoms TC_HALF_BYTES apart (LBO), 8-token groups 1024 B apart (SBO)
                    tc_mma_ts(d, a + kk * 8, umma_desc(vst + kk * 2048, TC_HALF_BYTES, 1024), IDESC_PV, (j > 0 || kk > 0));
                }
            };
#pragma unroll
            for (int t = 0; t < NQ; ++t)
                if (n_t[t] > 0) mbar_wait(q_full(t), 0);
            if (n_max > 0) {
                mbar_wait(k_full(0), 0);
                tc_fence_after();
#pragma unroll
                for (int t = 0; t < NQ; ++t)
                    if (n_t[t] > 0) { issue_qk(t, 0); tc_commit(s_full(t)); }
                tc_commit(k_empty(0));
            }
            for (int j = 0; j < n_max; ++j) {
                const int st = j % TC_KV_STAGES;
                const uint32_t par = (uint32_t)((j / TC_KV_STAGES) & 1);
                mbar_wait(v_full(st), par);
                if (j + 1 < n_max) mbar_wait(k_full((j + 1) % TC_KV_STAGES), (uint32_t)(((j + 1) / TC_KV_STAGES) & 1));
                tc_fence_after();
#pragma unroll
                for (int t = 0; t < NQ; ++t) {
                    if (j < n_t[t]) {
                        mbar_wait(p_full(t), (uint32_t)(j & 1));
                        tc_fence_after();
                        issue_pv(t, j);
                        if (j + 1 < n_t[t]) { issue_qk(t, j + 1); tc_commit(s_full(t)); }
                        else tc_commit(o_final(t));
                    }
                }
                tc_commit(v_empty(st));
                if (j + 1 < n_max) tc_commit(k_empty((j + 1) % TC_KV_STAGES));
            }
        }
    } else if (warp >= 4) {
        // ======================== softmax + correction + epilogue warpgroups =====================
        const int t = (warp - 4) >> 2;                              // query tile of this warpgroup
        const int wq = warp & 3;                                    // TMEM lane quarter this warp may touch
        const int row_in_tile = wq * 32 + lane;
        const int row = m0 + t * TC_BM + row_in_tile;               // query index within the sequence
        const uint32_t lane_addr = (uint32_t)(wq * 32) << 16;
        const uint32_t tS = tmem_base + t * 256 + lane_addr;
        const uint32_t tO = tS + 128;
        const int nt = n_t[t];
        const int r_begin = m0 + t * TC_BM;
        float m_used = -INFINITY, l_sum = 0.f;

        for (int j = 0; j < nt; ++j) {
            const int n0 = j * TC_BN;
            mbar_wait(s_full(t), (uint32_t)(j & 1));
            tc_fence_after();
            uint32_t s[4][32];
#pragma unroll
            for (int c = 0; c < 4; ++c) tmem_ld32(tS + c * 32, s[c]);
            tmem_wait_ld();
            const bool need_mask = (n0 + TC_BN > Lk) || (n0 + TC_BN - 1 > r_begin + offset);
            float mx = -INFINITY;
#pragma unroll
            for (int c = 0; c < 4; ++c)
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    float v = __uint_as_float(s[c][i]) * p.scale_log2;
                    if (need_mask) {
                        const int col = n0 + c * 32 + i;
                        if (col >= Lk || col > row + offset) v = -INFINITY;
                    }
                    s[c][i] = __float_as_uint(v);
                    mx = fmaxf(mx, v);
                }
            // lazy rescale: keep exponentiating against m_used until some row's max outgrows it by 2^8
            const bool grow = mx > m_used + TC_RESCALE_THRESHOLD || (m_used == -INFINITY && mx > -INFINITY);
            if (__any_sync(0xffffffffu, grow)) {
                const float m_new = fmaxf(m_used, mx);
                const float m_eff = m_new == -INFINITY ? 0.f : m_new;
                const float alpha = fast_exp2(m_used - m_eff);      // 0 when m_used = -inf
                l_sum *= alpha;
                m_used = m_new;
                if (j > 0) {                                        // PV(j-1) completed before s_full(j) fired
#pragma unroll
                    for (int c = 0; c < 4; ++c) {
                        uint32_t o[32];
                        tmem_ld32(tO + c * 32, o);
                        tmem_wait_ld();
#pragma unroll
                        for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
                        tmem_st32(tO + c * 32, o);
                    }
                }
            }
            const float m_eff = m_used == -INFINITY ? 0.f : m_used;
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                uint32_t pk[16];
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    const float p0 = fast_exp2(__uint_as_float(s[c][2 * i]) - m_eff);
                    const float p1 = fast_exp2(__uint_as_float(s[c][2 * i + 1]) - m_eff);
                    l_sum += p0 + p1;
                    pk[i] = pack_bf16x2(p0, p1);
                }
                tmem_st16(tS + c * 16, pk);                         // P over the first 64 columns of S
            }
            tmem_wait_st();
            if (n0 + TC_BN > Lk) {
                // Rows of the V tile past the context may hold NaN garbage (torch.empty cache):
                // 0 * NaN would poison O, so zero them in shared memory before PV reads them.
                const int st = j % TC_KV_STAGES;
                mbar_wait(v_full(st), (uint32_t)((j / TC_KV_STAGES) & 1));
                uint8_t* vt = smem_gen + L::V_OFF + st * TC_TILE_BYTES;
                const int first = Lk - n0;                          // first invalid row of this tile
                for (int idx = row_in_tile; idx < (TC_BN - first) * 16; idx += 128) {
                    const int r = first + (idx >> 4), c16 = idx & 15;
                    *reinterpret_cast<int4*>(vt + (c16 >> 3) * TC_HALF_BYTES + r * 128 + (c16 & 7) * 16) = make_int4(0, 0, 0, 0);
                }
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            }
            tc_fence_before();
            mbar_arrive(p_full(t));
        }

        // ---- epilogue: O / l -> bf16 -> global (rows beyond the sequence are dropped) ------------
        if (nt > 0) {
            mbar_wait(o_final(t), 0);
            tc_fence_after();
        }
        const float inv = l_sum > 0.f ? 1.f / l_sum : 0.f;
        __nv_bfloat16* orow = p.out + (int64_t)(q0 + row) * p.out_token_stride + (int64_t)head * B200_HD;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            uint32_t o[32];
            if (nt > 0) {
                tmem_ld32(tO + c * 32, o);
                tmem_wait_ld();
            } else {
#pragma unroll
                for (int i = 0; i < 32; ++i) o[i] = 0u;
            }
            if (row < Lq) {
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    int4 v;
                    v.x = (int)pack_bf16x2(__uint_as_float(o[8 * i + 0]) * inv, __uint_as_float(o[8 * i + 1]) * inv);
                    v.y = (int)pack_bf16x2(__uint_as_float(o[8 * i + 2]) * inv, __uint_as_float(o[8 * i + 3]) * inv);
                    v.z = (int)pack_bf16x2(__uint_as_float(o[8 * i + 4]) * inv, __uint_as_float(o[8 * i + 5]) * inv);
                    v.w = (int)pack_bf16x2(__uint_as_float(o[8 * i + 6]) * inv, __uint_as_float(o[8 * i + 7]) * inv);
                    *reinterpret_cast<int4*>(orow + c * 32 + i * 8) = v;
                }
            }
        }
    }

    // ---- teardown --------------------------------------------------------------------------------
    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
    }
}

bool prefill_tc_available() { return get_encode_fn() != nullptr; }

template <int NQ>
static int launch_tc(const PrefillArgs& a, cudaStream_t stream) {
    using L = TcSmem<NQ>;
    static PerDeviceOnce attr_once;                              // opt-in is per device
    if (attr_once.need()) {
        cudaError_t e = cudaFuncSetAttribute(prefill_tc_kernel<NQ>, cudaFuncAttributeMaxDynamicSharedMemorySize, L::ALLOC);
        if (e != cudaSuccess) return (int)e;
        attr_once.done();
    }
    CUtensorMap mq, mk, mv;
    const bool paged = a.block_tables != nullptr;
    if (!make_map(&mq, a.q, (uint64_t)a.total_q, (uint64_t)a.Hq, (uint64_t)a.q_token_stride)) return -6;
    if (paged) {
        const uint64_t rows = (uint64_t)a.num_cache_blocks * B200_PAGE;
        if (!make_map(&mk, a.k, rows, (uint64_t)a.Hkv, (uint64_t)a.Hkv * B200_HD)) return -6;
        if (!make_map(&mv, a.v, rows, (uint64_t)a.Hkv, (uint64_t)a.Hkv * B200_HD)) return -6;
    } else {
        // unpaged calls only happen when sum(Lk) == sum(Lq) (model_runner.py:483)
        if (!make_map(&mk, a.k, (uint64_t)a.total_q, (uint64_t)a.Hkv, (uint64_t)a.k_token_stride)) return -6;
        if (!make_map(&mv, a.v, (uint64_t)a.total_q, (uint64_t)a.Hkv, (uint64_t)a.v_token_stride)) return -6;
    }
    TcParams p;
    p.cu_q = a.cu_q; p.cu_k = a.cu_k; p.block_tables = a.block_tables; p.bt_row_stride = a.bt_row_stride;
    p.out = (__nv_bfloat16*)a.out; p.out_token_stride = a.out_token_stride;
    p.Hq = a.Hq; p.G = a.Hq / a.Hkv; p.scale_log2 = a.scale * 1.4426950408889634f;
    p.num_m_tiles = (a.max_q + NQ * TC_BM - 1) / (NQ * TC_BM);
    dim3 grid(p.num_m_tiles, a.Hq, a.batch);
    prefill_tc_kernel<NQ><<<grid, 128 + NQ * 128, L::ALLOC, stream>>>(mq, mk, mv, p);
    return (int)cudaGetLastError();
}

int launch_attn_prefill_tc(const PrefillArgs& a, cudaStream_t stream) {
    if (a.batch == 0 || a.total_q == 0 || a.max_q == 0) return 0;
    if (!prefill_tc_available()) return -6;
    return launch_tc<2>(a, stream);
}
int launch_attn_prefill_tc1(const PrefillArgs& a, cudaStream_t stream) {
    if (a.batch == 0 || a.total_q == 0 || a.max_q == 0) return 0;
    if (!prefill_tc_available()) return -6;
    return launch_tc<1>(a, stream);
}

}  // namespace b200
