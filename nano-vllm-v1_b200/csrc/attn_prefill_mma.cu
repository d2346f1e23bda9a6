// (2, variant 1) Varlen causal prefill on mma.sync -- the portable correctness baseline of the
// prefill piece and the fallback for shapes the tcgen05 kernel does not take.  The tensor-core
// product path for prefill is attn_prefill_tc.cu (tcgen05 + TMEM + TMA).
//
// CTA = 64 query rows of one (sequence, q head); 4 warps x 16 rows; K/V tiles of 64 tokens are
// double-buffered with cp.async; paged K/V tiles never straddle a 256-token page.
// Replaces flash_attn_varlen_func as called at attention.py:242-248 (paged and unpaged).
#include "common.cuh"
#include "launch.h"

namespace b200 {

constexpr int PF_BM = 64, PF_BN = 64, PF_THREADS = 128;
constexpr int PF_TILE_BYTES = 64 * B200_HD * 2;                 // 16 KB
constexpr int PF_SMEM_BYTES = PF_TILE_BYTES * 5;                // Q + 2 x (K, V) = 80 KB

struct PrefillParams {
    const __nv_bfloat16* q; int64_t q_token_stride;
    const uint8_t* k; const uint8_t* v; int64_t k_token_bytes, v_token_bytes;
    __nv_bfloat16* out; int64_t out_token_stride;
    const int32_t* cu_q; const int32_t* cu_k; const int32_t* block_tables; int32_t bt_row_stride;
    int32_t Hq, Hkv, G; float scale_log2;
};

__global__ void __launch_bounds__(PF_THREADS, 2) prefill_mma_kernel(const PrefillParams p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    const int b = blockIdx.z, head = blockIdx.y, kvh = head / p.G;
    const int q0 = p.cu_q[b], Lq = p.cu_q[b + 1] - q0;
    const int k0 = p.cu_k[b], Lk = p.cu_k[b + 1] - k0;
    const int m0 = blockIdx.x * PF_BM;
    if (m0 >= Lq) return;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int offset = Lk - Lq;                                   // bottom-right alignment
    const int kv_end = min(Lk, m0 + PF_BM + offset);              // exclusive bound of visible keys for this tile
    const int n_tiles = kv_end > 0 ? (kv_end + PF_BN - 1) / PF_BN : 0;

    const uint32_t sQ = smem_u32(smem);
    const uint32_t sK = sQ + PF_TILE_BYTES;                       // 2 stages
    const uint32_t sV = sK + 2 * PF_TILE_BYTES;                   // 2 stages
    const bool paged = p.block_tables != nullptr;
    const int32_t* bt = paged ? p.block_tables + (int64_t)b * p.bt_row_stride : nullptr;

    auto load_kv = [&](int j, int stage) {
        const int n0 = j * PF_BN;
        const uint8_t *kp, *vp;
        int64_t kstep, vstep;
        if (paged) {
            const int64_t blk = __ldg(bt + n0 / B200_PAGE);
            const int64_t row_bytes = (int64_t)p.Hkv * B200_HD * 2;
            const int64_t base = (blk * B200_PAGE + (n0 & (B200_PAGE - 1))) * row_bytes + (int64_t)kvh * (B200_HD * 2);
            kp = p.k + base; vp = p.v + base; kstep = vstep = row_bytes;
        } else {
            kp = p.k + (int64_t)(k0 + n0) * p.k_token_bytes + (int64_t)kvh * (B200_HD * 2);
            vp = p.v + (int64_t)(k0 + n0) * p.v_token_bytes + (int64_t)kvh * (B200_HD * 2);
            kstep = p.k_token_bytes; vstep = p.v_token_bytes;
        }
#pragma unroll
        for (int it = 0; it < 8; ++it) {
            const int idx = it * PF_THREADS + threadIdx.x;
            const int r = idx >> 4, c = idx & 15;
            const bool ok = n0 + r < Lk;
            const int rr = ok ? r : 0;                            // keep the address in bounds when zero-filling
            const uint32_t off = swz256(r, c);
            cp_async16(sK + stage * PF_TILE_BYTES + off, kp + rr * kstep + c * 16, ok ? 16u : 0u);
            cp_async16(sV + stage * PF_TILE_BYTES + off, vp + rr * vstep + c * 16, ok ? 16u : 0u);
        }
    };

    // Q tile (rows past Lq zero-filled)
    {
        const __nv_bfloat16* qb = p.q + (int64_t)q0 * p.q_token_stride + (int64_t)head * B200_HD;
#pragma unroll
        for (int it = 0; it < 8; ++it) {
            const int idx = it * PF_THREADS + threadIdx.x;
            const int r = idx >> 4, c = idx & 15;
            const bool ok = m0 + r < Lq;
            const int rr = ok ? m0 + r : m0;
            cp_async16(sQ + swz256(r, c), qb + (int64_t)rr * p.q_token_stride + c * 8, ok ? 16u : 0u);
        }
    }
    if (n_tiles > 0) load_kv(0, 0);
    cp_async_commit();

    const int r0 = lane >> 2, cq = (lane & 3) * 2;
    const int mat = lane >> 3, mr = lane & 7;
    const int row0 = m0 + warp * 16 + r0, row1 = row0 + 8;        // query index within the sequence
    uint32_t qa[8][4];
    float o[16][4];
#pragma unroll
    for (int i = 0; i < 16; ++i) { o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f; }
    float mrow0 = -INFINITY, mrow1 = -INFINITY, l0 = 0.f, l1 = 0.f;

    for (int j = 0; j < n_tiles; ++j) {
        if (j + 1 < n_tiles) { load_kv(j + 1, (j + 1) & 1); cp_async_commit(); cp_async_wait<1>(); }
        else cp_async_wait<0>();
        __syncthreads();
        if (j == 0) {
#pragma unroll
            for (int ks = 0; ks < 8; ++ks)
                ldmatrix_x4(qa[ks], sQ + swz256(warp * 16 + (mat & 1) * 8 + mr, ks * 2 + (mat >> 1)));
        }
        const uint32_t kst = sK + (j & 1) * PF_TILE_BYTES, vst = sV + (j & 1) * PF_TILE_BYTES;
        const int n0 = j * PF_BN;

        float sc[8][4];
#pragma unroll
        for (int i = 0; i < 8; ++i) { sc[i][0] = sc[i][1] = sc[i][2] = sc[i][3] = 0.f; }
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
#pragma unroll
            for (int np = 0; np < 4; ++np) {
                uint32_t kb[4];
                ldmatrix_x4(kb, kst + swz256(np * 16 + (mat >> 1) * 8 + mr, ks * 2 + (mat & 1)));
                mma_bf16_16816(sc[2 * np], qa[ks], kb[0], kb[1]);
                mma_bf16_16816(sc[2 * np + 1], qa[ks], kb[2], kb[3]);
            }
        }
        const bool need_mask = (n0 + PF_BN > Lk) || (n0 + PF_BN - 1 > m0 + warp * 16 + offset);
#pragma unroll
        for (int nt = 0; nt < 8; ++nt)
#pragma unroll
            for (int jx = 0; jx < 4; ++jx) {
                float v = sc[nt][jx] * p.scale_log2;
                if (need_mask) {
                    const int col = n0 + nt * 8 + cq + (jx & 1);
                    const int row = (jx & 2) ? row1 : row0;
                    if (col >= Lk || col > row + offset) v = -INFINITY;
                }
                sc[nt][jx] = v;
            }
        float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) {
            mx0 = fmaxf(mx0, fmaxf(sc[nt][0], sc[nt][1]));
            mx1 = fmaxf(mx1, fmaxf(sc[nt][2], sc[nt][3]));
        }
        mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
        mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
        const float mn0 = fmaxf(mrow0, mx0), mn1 = fmaxf(mrow1, mx1);
        const float mu0 = mn0 == -INFINITY ? 0.f : mn0, mu1 = mn1 == -INFINITY ? 0.f : mn1;   // all-masked rows
        const float a0 = fast_exp2(mrow0 - mu0), a1 = fast_exp2(mrow1 - mu1);
        mrow0 = mn0; mrow1 = mn1;
        float ps0 = 0.f, ps1 = 0.f;
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) {
            sc[nt][0] = fast_exp2(sc[nt][0] - mu0); sc[nt][1] = fast_exp2(sc[nt][1] - mu0);
            sc[nt][2] = fast_exp2(sc[nt][2] - mu1); sc[nt][3] = fast_exp2(sc[nt][3] - mu1);
            ps0 += sc[nt][0] + sc[nt][1];
            ps1 += sc[nt][2] + sc[nt][3];
        }
        l0 = l0 * a0 + ps0;
        l1 = l1 * a1 + ps1;
#pragma unroll
        for (int nt = 0; nt < 16; ++nt) { o[nt][0] *= a0; o[nt][1] *= a0; o[nt][2] *= a1; o[nt][3] *= a1; }
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
            uint32_t pa[4];
            pa[0] = pack_bf16x2(sc[2 * kk][0], sc[2 * kk][1]);
            pa[1] = pack_bf16x2(sc[2 * kk][2], sc[2 * kk][3]);
            pa[2] = pack_bf16x2(sc[2 * kk + 1][0], sc[2 * kk + 1][1]);
            pa[3] = pack_bf16x2(sc[2 * kk + 1][2], sc[2 * kk + 1][3]);
#pragma unroll
            for (int np = 0; np < 8; ++np) {
                uint32_t vb[4];
                ldmatrix_x4_trans(vb, vst + swz256(kk * 16 + (mat & 1) * 8 + mr, np * 2 + (mat >> 1)));
                mma_bf16_16816(o[2 * np], pa, vb[0], vb[1]);
                mma_bf16_16816(o[2 * np + 1], pa, vb[2], vb[3]);
            }
        }
        __syncthreads();                                          // stage (j&1) is overwritten by tile j+2
    }
    if (n_tiles == 0) cp_async_wait<0>();

    l0 += __shfl_xor_sync(0xffffffffu, l0, 1); l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 1); l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
    const float i0 = l0 > 0.f ? 1.f / l0 : 0.f, i1 = l1 > 0.f ? 1.f / l1 : 0.f;
    __nv_bfloat16* ob = p.out + (int64_t)head * B200_HD;
#pragma unroll
    for (int nt = 0; nt < 16; ++nt) {
        if (row0 < Lq)
            *reinterpret_cast<uint32_t*>(ob + (int64_t)(q0 + row0) * p.out_token_stride + nt * 8 + cq) = pack_bf16x2(o[nt][0] * i0, o[nt][1] * i0);
        if (row1 < Lq)
            *reinterpret_cast<uint32_t*>(ob + (int64_t)(q0 + row1) * p.out_token_stride + nt * 8 + cq) = pack_bf16x2(o[nt][2] * i1, o[nt][3] * i1);
    }
}

int launch_attn_prefill_mma(const PrefillArgs& a, cudaStream_t stream) {
    static PerDeviceOnce attr_once;                              // opt-in is per device
    if (attr_once.need()) {
        cudaError_t e = cudaFuncSetAttribute(prefill_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, PF_SMEM_BYTES);
        if (e != cudaSuccess) return (int)e;
        attr_once.done();
    }
    if (a.batch == 0 || a.total_q == 0 || a.max_q == 0) return 0;
    PrefillParams p;
    p.q = (const __nv_bfloat16*)a.q; p.q_token_stride = a.q_token_stride;
    p.k = (const uint8_t*)a.k; p.v = (const uint8_t*)a.v;
    p.k_token_bytes = a.k_token_stride * 2; p.v_token_bytes = a.v_token_stride * 2;
    p.out = (__nv_bfloat16*)a.out; p.out_token_stride = a.out_token_stride;
    p.cu_q = a.cu_q; p.cu_k = a.cu_k; p.block_tables = a.block_tables; p.bt_row_stride = a.bt_row_stride;
    p.Hq = a.Hq; p.Hkv = a.Hkv; p.G = a.Hq / a.Hkv;
    p.scale_log2 = a.scale * 1.4426950408889634f;
    dim3 grid((a.max_q + PF_BM - 1) / PF_BM, a.Hq, a.batch);
    prefill_mma_kernel<<<grid, PF_THREADS, PF_SMEM_BYTES, stream>>>(p);
    return (int)cudaGetLastError();
}

}  // namespace b200
