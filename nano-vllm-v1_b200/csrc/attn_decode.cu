// (3) Split-K paged decode attention + LSE reduce.  HBM-bound: every K/V byte of every
// context token is read exactly once per KV head and shared by the whole GQA group.
//
// Work decomposition
//   CTA  = (sequence b, kv head h, split s): a contiguous run of 16-token chunks.
//   warp = private software pipeline over every 4th chunk of the CTA's run: its own
//          cp.async ring (DEC_STAGES x 8 KB), its own online-softmax state.  There is no
//          CTA-wide barrier in the main loop; the four warps merge once at the end.
//   GQA packing: the G = Hq/Hkv query heads of kv head h are the M rows (padded to 16) of
//          mma.sync m16n8k16, so a K/V chunk staged once in shared memory serves all of them.
// The M dimension is <= 8 real rows, far below a tcgen05 tile; the kernel is judged on HBM
// GB/s, not tensor throughput, so the legacy mma path is the right tool here.
//
// Replaces flash_attn_with_kvcache (attention.py:229-232) and the q_len==1 case of
// flash_attn_varlen_func (attention.py:242-248).
#include <stdlib.h>
#include <string.h>

#include "common.cuh"
#include "launch.h"

namespace b200 {

// Two CTA shapes, same math (template parameters of the kernel):
//   <4 warps, 3 stages>  96 KB, 2 CTAs/SM: 192 KB of loads in flight per SM, long contexts;
//   <2 warps, 3 stages>  48 KB, 4 CTAs/SM: short contexts / many small CTAs, where what limits the SM
//                        is the number of CTAs whose prologue (length -> block id -> first K bytes) and
//                        merge epilogue can overlap, not the bytes in flight.
constexpr int DEC_TOK = 16;                            // tokens per warp step
constexpr int DEC_TILE_BYTES = DEC_TOK * B200_HD * 2;  // 4 KB (K) ; same for V
constexpr int DEC_STAGE_BYTES = 2 * DEC_TILE_BYTES;    // 8 KB
constexpr int dec_smem_bytes(int warps, int stages) { return warps * stages * DEC_STAGE_BYTES; }
constexpr int DEC_MIN_SPLIT_TOKENS = 256;              // never split finer than one page
constexpr int DEC_CHUNKS_PER_PAGE = B200_PAGE / DEC_TOK;

__host__ __device__ __forceinline__ int dec_eff_splits(int len, int num_splits) {
    int m = (len + DEC_MIN_SPLIT_TOKENS - 1) / DEC_MIN_SPLIT_TOKENS;
    if (m < 1) m = 1;
    return m < num_splits ? m : num_splits;
}

struct DecodeParams {
    const __nv_bfloat16* q; int64_t q_token_stride;
    const uint8_t* k_cache; const uint8_t* v_cache;
    __nv_bfloat16* out; int64_t out_token_stride;
    const int32_t* context_lens; const int32_t* block_tables; int32_t bt_row_stride;
    int32_t Hq, Hkv, G; float scale_log2; int32_t num_splits;
    float* ws_o; float* ws_lse;
    int32_t* tickets;       // non-null: the last split of a (row, kv head) to arrive merges (no reduce launch)
    // varlen form (mixed batches): rows come from a device-built list, lengths from cu_seqlens
    const int32_t* cu_q; const int32_t* cu_k; const int32_t* row_list; const int32_t* n_rows;
};

// Resolve (list index -> sequence, its context length, the token index of its single query).
struct DecodeRow { int b, L, tok; bool valid; };
__device__ __forceinline__ DecodeRow decode_row(const DecodeParams& p, int idx) {
    DecodeRow r;
    r.valid = true;
    if (p.row_list != nullptr) {
        if (idx >= __ldg(p.n_rows)) { r.valid = false; r.b = 0; r.L = 0; r.tok = 0; return r; }
        r.b = __ldg(p.row_list + idx);
        r.L = __ldg(p.cu_k + r.b + 1) - __ldg(p.cu_k + r.b);
        r.tok = __ldg(p.cu_q + r.b);
    } else {
        r.b = idx; r.L = p.context_lens[idx]; r.tok = idx;
    }
    return r;
}

template <int DEC_WARPS, int DEC_STAGES, int MIN_CTAS>
__global__ void __launch_bounds__(DEC_WARPS * 32, MIN_CTAS) paged_decode_kernel(const DecodeParams p) {
    constexpr int DEC_THREADS = DEC_WARPS * 32;
    extern __shared__ __align__(1024) uint8_t smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int S = p.num_splits;
    const int s = blockIdx.x % S;
    const int bh = blockIdx.x / S;
    const int h = bh % p.Hkv;
    const int bi = bh / p.Hkv;                                  // row index (workspace slot)
    const int G = p.G;
    const DecodeRow row = decode_row(p, bi);
    if (!row.valid) return;
    const int b = row.b, L = row.L;
    const int64_t tok = row.tok;

    const int chunks_total = (L + DEC_TOK - 1) / DEC_TOK;
    const int n_eff = dec_eff_splits(L, S);
    const int per = (chunks_total + n_eff - 1) / n_eff;
    const int c_begin = s * per;
    const int c_end = min(chunks_total, c_begin + per);
    // splits of this row that hold tokens; the fused merge counts exactly these arrivals
    const int n_part = per > 0 ? (chunks_total + per - 1) / per : 0;
    const bool fused = p.tickets != nullptr;
    const bool direct = S == 1 || (fused && n_part == 1);       // this CTA alone owns the row: write `out` itself

    if (c_begin >= c_end) {                                   // empty split or padding row (len 0)
        if (S == 1 || (fused && s == 0)) {                    // (fused: s == 0 is empty only when the row is)
            for (int r = 0; r < G; ++r)
                for (int t = threadIdx.x; t < B200_HD; t += DEC_THREADS)
                    p.out[tok * p.out_token_stride + (int64_t)(h * G + r) * B200_HD + t] = __float2bfloat16(0.f);
        } else if (!fused && threadIdx.x < G) {
            p.ws_lse[((int64_t)bi * S + s) * p.Hq + h * G + threadIdx.x] = -INFINITY;
        }
        return;
    }

    // ---- Q fragments: rows = query heads of this kv head (zero rows beyond G) ------------------
    const int r0 = lane >> 2, r1 = r0 + 8, cq = (lane & 3) * 2;
    uint32_t qa[8][4];
    {
        const __nv_bfloat16* qb = p.q + tok * p.q_token_stride + (int64_t)h * G * B200_HD;
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
            qa[ks][0] = r0 < G ? *reinterpret_cast<const uint32_t*>(qb + r0 * B200_HD + ks * 16 + cq) : 0u;
            qa[ks][1] = r1 < G ? *reinterpret_cast<const uint32_t*>(qb + r1 * B200_HD + ks * 16 + cq) : 0u;
            qa[ks][2] = r0 < G ? *reinterpret_cast<const uint32_t*>(qb + r0 * B200_HD + ks * 16 + 8 + cq) : 0u;
            qa[ks][3] = r1 < G ? *reinterpret_cast<const uint32_t*>(qb + r1 * B200_HD + ks * 16 + 8 + cq) : 0u;
        }
    }

    const int64_t row_bytes = (int64_t)p.Hkv * B200_HD * 2;       // one token, all kv heads
    const int32_t* bt = p.block_tables + (int64_t)b * p.bt_row_stride;
    const uint32_t ring = smem_u32(smem) + warp * (DEC_STAGES * DEC_STAGE_BYTES);
    const int page_end = (c_end + DEC_CHUNKS_PER_PAGE - 1) / DEC_CHUNKS_PER_PAGE;
    int cur_page = -1, cur_block = 0, nxt_block = 0;

    const int n_my = (c_end - c_begin - warp + DEC_WARPS - 1) / DEC_WARPS;   // may be <= 0
    const int lr = lane >> 4, lc = lane & 15;

    auto issue = [&](int i, int stage) {
        const int chunk = c_begin + warp + i * DEC_WARPS;
        const int page = chunk / DEC_CHUNKS_PER_PAGE;
        if (page != cur_page) {                                   // one-page-ahead block id prefetch
            cur_block = (page == cur_page + 1 && cur_page >= 0) ? nxt_block : __ldg(bt + page);
            cur_page = page;
            nxt_block = (page + 1 < page_end) ? __ldg(bt + page + 1) : 0;
        }
        const int tok0 = chunk * DEC_TOK;
        const int64_t base = ((int64_t)cur_block * B200_PAGE + (tok0 & (B200_PAGE - 1))) * row_bytes + (int64_t)h * (B200_HD * 2);
        const uint8_t* kp = p.k_cache + base;
        const uint8_t* vp = p.v_cache + base;
        const uint32_t st = ring + stage * DEC_STAGE_BYTES;
#pragma unroll
        for (int it = 0; it < DEC_TOK / 2; ++it) {
            const int r = it * 2 + lr;
            const uint32_t nbytes = (tok0 + r < L) ? 16u : 0u;    // zero-fill rows past the context
            const uint32_t off = swz256(r, lc);
            cp_async16(st + off, kp + r * row_bytes + lc * 16, nbytes);
            cp_async16(st + DEC_TILE_BYTES + off, vp + r * row_bytes + lc * 16, nbytes);
        }
    };

    float o[16][4];
#pragma unroll
    for (int i = 0; i < 16; ++i) { o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f; }
    float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;

#pragma unroll
    for (int i = 0; i < DEC_STAGES - 1; ++i) {
        if (i < n_my) issue(i, i);
        cp_async_commit();
    }

    const int mat = lane >> 3, mr = lane & 7;
    for (int i = 0; i < n_my; ++i) {
        cp_async_wait<DEC_STAGES - 2>();
        __syncwarp();
        {
            const int nx = i + DEC_STAGES - 1;
            if (nx < n_my) issue(nx, nx % DEC_STAGES);
            cp_async_commit();
        }
        const uint32_t kst = ring + (i % DEC_STAGES) * DEC_STAGE_BYTES;
        const uint32_t vst = kst + DEC_TILE_BYTES;
        const int tok0 = (c_begin + warp + i * DEC_WARPS) * DEC_TOK;

        // ---- S = Q K^T  (16 heads x 16 tokens) -------------------------------------------------
        float sc[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
            uint32_t kb[4];
            ldmatrix_x4(kb, kst + swz256((mat >> 1) * 8 + mr, ks * 2 + (mat & 1)));
            mma_bf16_16816(sc[0], qa[ks], kb[0], kb[1]);
            mma_bf16_16816(sc[1], qa[ks], kb[2], kb[3]);
        }
        // ---- online softmax (base-2, scale folded) ---------------------------------------------
#pragma unroll
        for (int nt = 0; nt < 2; ++nt)
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int tok = tok0 + nt * 8 + cq + (j & 1);
                sc[nt][j] = tok < L ? sc[nt][j] * p.scale_log2 : -INFINITY;
            }
        float mx0 = fmaxf(fmaxf(sc[0][0], sc[0][1]), fmaxf(sc[1][0], sc[1][1]));
        float mx1 = fmaxf(fmaxf(sc[0][2], sc[0][3]), fmaxf(sc[1][2], sc[1][3]));
        mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
        mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
        const float mn0 = fmaxf(m0, mx0), mn1 = fmaxf(m1, mx1);          // finite: chunk has >= 1 valid token
        const float a0 = fast_exp2(m0 - mn0), a1 = fast_exp2(m1 - mn1);
        m0 = mn0; m1 = mn1;
        float ps0 = 0.f, ps1 = 0.f;
#pragma unroll
        for (int nt = 0; nt < 2; ++nt) {
            sc[nt][0] = fast_exp2(sc[nt][0] - mn0); sc[nt][1] = fast_exp2(sc[nt][1] - mn0);
            sc[nt][2] = fast_exp2(sc[nt][2] - mn1); sc[nt][3] = fast_exp2(sc[nt][3] - mn1);
            ps0 += sc[nt][0] + sc[nt][1];
            ps1 += sc[nt][2] + sc[nt][3];
        }
        l0 = l0 * a0 + ps0;
        l1 = l1 * a1 + ps1;
#pragma unroll
        for (int nt = 0; nt < 16; ++nt) { o[nt][0] *= a0; o[nt][1] *= a0; o[nt][2] *= a1; o[nt][3] *= a1; }
        // ---- O += P V  (P rounded to bf16, as flash-attn does) ---------------------------------
        uint32_t pa[4];
        pa[0] = pack_bf16x2(sc[0][0], sc[0][1]);
        pa[1] = pack_bf16x2(sc[0][2], sc[0][3]);
        pa[2] = pack_bf16x2(sc[1][0], sc[1][1]);
        pa[3] = pack_bf16x2(sc[1][2], sc[1][3]);
#pragma unroll
        for (int np = 0; np < 8; ++np) {
            uint32_t vb[4];
            ldmatrix_x4_trans(vb, vst + swz256((mat & 1) * 8 + mr, np * 2 + (mat >> 1)));
            mma_bf16_16816(o[2 * np], pa, vb[0], vb[1]);
            mma_bf16_16816(o[2 * np + 1], pa, vb[2], vb[3]);
        }
    }
    cp_async_wait<0>();

    l0 += __shfl_xor_sync(0xffffffffu, l0, 1); l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 1); l1 += __shfl_xor_sync(0xffffffffu, l1, 2);

    // ---- merge the four warps through shared memory (ring is dead now) --------------------------
    __syncthreads();
    float* sm_m = reinterpret_cast<float*>(smem);                 // [4][16]
    float* sm_l = sm_m + DEC_WARPS * 16;                          // [4][16]
    float* sm_o = sm_l + DEC_WARPS * 16;                          // [4][16][128]
    if ((lane & 3) == 0) {
        sm_m[warp * 16 + r0] = m0; sm_l[warp * 16 + r0] = l0;
        sm_m[warp * 16 + r1] = m1; sm_l[warp * 16 + r1] = l1;
    }
#pragma unroll
    for (int nt = 0; nt < 16; ++nt) {
        if (r0 < G) *reinterpret_cast<float2*>(sm_o + (warp * 16 + r0) * B200_HD + nt * 8 + cq) = make_float2(o[nt][0], o[nt][1]);
        if (r1 < G) *reinterpret_cast<float2*>(sm_o + (warp * 16 + r1) * B200_HD + nt * 8 + cq) = make_float2(o[nt][2], o[nt][3]);
    }
    __syncthreads();
    for (int idx = threadIdx.x; idx < G * B200_HD; idx += DEC_THREADS) {   // one thread per (head, head-dim element)
        const int r = idx >> 7, t = idx & (B200_HD - 1);
        float ms = -INFINITY;
#pragma unroll
        for (int w = 0; w < DEC_WARPS; ++w) ms = fmaxf(ms, sm_m[w * 16 + r]);
        float acc = 0.f, den = 0.f;
#pragma unroll
        for (int w = 0; w < DEC_WARPS; ++w) {
            const float wgt = fast_exp2(sm_m[w * 16 + r] - ms);   // 0 for idle warps (m = -inf)
            acc += wgt * sm_o[(w * 16 + r) * B200_HD + t];
            den += wgt * sm_l[w * 16 + r];
        }
        const float val = acc / den;
        const int head = h * G + r;
        if (direct) {
            p.out[tok * p.out_token_stride + (int64_t)head * B200_HD + t] = __float2bfloat16_rn(val);
        } else {
            const int64_t widx = ((int64_t)bi * S + s) * p.Hq + head;
            p.ws_o[widx * B200_HD + t] = val;
            if (t == 0) p.ws_lse[widx] = ms + log2f(den);
        }
    }
    if (direct || !fused) return;

    // ---- fused split merge: the last of this (row, kv head)'s n_part CTAs to arrive reduces them all ----
    volatile int* s_last = reinterpret_cast<volatile int*>(smem);   // the warp-merge scratch is dead after the barrier
    __syncthreads();                                          // every thread's partials are ordered before thread 0 here,
    if (threadIdx.x == 0) {
        __threadfence();                                      // ... so ONE cumulative fence releases them with the ticket
        int32_t* ticket = p.tickets + (int64_t)bi * p.Hkv + h;
        const int t = atomicAdd(ticket, 1);
        *s_last = (t == n_part - 1);
        if (t == n_part - 1) *ticket = 0;                     // self re-arming: every launch finds zeros (graph replay)
        __threadfence();
    }
    __syncthreads();
    if (!*s_last) return;
    // Merge with every load in flight at once (a chain of dependent L2 round trips here would serialise the tail
    // of the whole launch): (1) all LSEs into shared memory, (2) one thread per head turns them into weights,
    // (3) 16-byte partial-O loads, unrolled over the splits.
    float* sm_w = reinterpret_cast<float*>(smem) + 32;            // [n_part][16] weights, normalised (sum to 1 per head)
    __syncthreads();                                              // everyone has read *s_last
    for (int i = threadIdx.x; i < n_part * G; i += DEC_THREADS) {
        const int ss = i / G, r = i - ss * G;
        sm_w[ss * 16 + r] = __ldcg(p.ws_lse + ((int64_t)bi * S + ss) * p.Hq + h * G + r);
    }
    __syncthreads();
    if (threadIdx.x < G) {
        const int r = threadIdx.x;
        float mx = -INFINITY;
        for (int ss = 0; ss < n_part; ++ss) mx = fmaxf(mx, sm_w[ss * 16 + r]);
        float den = 0.f;
        for (int ss = 0; ss < n_part; ++ss) {
            const float w = fast_exp2(sm_w[ss * 16 + r] - mx);
            sm_w[ss * 16 + r] = w;
            den += w;
        }
        const float inv = 1.f / den;
        for (int ss = 0; ss < n_part; ++ss) sm_w[ss * 16 + r] *= inv;
    }
    __syncthreads();
    for (int idx = threadIdx.x; idx < G * (B200_HD / 4); idx += DEC_THREADS) {    // one thread per (head, 4 dims)
        const int r = idx >> 5, t4 = idx & 31;
        const int head = h * G + r;
        const float4* src = reinterpret_cast<const float4*>(p.ws_o + (((int64_t)bi * S) * p.Hq + head) * B200_HD) + t4;
        const int64_t split_stride = (int64_t)p.Hq * (B200_HD / 4);               // float4 units between splits
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        int ss = 0;
        for (; ss + 8 <= n_part; ss += 8) {
            float4 v[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) v[u] = __ldcg(src + (ss + u) * split_stride);
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const float w = sm_w[(ss + u) * 16 + r];
                acc.x += w * v[u].x; acc.y += w * v[u].y; acc.z += w * v[u].z; acc.w += w * v[u].w;
            }
        }
        {
            float4 v[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) if (ss + u < n_part) v[u] = __ldcg(src + (ss + u) * split_stride);
#pragma unroll
            for (int u = 0; u < 8; ++u)
                if (ss + u < n_part) {
                    const float w = sm_w[(ss + u) * 16 + r];
                    acc.x += w * v[u].x; acc.y += w * v[u].y; acc.z += w * v[u].z; acc.w += w * v[u].w;
                }
        }
        uint2 o;
        o.x = pack_bf16x2(acc.x, acc.y); o.y = pack_bf16x2(acc.z, acc.w);
        *reinterpret_cast<uint2*>(p.out + tok * p.out_token_stride + (int64_t)head * B200_HD + t4 * 4) = o;
    }
}

// Merge the splits of one (sequence, head): out = sum_s 2^(lse_s - max) o_s / sum_s 2^(lse_s - max).
__global__ void __launch_bounds__(B200_HD) decode_reduce_kernel(const DecodeParams p, int32_t batch) {
    const int bi = blockIdx.x / p.Hq, head = blockIdx.x % p.Hq, t = threadIdx.x;
    const int S = p.num_splits;
    const DecodeRow row = decode_row(p, bi);
    if (!row.valid) return;
    const int n_eff = dec_eff_splits(row.L, S);
    const float* lse = p.ws_lse + ((int64_t)bi * S) * p.Hq + head;
    float mx = -INFINITY;
    for (int s = 0; s < n_eff; ++s) mx = fmaxf(mx, lse[(int64_t)s * p.Hq]);
    float acc = 0.f, den = 0.f;
    if (mx > -INFINITY) {
        for (int s = 0; s < n_eff; ++s) {
            const float ls = lse[(int64_t)s * p.Hq];
            if (ls == -INFINITY) continue;
            const float w = fast_exp2(ls - mx);
            acc += w * p.ws_o[(((int64_t)bi * S + s) * p.Hq + head) * B200_HD + t];
            den += w;
        }
    }
    const float val = den > 0.f ? acc / den : 0.f;
    p.out[(int64_t)row.tok * p.out_token_stride + (int64_t)head * B200_HD + t] = __float2bfloat16_rn(val);
}

template <int W, int ST, int MC>
static int launch_decode_shape(const DecodeParams& p, int64_t grid, cudaStream_t stream) {
    static PerDeviceOnce once;                                 // the opt-in is per device, not per process
    constexpr int SMEM = dec_smem_bytes(W, ST);
    if (once.need()) {
        cudaError_t e = cudaFuncSetAttribute(paged_decode_kernel<W, ST, MC>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
        if (e != cudaSuccess) return (int)e;
        once.done();
    }
    paged_decode_kernel<W, ST, MC><<<(unsigned)grid, W * 32, SMEM, stream>>>(p);
    return (int)cudaGetLastError();
}

// CTA shape of a launch: host-only (static under CUDA graphs).  B200ATTN_DECODE_SHAPE=w4s3|w2s3|w4s2|w2s2 overrides.
int decode_shape(int64_t grid_ctas) {
    static int forced = -2;
    if (forced == -2) {
        const char* env = getenv("B200ATTN_DECODE_SHAPE");
        forced = -1;
        if (env) {
            if (!strcmp(env, "w4s3")) forced = 0;
            else if (!strcmp(env, "w2s3")) forced = 1;
            else if (!strcmp(env, "w4s2")) forced = 2;
            else if (!strcmp(env, "w2s2")) forced = 3;
        }
    }
    if (forced >= 0) return forced;
    (void)grid_ctas;
    return 0;
}

int launch_attn_decode(const DecodeArgs& a, cudaStream_t stream) {
    DecodeParams p;
    p.q = (const __nv_bfloat16*)a.q; p.q_token_stride = a.q_token_stride;
    p.k_cache = (const uint8_t*)a.k_cache; p.v_cache = (const uint8_t*)a.v_cache;
    p.out = (__nv_bfloat16*)a.out; p.out_token_stride = a.out_token_stride;
    p.context_lens = a.context_lens; p.block_tables = a.block_tables; p.bt_row_stride = a.bt_row_stride;
    p.Hq = a.Hq; p.Hkv = a.Hkv; p.G = a.Hq / a.Hkv;
    p.scale_log2 = a.scale * 1.4426950408889634f;
    p.num_splits = a.num_splits; p.ws_o = a.ws_o; p.ws_lse = a.ws_lse;
    p.tickets = a.num_splits > 1 ? a.tickets : nullptr;
    p.cu_q = a.cu_q; p.cu_k = a.cu_k; p.row_list = a.row_list; p.n_rows = a.n_rows;
    if (a.batch == 0) return 0;
    const int64_t grid = (int64_t)a.batch * a.Hkv * a.num_splits;
    int rc;
    switch (decode_shape(grid)) {
        case 1: rc = launch_decode_shape<2, 3, 4>(p, grid, stream); break;
        case 2: rc = launch_decode_shape<4, 2, 2>(p, grid, stream); break;
        case 3: rc = launch_decode_shape<2, 2, 5>(p, grid, stream); break;
        default: rc = launch_decode_shape<4, 3, 2>(p, grid, stream); break;
    }
    if (rc != 0) return rc;
    cudaError_t e = cudaSuccess;
    if (a.num_splits > 1 && p.tickets == nullptr) {           // legacy two-launch form (B200ATTN_DECODE_FUSED_REDUCE=0)
        decode_reduce_kernel<<<(unsigned)(a.batch * a.Hq), B200_HD, 0, stream>>>(p, a.batch);
        e = cudaGetLastError();
    }
    return (int)e;
}

}  // namespace b200
