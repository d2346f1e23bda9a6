// (3) Split-K paged decode attention + LSE reduce.  HBM-bound: every K/V byte of every
// context token is read exactly once per KV head and shared by the whole GQA group.
//
// Work decomposition
//   item = (sequence b, kv head h, split s): a contiguous run of 16-token chunks.
//   CTA  = persistent: 2 CTAs per SM loop over items.  The first three rounds are taken statically
//          (item = round * grid + cta), later ones from an atomic counter in the workspace header, so
//          ragged batches balance dynamically.  Metadata runs three items deep (cur: streaming,
//          next: length known + first block id loading, next-next: length loading), so the
//          dependent chain  context length -> block id -> first K bytes  of an item is paid while the
//          previous item streams, and a warp's cp.async ring never drains between items.
//   warp = private software pipeline over every 4th chunk of an item: its own cp.async ring
//          (STAGES x 8 KB) that runs ahead ACROSS item boundaries, its own online-softmax state, its
//          own staged copy of the item's query rows.  No CTA-wide barrier in the main loop; the four
//          warps merge once per item through a small dedicated shared-memory buffer.
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

constexpr int DEC_WARPS = 4;
constexpr int DEC_THREADS = DEC_WARPS * 32;
constexpr int DEC_TOK = 16;                            // tokens per warp step
constexpr int DEC_TILE_BYTES = DEC_TOK * B200_HD * 2;  // 4 KB (K) ; same for V
constexpr int DEC_STAGE_BYTES = 2 * DEC_TILE_BYTES;    // 8 KB
constexpr int DEC_MIN_SPLIT_TOKENS = 256;              // never split finer than one page
constexpr int DEC_CHUNKS_PER_PAGE = B200_PAGE / DEC_TOK;
constexpr int DEC_STATIC_ROUNDS = 3;                   // rounds assigned without the atomic counter

// shared memory of one CTA: [rings: WARPS x STAGES x 8 KB][q: WARPS x G x 256 B][merge m,l: 2 x WARPS x 16 f32]
//                           [merge o: WARPS x G x 128 f32][claim slots: 4 x int]
__host__ __device__ constexpr int dec_ring_bytes(int stages) { return DEC_WARPS * stages * DEC_STAGE_BYTES; }
__host__ __device__ inline int dec_smem_bytes(int stages, int G) {
    return dec_ring_bytes(stages) + DEC_WARPS * G * B200_HD * 2 + 2 * DEC_WARPS * 16 * 4 + DEC_WARPS * G * B200_HD * 4 + 16;
}

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
    uint32_t* sched;                                   // [0] next dynamic item, [1] CTAs done (both 0 between launches)
    int32_t total_items;                               // batch * Hkv * num_splits
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

// Everything a warp needs to stream one item.
struct DecItem {
    int valid;                 // 0 => end of the work stream (items are claimed in increasing order)
    int bi, b, h, s, L, tok;   // workspace row, sequence, kv head, split, context length, query token
    int c_begin, n_my;         // first chunk of the item; chunks this warp owns (may be 0)
    int page_end;              // one past the last page the item touches
    int empty;                 // no chunk at all (padding row or empty split)
    int blk0;                  // block id of the page holding this warp's first chunk
};

// phase 1 (ids + the loads that depend only on the item index)
__device__ __forceinline__ void dec_item_load(const DecodeParams& p, int item, DecItem& it) {
    it.valid = item < p.total_items;
    const int S = p.num_splits;
    const int bh = item / S;
    it.s = item - bh * S;
    it.bi = bh / p.Hkv;
    it.h = bh - it.bi * p.Hkv;
    it.b = 0; it.L = 0; it.tok = 0;
    if (it.valid) {
        const DecodeRow r = decode_row(p, it.bi);
        it.valid = r.valid;
        it.b = r.b; it.L = r.L; it.tok = r.tok;
    }
}
// phase 2 (needs L): chunk range of this warp and the load of its first block id
__device__ __forceinline__ void dec_item_geometry(const DecodeParams& p, int warp, DecItem& it) {
    const int chunks_total = (it.L + DEC_TOK - 1) / DEC_TOK;
    const int n_eff = dec_eff_splits(it.L, p.num_splits);
    const int per = (chunks_total + n_eff - 1) / n_eff;
    it.c_begin = it.s * per;
    const int c_end = min(chunks_total, it.c_begin + per);
    it.empty = (!it.valid) || it.c_begin >= c_end;
    it.n_my = it.empty ? 0 : max(0, (c_end - it.c_begin - warp + DEC_WARPS - 1) / DEC_WARPS);
    it.page_end = (c_end + DEC_CHUNKS_PER_PAGE - 1) / DEC_CHUNKS_PER_PAGE;
    it.blk0 = 0;
    if (it.n_my > 0)
        it.blk0 = __ldg(p.block_tables + (int64_t)it.b * p.bt_row_stride + (it.c_begin + warp) / DEC_CHUNKS_PER_PAGE);
}

template <int STAGES>
__global__ void __launch_bounds__(DEC_THREADS, 2) paged_decode_kernel(const DecodeParams p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int G = p.G, S = p.num_splits;
    const int64_t row_bytes = (int64_t)p.Hkv * B200_HD * 2;       // one token, all kv heads

    const uint32_t ring = smem_u32(smem) + warp * (STAGES * DEC_STAGE_BYTES);
    uint8_t* q_gen = smem + dec_ring_bytes(STAGES) + warp * (G * B200_HD * 2);   // this warp's staged query rows
    const uint32_t q_s = smem_u32(q_gen);
    float* sm_m = reinterpret_cast<float*>(smem + dec_ring_bytes(STAGES) + DEC_WARPS * G * B200_HD * 2);   // [4][16]
    float* sm_l = sm_m + DEC_WARPS * 16;                          // [4][16]
    float* sm_o = sm_l + DEC_WARPS * 16;                          // [4][G][128]
    volatile int* claim = reinterpret_cast<volatile int*>(sm_o + DEC_WARPS * G * B200_HD);   // [4]

    const int grid = (int)gridDim.x;
    DecItem cur, nx, nn;
    dec_item_load(p, (int)blockIdx.x, cur);
    dec_item_load(p, grid + (int)blockIdx.x, nx);
    dec_item_load(p, 2 * grid + (int)blockIdx.x, nn);
    dec_item_geometry(p, warp, cur);
    dec_item_geometry(p, warp, nx);

    const int r0 = lane >> 2, r1 = r0 + 8, cq = (lane & 3) * 2;
    const int lr = lane >> 4, lc = lane & 15;
    const int mat = lane >> 3, mr = lane & 7;

    // ---- producer: this warp's cp.async stream, running ahead of the consumer across items ------
    int p_next = 0;                    // 0: issuing chunks of `cur`, 1: of `nx`
    int pc = 0;                        // next chunk (of this warp) to issue within that item
    int p_page = -1, p_blk = 0, p_nblk = 0;
    int slot = 0, cslot = 0;           // ring slot of the next issue / next consume

    auto stage_q = [&](const DecItem& it) {                       // G rows x 256 B -> this warp's q buffer
        if (!it.valid) return;
        const uint8_t* qb = reinterpret_cast<const uint8_t*>(p.q + (int64_t)it.tok * p.q_token_stride + (int64_t)it.h * G * B200_HD);
        for (int c = lane; c < G * 16; c += 32) cp_async16(q_s + c * 16, qb + c * 16, 16u);
    };
    auto issue = [&](const DecItem& it, int i, int stage) {
        const int chunk = it.c_begin + warp + i * DEC_WARPS;
        const int page = chunk / DEC_CHUNKS_PER_PAGE;
        const int32_t* bt = p.block_tables + (int64_t)it.b * p.bt_row_stride;
        if (i == 0 || page != p_page) {                           // one-page-ahead block id prefetch
            p_blk = (i == 0) ? it.blk0 : ((page == p_page + 1) ? p_nblk : __ldg(bt + page));
            p_page = page;
            p_nblk = (page + 1 < it.page_end) ? __ldg(bt + page + 1) : 0;
        }
        const int tok0 = chunk * DEC_TOK;
        const int64_t base = ((int64_t)p_blk * B200_PAGE + (tok0 & (B200_PAGE - 1))) * row_bytes + (int64_t)it.h * (B200_HD * 2);
        const uint8_t* kp = p.k_cache + base;
        const uint8_t* vp = p.v_cache + base;
        const uint32_t st = ring + stage * DEC_STAGE_BYTES;
#pragma unroll
        for (int t = 0; t < DEC_TOK / 2; ++t) {
            const int r = t * 2 + lr;
            const uint32_t nbytes = (tok0 + r < it.L) ? 16u : 0u; // zero-fill rows past the context
            const uint32_t off = swz256(r, lc);
            cp_async16(st + off, kp + r * row_bytes + lc * 16, nbytes);
            cp_async16(st + DEC_TILE_BYTES + off, vp + r * row_bytes + lc * 16, nbytes);
        }
    };
    auto produce_one = [&]() {
        if (p_next == 0 && pc >= cur.n_my) { p_next = 1; pc = 0; }
        if (p_next == 0) {
            issue(cur, pc, slot);
        } else {
            if (pc >= nx.n_my) return;                            // nothing further to run ahead into
            issue(nx, pc, slot);
        }
        ++pc;
        slot = (slot + 1 == STAGES) ? 0 : slot + 1;
    };

    // ---- prologue: query rows of the first item, then fill the ring ----------------------------
    stage_q(cur);
    cp_async_commit();
    cp_async_wait<0>();
    __syncwarp();

    for (int round = 0; cur.valid; ++round) {
        // claim the item of round + 3 (visible to everyone after this item's barriers)
        if (threadIdx.x == 0)
            claim[(round + 3) & 3] = nn.valid ? (int)(DEC_STATIC_ROUNDS * grid + atomicAdd(p.sched, 1u)) : p.total_items;

        // ---- Q fragments of `cur` from the staged copy; then stage the next item's query rows ----
        uint32_t qa[8][4];
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
            qa[ks][0] = r0 < G ? *reinterpret_cast<const uint32_t*>(q_gen + (r0 * B200_HD + ks * 16 + cq) * 2) : 0u;
            qa[ks][1] = r1 < G ? *reinterpret_cast<const uint32_t*>(q_gen + (r1 * B200_HD + ks * 16 + cq) * 2) : 0u;
            qa[ks][2] = r0 < G ? *reinterpret_cast<const uint32_t*>(q_gen + (r0 * B200_HD + ks * 16 + 8 + cq) * 2) : 0u;
            qa[ks][3] = r1 < G ? *reinterpret_cast<const uint32_t*>(q_gen + (r1 * B200_HD + ks * 16 + 8 + cq) * 2) : 0u;
        }
        __syncwarp();
        stage_q(nx);                                              // rides in the next commit group
        if (round == 0) {
#pragma unroll
            for (int i = 0; i < STAGES - 1; ++i) { produce_one(); cp_async_commit(); }
        }

        float o[16][4];
#pragma unroll
        for (int i = 0; i < 16; ++i) { o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f; }
        float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;

        const int n_my = cur.n_my;
        for (int i = 0; i < n_my; ++i) {
            cp_async_wait<STAGES - 2>();
            __syncwarp();
            produce_one();
            cp_async_commit();
            const uint32_t kst = ring + cslot * DEC_STAGE_BYTES;
            const uint32_t vst = kst + DEC_TILE_BYTES;
            cslot = (cslot + 1 == STAGES) ? 0 : cslot + 1;
            const int tok0 = (cur.c_begin + warp + i * DEC_WARPS) * DEC_TOK;

            // ---- S = Q K^T  (16 heads x 16 tokens) ---------------------------------------------
            float sc[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
#pragma unroll
            for (int ks = 0; ks < 8; ++ks) {
                uint32_t kb[4];
                ldmatrix_x4(kb, kst + swz256((mat >> 1) * 8 + mr, ks * 2 + (mat & 1)));
                mma_bf16_16816(sc[0], qa[ks], kb[0], kb[1]);
                mma_bf16_16816(sc[1], qa[ks], kb[2], kb[3]);
            }
            // ---- online softmax (base-2, scale folded) -----------------------------------------
#pragma unroll
            for (int nt = 0; nt < 2; ++nt)
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int tk = tok0 + nt * 8 + cq + (j & 1);
                    sc[nt][j] = tk < cur.L ? sc[nt][j] * p.scale_log2 : -INFINITY;
                }
            float mx0 = fmaxf(fmaxf(sc[0][0], sc[0][1]), fmaxf(sc[1][0], sc[1][1]));
            float mx1 = fmaxf(fmaxf(sc[0][2], sc[0][3]), fmaxf(sc[1][2], sc[1][3]));
            mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
            mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
            mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
            mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
            const float mn0 = fmaxf(m0, mx0), mn1 = fmaxf(m1, mx1);      // finite: chunk has >= 1 valid token
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
            // ---- O += P V  (P rounded to bf16, as flash-attn does) -----------------------------
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
        if (n_my < STAGES) {           // the staged query rows of the next item may still be in flight
            cp_async_commit();
            cp_async_wait<0>();
        }

        // ---- merge the four warps (dedicated buffer: the rings already hold the next item) --------
        const int64_t tok = cur.tok;
        const int h = cur.h;
        if (cur.empty) {                                          // padding row (len 0) or empty split
            if (S == 1) {
                for (int idx = threadIdx.x; idx < G * B200_HD; idx += DEC_THREADS)
                    p.out[tok * p.out_token_stride + (int64_t)h * G * B200_HD + idx] = __float2bfloat16(0.f);
            } else if (threadIdx.x < G) {
                p.ws_lse[((int64_t)cur.bi * S + cur.s) * p.Hq + h * G + threadIdx.x] = -INFINITY;
            }
            __syncthreads();                                      // publishes the claim slot
        } else {
            l0 += __shfl_xor_sync(0xffffffffu, l0, 1); l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
            l1 += __shfl_xor_sync(0xffffffffu, l1, 1); l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
            if ((lane & 3) == 0) {
                sm_m[warp * 16 + r0] = m0; sm_l[warp * 16 + r0] = l0;
                sm_m[warp * 16 + r1] = m1; sm_l[warp * 16 + r1] = l1;
            }
#pragma unroll
            for (int nt = 0; nt < 16; ++nt) {
                if (r0 < G) *reinterpret_cast<float2*>(sm_o + (warp * G + r0) * B200_HD + nt * 8 + cq) = make_float2(o[nt][0], o[nt][1]);
                if (r1 < G) *reinterpret_cast<float2*>(sm_o + (warp * G + r1) * B200_HD + nt * 8 + cq) = make_float2(o[nt][2], o[nt][3]);
            }
            __syncthreads();
            for (int idx = threadIdx.x; idx < G * B200_HD; idx += DEC_THREADS) {   // one thread per (head, element)
                const int r = idx >> 7, t = idx & (B200_HD - 1);
                float ms = -INFINITY;
#pragma unroll
                for (int w = 0; w < DEC_WARPS; ++w) ms = fmaxf(ms, sm_m[w * 16 + r]);
                float acc = 0.f, den = 0.f;
#pragma unroll
                for (int w = 0; w < DEC_WARPS; ++w) {
                    const float wgt = fast_exp2(sm_m[w * 16 + r] - ms);   // 0 for idle warps (m = -inf)
                    acc += wgt * sm_o[(w * G + r) * B200_HD + t];
                    den += wgt * sm_l[w * 16 + r];
                }
                const float val = acc / den;
                const int head = h * G + r;
                if (S == 1) {
                    p.out[tok * p.out_token_stride + (int64_t)head * B200_HD + t] = __float2bfloat16_rn(val);
                } else {
                    const int64_t widx = ((int64_t)cur.bi * S + cur.s) * p.Hq + head;
                    p.ws_o[widx * B200_HD + t] = val;
                    if (t == 0) p.ws_lse[widx] = ms + log2f(den);
                }
            }
            __syncthreads();                                      // buffer free for the next item; claim slot published
        }

        // ---- shift the item pipeline ---------------------------------------------------------------
        const bool crossed = p_next == 1;                         // the producer already runs inside `nx`
        cur = nx;
        nx = nn;
        dec_item_geometry(p, warp, nx);
        dec_item_load(p, claim[(round + 3) & 3], nn);
        if (crossed) {
            p_next = 0;                                           // pc / page state carry over
        } else {
            pc = 0;                                               // (cur had all its chunks issued; none of nx yet)
        }
    }

    cp_async_wait<0>();
    if (threadIdx.x == 0) {                                       // the last CTA out re-arms the scheduler words
        __threadfence();
        const uint32_t done = atomicAdd(p.sched + 1, 1u);
        if (done == gridDim.x - 1) {
            p.sched[0] = 0u;
            p.sched[1] = 0u;
        }
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

template <int STAGES>
static int launch_decode_stages(const DecodeParams& p, int grid, int smem_bytes, cudaStream_t stream) {
    static int attr_bytes = 0;
    if (smem_bytes > attr_bytes) {
        cudaError_t e = cudaFuncSetAttribute(paged_decode_kernel<STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
        if (e != cudaSuccess) return (int)e;
        attr_bytes = smem_bytes;
    }
    paged_decode_kernel<STAGES><<<(unsigned)grid, DEC_THREADS, smem_bytes, stream>>>(p);
    return (int)cudaGetLastError();
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
    p.sched = a.sched;
    p.cu_q = a.cu_q; p.cu_k = a.cu_k; p.row_list = a.row_list; p.n_rows = a.n_rows;
    if (a.batch == 0) return 0;
    const int64_t items = (int64_t)a.batch * a.Hkv * a.num_splits;
    if (items > 0x7fffffff / 2) return -5;
    p.total_items = (int32_t)items;
    // 3 ring stages (96 KB) while two CTAs still fit an SM next to the q/merge buffers, else 2 stages
    int stages = 3;
    int smem = dec_smem_bytes(3, p.G);
    if (2 * (smem + 1024) > 227 * 1024) { stages = 2; smem = dec_smem_bytes(2, p.G); }
    const char* env = getenv("B200ATTN_DECODE_STAGES");
    if (env && atoi(env) == 2) { stages = 2; smem = dec_smem_bytes(2, p.G); }
    int64_t grid = (int64_t)sm_count() * 2;
    if (grid > items) grid = items;
    int rc = stages == 3 ? launch_decode_stages<3>(p, (int)grid, smem, stream) : launch_decode_stages<2>(p, (int)grid, smem, stream);
    if (rc != 0) return rc;
    cudaError_t e = cudaSuccess;
    if (a.num_splits > 1) {
        decode_reduce_kernel<<<(unsigned)(a.batch * a.Hq), B200_HD, 0, stream>>>(p, a.batch);
        e = cudaGetLastError();
    }
    return (int)e;
}

}  // namespace b200
