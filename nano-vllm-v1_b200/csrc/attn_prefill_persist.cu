// (2, variant 4) Persistent varlen causal prefill / chunked prefill on Blackwell tensor cores.
//
// Same tile math as attn_prefill_tc.cu (TMA-staged Q/K/V, tcgen05.mma with S and O in TMEM, P fed
// back to the tensor core from TMEM, two 128-row query tiles ping-ponging per CTA), restructured
// around a device-built work list (attn_plan.cu):
//   * one CTA per SM loops over work units (two 128-row query tiles sharing one K/V stream: the same
//     rows of two q heads of a GQA group, or two consecutive row tiles of one head) taken round-robin from a list sorted by descending KV length, so a causal batch is load-balanced
//     and rows with a single query token never reach this kernel (they go to the decode kernel);
//   * barrier phases run on counters that survive across units, so the producer prefetches the
//     next unit's Q/K/V while the current unit drains and TMEM/smem/barriers are set up once;
//   * setmaxnreg moves registers from the producer/MMA warps to the softmax warpgroups
//     (a softmax thread keeps its whole 128-column score row in registers, no spills);
//   * softmax issue slots are cut with packed f32x2 FMA/ADD, 3-input max and a single FFMA that
//     folds scale and row maximum into the exponent argument.
//
// Warp roles (384 threads): warp 0 TMA producer, warp 1 MMA issuer, warp 2 TMEM allocator,
// warp 3 idle, warps 4-7 / 8-11 softmax + correction + epilogue of query tile 0 / 1.
//
// Replaces flash_attn_varlen_func (attention.py:242-248), paged and unpaged.
#include "launch.h"
#include "tc_ptx.cuh"

namespace b200 {

struct PersistParams {
    const int32_t* cu_q; const int32_t* cu_k; const int32_t* block_tables; int32_t bt_row_stride;
    __nv_bfloat16* out; int64_t out_token_stride;
    int32_t Hq, G, pair_heads; float scale_log2;
    uint32_t g_magic16;                            // ceil(2^16 / G): head / G by multiply-high
    const int32_t* plan; const int32_t* items;
    long long* trace; int32_t trace_tiles;      // TRACE builds only: clock64 stamps of CTA 0 (b200_debug_prefill_trace)
};

#ifndef PS_PACK
#define PS_PACK 1        // two q heads per 128-row tile for query tiles with <= 64 valid rows (build knob)
#endif
constexpr int PREFILL_EMU_DEFAULT = 0;
constexpr int PREFILL_WGS_DEFAULT = 1;

struct PsSmem {
    static constexpr int NQ = 2;
    static constexpr int Q_OFF = 0;
    static constexpr int K_OFF = Q_OFF + NQ * TC_TILE_BYTES;
    static constexpr int V_OFF = K_OFF + TC_KV_STAGES * TC_TILE_BYTES;
    static constexpr int BAR_OFF = V_OFF + TC_KV_STAGES * TC_TILE_BYTES;
    // q_full[2] q_empty[2] k_full[S] k_empty[S] v_full[S] v_empty[S] s_full[2] p_full[2] o_final[2] p_half[2]
    static constexpr int NUM_BARS = 6 * NQ + 4 * TC_KV_STAGES;
    static constexpr int TMEM_PTR_OFF = BAR_OFF + NUM_BARS * 8;
    // WGS == 2: row-maximum / row-sum exchange between the two half-row warpgroups of a query tile,
    // float[2 buffers][2 tiles][2 halves][128 rows]
    static constexpr int XCH_OFF = TMEM_PTR_OFF + 16;
    // WGS == 1: output staging, one [128 rows][64 bf16] half tile (16 KB, 16-byte chunks XOR-swizzled by row) per
    // softmax warpgroup, so the epilogue's global stores are whole 128-byte lines.  Same bytes as XCH (never both).
    static constexpr int STAGE_OFF = (XCH_OFF + 127) & ~127;
    static constexpr int STAGE_BYTES = 128 * 128;
    static constexpr int TOTAL = STAGE_OFF + 2 * STAGE_BYTES;
    static constexpr int ALLOC = TOTAL + 1024;
};

// One work unit = two 128-row query tiles that share one stream of K/V tiles:
//   head-pair mode (even GQA group): the SAME rows of two q heads of one kv head (128-row items);
//   row-pair mode  (odd group, e.g. 5): two consecutive row tiles of one q head (256-row items).
//   packed head-pair mode (group divisible by 4, a query tile with <= 64 rows inside its sequence -- the tail tile of
//   a chunk, a 64-token branch of a shared prefix): every 128-row tile holds 64 rows of TWO q heads (rows 0..63 head h,
//   rows 64..127 head h+1), so the unit covers four q heads and no MMA row is wasted on padding; the units of the
//   item's upper half (sub >= Hq/4) are empty.
struct Unit {
    int b, q0, Lq, k0, Lk, kvh, head0, head1, r0, r1, n_t0, n_t1, n_max, pack;
    __device__ __forceinline__ int nt(int t) const { return t == 0 ? n_t0 : n_t1; }
    __device__ __forceinline__ int head(int t) const { return t == 0 ? head0 : head1; }
    __device__ __forceinline__ int rbegin(int t) const { return t == 0 ? r0 : r1; }
};

// Round k of CTA c takes unit snake(k, c): the list is sorted heaviest-first, so alternating the
// direction of every other round pairs a CTA's heavy units with light ones (static LPT-like balance).
__device__ __forceinline__ int snake(int k, int c, int g) { return (k & 1) ? (k + 1) * g - 1 - c : k * g + c; }

// Work unit w of this launch.  Every role walks the same list on its own.  A unit's record is fetched in one
// load level (two 16-byte loads of the plan's item record) one unit AHEAD: on this part an L2 hit costs
// ~800 cycles, and three dependent levels (count -> item -> cu_seqlens) in front of every unit left the tensor
// pipe idle for ~3,000 cycles per unit while the MMA issuer found out what to do next.
struct RawUnit { int4 a; int4 b; };          // {sequence, tile, q0, Lq} {k0, Lk, sub, pack}: one record of the plan's unit table
__device__ __forceinline__ RawUnit fetch_unit(const PersistParams& p, int w, int n_units) {
    RawUnit r;
    r.a = make_int4(0, 0, 0, 0); r.b = r.a;
    if (w < n_units) {
        // volatile asm: an ordinary read-only load is free to be sunk to its first use (the next unit's start),
        // which is exactly the latency this fetch-ahead exists to hide
        const int4* rec = reinterpret_cast<const int4*>(p.items) + 2 * w;
        asm volatile("ld.global.nc.v4.s32 {%0,%1,%2,%3}, [%4];" : "=r"(r.a.x), "=r"(r.a.y), "=r"(r.a.z), "=r"(r.a.w) : "l"(rec));
        asm volatile("ld.global.nc.v4.s32 {%0,%1,%2,%3}, [%4];" : "=r"(r.b.x), "=r"(r.b.y), "=r"(r.b.z), "=r"(r.b.w) : "l"(rec + 1));
    }
    return r;
}
__device__ __forceinline__ void decode_unit(const PersistParams& p, const RawUnit& r, Unit& u) {
    u.b = r.a.x;
    const int mt = r.a.y, sub = r.b.z;
    u.q0 = r.a.z; u.Lq = r.a.w; u.k0 = r.b.x; u.Lk = r.b.y;
    u.pack = r.b.w;
    if (p.pair_heads) {
        u.r0 = u.r1 = mt * TC_BM;
        if (u.pack == 2) { u.head0 = 4 * sub; u.head1 = 4 * sub + 2; }
        else { u.head0 = 2 * sub; u.head1 = 2 * sub + 1; }
    } else {
        u.head0 = u.head1 = sub;
        u.r0 = mt * 2 * TC_BM; u.r1 = u.r0 + TC_BM;
    }
    u.kvh = (int)(((uint32_t)u.head0 * p.g_magic16) >> 16);       // head0 / G, exact for head0, G <= 128
    const int offset = u.Lk - u.Lq;
    const int rows = TC_BM / u.pack;                        // query rows per head in a tile
    int kv_end0 = min(u.Lk, u.r0 + rows + offset);
    if (u.r0 >= u.Lq || kv_end0 < 0) kv_end0 = 0;
    int kv_end1 = min(u.Lk, u.r1 + rows + offset);
    if (u.r1 >= u.Lq || kv_end1 < 0) kv_end1 = 0;
    u.n_t0 = (kv_end0 + TC_BN - 1) / TC_BN;
    u.n_t1 = (kv_end1 + TC_BN - 1) / TC_BN;
    u.n_max = max(u.n_t0, u.n_t1);
}

// A CTA stops at its first out-of-range round: if snake(k, c) >= n then n < (k+1)*g, so every index of
// every later round is out of range as well.
struct UnitWalk {
    int n_units, k; RawUnit raw;
    __device__ __forceinline__ void begin(const PersistParams& p) {
        n_units = __ldg(p.plan);
        k = 0;
        raw = fetch_unit(p, snake(0, (int)blockIdx.x, (int)gridDim.x), n_units);
    }
    // Decodes the unit of this round into u (false: no more work) and starts the fetch of the next round's record.
    __device__ __forceinline__ bool next(const PersistParams& p, Unit& u) {
        const int w = snake(k, (int)blockIdx.x, (int)gridDim.x);
        if (w >= n_units) return false;
        decode_unit(p, raw, u);
        ++k;
        raw = fetch_unit(p, snake(k, (int)blockIdx.x, (int)gridDim.x), n_units);
        return true;
    }
};

// ---- packed f32x2 arithmetic (sm_100: FFMA2 / FADD2) and 3-input max ---------------------------
__device__ __forceinline__ uint64_t pack2(float lo, float hi) {
    uint64_t r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void unpack2(uint64_t v, float& lo, float& hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ uint64_t fma2(uint64_t a, uint64_t b, uint64_t c) {
    uint64_t d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
__device__ __forceinline__ uint64_t add2(uint64_t a, uint64_t b) {
    uint64_t d;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ uint64_t mul2(uint64_t a, uint64_t b) {
    uint64_t d;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ float max3(float a, float b, float c) {
    float d;
    asm("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
    return d;
}
__device__ __forceinline__ uint32_t cvt_bf16x2(float lo, float hi) {
    uint32_t r;
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
    return r;
}

// Named barriers over a subset of the CTA (ids 1..15; 0 is __syncthreads).
__device__ __forceinline__ bool bar_red_or(uint32_t id, uint32_t nthreads, bool pred) {
    uint32_t r;
    asm volatile(
        "{\n\t.reg .pred p, q;\n\tsetp.ne.b32 q, %3, 0;\n\t"
        "bar.red.or.pred p, %1, %2, q;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(r) : "r"(id), "r"(nthreads), "r"((uint32_t)pred) : "memory");
    return r != 0;
}
__device__ __forceinline__ void bar_sync_n(uint32_t id, uint32_t nthreads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// 2^x for two values on the FMA/ALU pipes instead of MUFU.ex2 (which is co-critical with the tensor
// pipe: 16 ex2/clk/SM).  x = n + r with n = floor(x) taken from the low mantissa bits of
// (x + 1.5*2^23) rounded towards -inf, 2^r on [0,1) by a degree-3 polynomial (max relative error
// 8.8e-5, far below the bf16 rounding of P), n added straight into the exponent field.  x is clamped
// to >= -126 first, so masked (-inf) scores come out as ~2^-126 instead of 0; the epilogue treats a
// row sum below 1e-30 as "nothing visible" (a row with anything visible has a sum >= 2^-8).
__device__ __forceinline__ void exp2_poly2(float x0, float x1, float& p0, float& p1) {
    constexpr float MAGIC = 12582912.f;                             // 2^23 + 2^22
    const uint64_t x2 = pack2(fmaxf(x0, -126.f), fmaxf(x1, -126.f));
    uint64_t t2;
    asm("add.rm.f32x2 %0, %1, %2;" : "=l"(t2) : "l"(x2), "l"(pack2(MAGIC, MAGIC)));
    const uint64_t n2 = add2(t2, pack2(-MAGIC, -MAGIC));            // floor(x), exact
    const uint64_t r2 = fma2(n2, pack2(-1.f, -1.f), x2);            // x - floor(x) in [0,1)
    uint64_t q2 = fma2(r2, pack2(0.077119089663028717041015625f, 0.077119089663028717041015625f),
                       pack2(0.227564394474029541015625f, 0.227564394474029541015625f));
    q2 = fma2(q2, r2, pack2(0.695146143436431884765625f, 0.695146143436431884765625f));
    q2 = fma2(q2, r2, pack2(1.f, 1.f));
    float q0, q1, t0, t1;
    unpack2(q2, q0, q1);
    unpack2(t2, t0, t1);
    p0 = __uint_as_float(__float_as_uint(q0) + (__float_as_uint(t0) << 23));
    p1 = __uint_as_float(__float_as_uint(q1) + (__float_as_uint(t1) << 23));
}

constexpr int ps_threads(int wgs) { return wgs == 2 ? 640 : 384; }
#ifndef PS_W2_LOW
#define PS_W2_LOW 64
#define PS_W2_HIGH 104
#endif
#ifndef PS_W1_LOW
#define PS_W1_LOW 72
#define PS_W1_HIGH 216
#endif
#ifndef PS_P_HALF
#define PS_P_HALF 1      // publish the first 64 columns of P early (build knob; 0: one publication per tile)
#endif
#ifndef PS_CROSS_UNIT
#define PS_CROSS_UNIT 0
#endif
#define PS_STR2(x) #x
#define PS_STR(x) PS_STR2(x)

// EMU = how many of every 4 score pairs take the polynomial exp2 instead of MUFU.ex2 (0..4).
// WGS = softmax warpgroups per query tile.
//   1: one thread owns a whole 128-column score row; 384 threads: warp 0 producer, 1 MMA, 2 TMEM allocator,
//      4-11 softmax; setmaxnreg 80 / 208.
//   2: two warpgroups split every row in halves of 64 columns; 576 threads: warps 0-15 softmax (tile = wg >> 1,
//      half = wg & 1), 16 producer, 17 MMA + TMEM allocator; 112 registers each, no setmaxnreg.  The row maximum,
//      the packing and the TMEM traffic of a tile take half as long, which shortens the
//      S -> softmax -> P -> PV -> QK -> S chain that bounds this kernel.  Half h keeps its P over the first 32
//      columns of ITS OWN 64 score columns (PV K-steps 4..7 read their A operand at +64), so the halves never
//      touch each other's TMEM; they meet on a 64-thread named barrier that ORs the "row maximum outgrew the
//      running one" predicate and exchange maxima through shared memory only when it fires.
template <int EMU, bool TRACE = false, int WGS = 1>
__global__ void __launch_bounds__(ps_threads(WGS), 1) prefill_persist_kernel(
    const __grid_constant__ CUtensorMap map_q, const __grid_constant__ CUtensorMap map_k,
    const __grid_constant__ CUtensorMap map_v, const __grid_constant__ CUtensorMap map_q64, const PersistParams p) {
    using L = PsSmem;
    constexpr int S = TC_KV_STAGES;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    const uint32_t sQ = smem_base + L::Q_OFF, sK = smem_base + L::K_OFF, sV = smem_base + L::V_OFF;
    const uint32_t bars = smem_base + L::BAR_OFF;
    auto q_full = [&](int t) { return bars + 8 * t; };
    auto q_empty = [&](int t) { return bars + 8 * (2 + t); };
    auto k_full = [&](int s) { return bars + 8 * (4 + s); };
    auto k_empty = [&](int s) { return bars + 8 * (4 + S + s); };
    auto v_full = [&](int s) { return bars + 8 * (4 + 2 * S + s); };
    auto v_empty = [&](int s) { return bars + 8 * (4 + 3 * S + s); };
    auto s_full = [&](int t) { return bars + 8 * (4 + 4 * S + t); };
    auto p_full = [&](int t) { return bars + 8 * (6 + 4 * S + t); };
    auto o_final = [&](int t) { return bars + 8 * (8 + 4 * S + t); };
    auto p_half = [&](int t) { return bars + 8 * (10 + 4 * S + t); };
    volatile uint32_t* tmem_ptr_smem = reinterpret_cast<volatile uint32_t*>(smem_gen + L::TMEM_PTR_OFF);
    constexpr uint32_t TMEM_COLS = 512;

    constexpr int PROD_WARP = WGS == 2 ? 16 : 0, MMA_WARP = WGS == 2 ? 17 : 1, ALLOC_WARP = WGS == 2 ? 17 : 2;
    constexpr int INIT_WARP = WGS == 2 ? 16 : 1;
    const bool is_softmax = WGS == 2 ? warp < 16 : warp >= 4;

    if (warp == INIT_WARP && lane == 0) {
#pragma unroll
        for (int t = 0; t < 2; ++t) {
            mbar_init(q_full(t), 1); mbar_init(q_empty(t), 1); mbar_init(s_full(t), 1);
            mbar_init(p_full(t), 128); mbar_init(p_half(t), 128); mbar_init(o_final(t), 1);
        }
#pragma unroll
        for (int s = 0; s < S; ++s) {
            mbar_init(k_full(s), 1); mbar_init(k_empty(s), 1); mbar_init(v_full(s), 1); mbar_init(v_empty(s), 1);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == ALLOC_WARP) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_base + L::TMEM_PTR_OFF), "r"(TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    if (warp == PROD_WARP && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_q) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_k) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_v) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_q64) : "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr_smem;

    if (!is_softmax) {
        if constexpr (WGS == 1) asm volatile("setmaxnreg.dec.sync.aligned.u32 " PS_STR(PS_W1_LOW) ";");
        else asm volatile("setmaxnreg.dec.sync.aligned.u32 " PS_STR(PS_W2_LOW) ";");
        if (warp == PROD_WARP) {
            // ===================================== TMA producer =====================================
            if (lane == 0) {
                const bool paged = p.block_tables != nullptr;
                uint32_t kv_it = 0, q_use0 = 0, q_use1 = 0;
                int tiles0 = 0;                                  // TRACE: tile index (of query tile 0) this unit starts at
                auto pstamp = [&](int e) {
                    if (TRACE && blockIdx.x == 0 && tiles0 < p.trace_tiles) p.trace[(3 * p.trace_tiles + tiles0) * 8 + e] = clock64();
                };
                // Q tile t: 128 rows of one head, or (packed) 64 rows of head and of head + 1 stacked -- two 64-row boxes
                // per 64-dim half; 64 rows x 128 B = 8 swizzle atoms, so the second box keeps the first one's phase
                auto load_q = [&](int t, int head, int row, int pack) {
                    const uint32_t dst = sQ + t * TC_TILE_BYTES;
                    if (pack == 1) {
                        tma_load_3d(dst, &map_q, q_full(t), 0, head, row);
                        tma_load_3d(dst + TC_HALF_BYTES, &map_q, q_full(t), 64, head, row);
                    } else {
#pragma unroll
                        for (int i = 0; i < 2; ++i) {
                            tma_load_3d(dst + i * (TC_HALF_BYTES / 2), &map_q64, q_full(t), 0, head + i, row);
                            tma_load_3d(dst + TC_HALF_BYTES + i * (TC_HALF_BYTES / 2), &map_q64, q_full(t), 64, head + i, row);
                        }
                    }
                };
                Unit u;
                UnitWalk walk;
                walk.begin(p);
                while (walk.next(p, u)) {
                    const int kvh = u.kvh;
                    pstamp(6);                                   // producer holds this unit's descriptor
                    if (u.n_t0 > 0) {
                        mbar_wait(q_empty(0), (q_use0 & 1u) ^ 1u);
                        pstamp(4);                               // Q slot 0 free -> Q load issued
                        ++q_use0;
                        mbar_expect_tx(q_full(0), TC_TILE_BYTES);
                        load_q(0, u.head0, u.q0 + u.r0, u.pack);
                    }
                    if (u.n_t1 > 0) {
                        mbar_wait(q_empty(1), (q_use1 & 1u) ^ 1u);
                        ++q_use1;
                        mbar_expect_tx(q_full(1), TC_TILE_BYTES);
                        load_q(1, u.head1, u.q0 + u.r1, u.pack);
                    }
                    const int32_t* bt = paged ? p.block_tables + (int64_t)u.b * p.bt_row_stride : nullptr;
                    for (int j = 0; j < u.n_max; ++j, ++kv_it) {
                        const int n0 = j * TC_BN;
                        const int row = paged ? __ldg(bt + n0 / B200_PAGE) * B200_PAGE + (n0 & (B200_PAGE - 1)) : u.k0 + n0;
                        const uint32_t st = kv_it % S, par = (kv_it / S) & 1u;
                        mbar_wait(k_empty(st), par ^ 1u);
                        if (j == 0) pstamp(5);                   // first K tile of the unit issued
                        mbar_expect_tx(k_full(st), TC_TILE_BYTES);
                        tma_load_3d(sK + st * TC_TILE_BYTES, &map_k, k_full(st), 0, kvh, row);
                        tma_load_3d(sK + st * TC_TILE_BYTES + TC_HALF_BYTES, &map_k, k_full(st), 64, kvh, row);
                        mbar_wait(v_empty(st), par ^ 1u);
                        mbar_expect_tx(v_full(st), TC_TILE_BYTES);
                        tma_load_3d(sV + st * TC_TILE_BYTES, &map_v, v_full(st), 0, kvh, row);
                        tma_load_3d(sV + st * TC_TILE_BYTES + TC_HALF_BYTES, &map_v, v_full(st), 64, kvh, row);
                    }
                    tiles0 += u.n_t0;
                }
            }
        } else if (warp == MMA_WARP) {
            // ===================================== MMA issuer =======================================
            if (lane == 0) {
                constexpr uint32_t IDESC_QK = umma_idesc(0);           // B = K tile, K-major
                constexpr uint32_t IDESC_PV = umma_idesc(1);           // B = V tile, MN-major
                constexpr uint32_t DESC_HI = umma_desc_hi(1024);
                auto issue_qk = [&](int t, uint32_t st) {
                    const uint32_t k_lo = umma_desc_lo(sK + st * TC_TILE_BYTES, 0), q_lo = umma_desc_lo(sQ + t * TC_TILE_BYTES, 0);
                    const uint32_t d = tmem_base + t * 256;
#pragma unroll
                    for (int ks = 0; ks < 8; ++ks) {
                        const uint32_t off = ((ks >> 2) * TC_HALF_BYTES + (ks & 3) * 32) >> 4;
                        tc_mma_ss2(d, q_lo + off, k_lo + off, DESC_HI, IDESC_QK, ks > 0);
                    }
                };
                auto issue_pv = [&](int t, uint32_t st, bool first, int half) {     // K-steps 4*half .. 4*half+3
                    const uint32_t v_lo = umma_desc_lo(sV + st * TC_TILE_BYTES, TC_HALF_BYTES);
                    const uint32_t d = tmem_base + t * 256 + 128, a = tmem_base + t * 256;
#pragma unroll
                    for (int kk = 4 * half; kk < 4 * half + 4; ++kk)
                        tc_mma_ts2(d, a + (WGS == 2 && kk >= 4 ? 64 + (kk - 4) * 8 : kk * 8), v_lo + kk * (2048 >> 4), DESC_HI, IDESC_PV, (!first || kk > 0));
                };
                uint32_t kv_it = 0, q_use[2] = {0, 0}, p_cnt[2] = {0, 0};
                auto stamp = [&](int t, uint32_t n, int e) {
                    if (TRACE && blockIdx.x == 0 && (int)n < p.trace_tiles) p.trace[((2 + t) * p.trace_tiles + n) * 8 + e] = clock64();
                };
                // Cross-unit pipelining: in a unit's final K/V iteration, right after a query tile's last PV, the
                // FIRST QK of the next unit for that tile is issued (its Q tile and K tile 0 were staged by the
                // producer about one tile earlier), so the softmax warpgroup finds S ready when it returns from its
                // epilogue instead of waiting for the other tile to finish and for this thread to walk to the next
                // unit.  Only the final iteration may do this: earlier, the producer can still be held up by a
                // barrier that this thread commits later in the same iteration (deadlock).
                Unit u, un;
                UnitWalk walk;
                walk.begin(p);
                bool have = walk.next(p, u);
                bool pre[2] = {false, false};                   // first QK of unit `u`, tile t, already issued
                while (have) {
                    const bool have_n = walk.next(p, un);       // record fetched one unit ago
                    stamp(0, p_cnt[0], 7);                       // issuer holds this unit's descriptor
                    if (u.n_max > 0) {
                        const uint32_t st0 = kv_it % S;
                        bool waited_k = false;
#pragma unroll
                        for (int t = 0; t < 2; ++t)
                            if (u.nt(t) > 0 && !pre[t]) {
                                mbar_wait(q_full(t), q_use[t] & 1u); ++q_use[t];
                                if (!waited_k) { mbar_wait(k_full(st0), (kv_it / S) & 1u); waited_k = true; }
                                tc_fence_after();
                                issue_qk(t, st0);
                                tc_commit(s_full(t));
                                if (u.nt(t) == 1) tc_commit(q_empty(t));
                            }
                        stamp(1, p_cnt[0], 7);
                        tc_commit(k_empty(st0));                 // covers the QKs issued ahead during the previous unit too
                    }
                    pre[0] = pre[1] = false;
                    const uint32_t st0n = (kv_it + u.n_max) % S, par0n = ((kv_it + u.n_max) / S) & 1u;
                    for (int j = 0; j < u.n_max; ++j) {
                        const uint32_t it = kv_it + j, st = it % S, par = (it / S) & 1u;
                        const bool final_it = j + 1 == u.n_max;
                        stamp(0, p_cnt[0], 4);
                        mbar_wait(v_full(st), par);
                        stamp(0, p_cnt[0], 5);
                        if (!final_it) mbar_wait(k_full((it + 1) % S), ((it + 1) / S) & 1u);
                        stamp(0, p_cnt[0], 6);
                        tc_fence_after();
#pragma unroll
                        for (int t = 0; t < 2; ++t) {
                            if (j < u.nt(t)) {
                                stamp(t, p_cnt[t], 0);
                                if (PS_P_HALF || WGS == 2) {
                                    mbar_wait(p_half(t), p_cnt[t] & 1u);
                                    tc_fence_after();
                                    issue_pv(t, st, j == 0, 0);
                                }
                                stamp(t, p_cnt[t], 3);
                                mbar_wait(p_full(t), p_cnt[t] & 1u);
                                stamp(t, p_cnt[t], 1);
                                ++p_cnt[t];
                                tc_fence_after();
                                if (!(PS_P_HALF || WGS == 2)) issue_pv(t, st, j == 0, 0);
                                issue_pv(t, st, false, 1);
                                if (j + 1 < u.nt(t)) {
                                    issue_qk(t, (it + 1) % S);
                                    tc_commit(s_full(t));
                                    if (j + 2 == u.nt(t)) tc_commit(q_empty(t));   // that was the last Q read
                                } else {
                                    tc_commit(o_final(t));
                                }
                                stamp(t, p_cnt[t] - 1, 2);
                            }
                            if (PS_CROSS_UNIT && final_it && have_n && un.nt(t) > 0) {
                                mbar_wait(q_full(t), q_use[t] & 1u); ++q_use[t];
                                mbar_wait(k_full(st0n), par0n);
                                tc_fence_after();
                                issue_qk(t, st0n);
                                tc_commit(s_full(t));
                                if (un.nt(t) == 1) tc_commit(q_empty(t));
                                pre[t] = true;
                            }
                        }
                        tc_commit(v_empty(st));
                        if (!final_it) tc_commit(k_empty((it + 1) % S));
                    }
                    kv_it += u.n_max;
                    u = un; have = have_n;
                }
            }
        }
    } else if constexpr (WGS == 2) {
        asm volatile("setmaxnreg.inc.sync.aligned.u32 " PS_STR(PS_W2_HIGH) ";");
        // ============ softmax + correction + epilogue, two half-row warpgroups per query tile ============
        const int wg = warp >> 2;                                   // 0..3
        const int t = wg >> 1, h = wg & 1;                          // query tile, column half
        const int wq = warp & 3;                                    // TMEM lane quarter this warp may touch
        const int row_in_tile = wq * 32 + lane;
        // my 64 score columns; my P (bf16x2) overlays the first 32 of them; my 64 output columns are at +128
        const uint32_t tB = tmem_base + t * 256 + 64 * h + ((uint32_t)(wq * 32) << 16);
        const uint32_t pair_bar = 1 + t * 4 + wq;                   // the two warps that share these 32 rows
        const uint32_t my_p_bar = h == 0 ? p_half(t) : p_full(t);   // K-steps 0..3 of PV read half 0, 4..7 half 1
        float* xch = reinterpret_cast<float*>(smem_gen + L::XCH_OFF) + t * 256 + row_in_tile;
        uint32_t xpar = 0;                                          // exchange buffer parity (same in both halves)
        const uint64_t scale2 = pack2(p.scale_log2, p.scale_log2);
        uint32_t s_cnt = 0, o_cnt = 0, kv_it = 0;
        auto stamp = [&](uint32_t n, int e) {
            if (TRACE && blockIdx.x == 0 && h == 0 && row_in_tile == 0 && (int)n < p.trace_tiles) p.trace[(t * p.trace_tiles + n) * 8 + e] = clock64();
        };
        Unit u;
        UnitWalk walk;
        walk.begin(p);
        while (walk.next(p, u)) {
            const int nt = u.nt(t);
            const int r_begin = u.rbegin(t);
            const int row = r_begin + row_in_tile;                  // query index within the sequence
            const int offset = u.Lk - u.Lq;
            float m_used = -INFINITY;                               // running maximum, scaled (log2) domain
            uint64_t l2 = pack2(0.f, 0.f);                          // two partial sums of MY half of the row

            for (int j = 0; j < nt; ++j) {
                const int n0 = j * TC_BN;
                stamp(s_cnt, 0);
                mbar_wait(s_full(t), s_cnt & 1u);
                stamp(s_cnt, 1);
                ++s_cnt;
                tc_fence_after();
                uint32_t s[64];
                tmem_ld32(tB, *reinterpret_cast<uint32_t(*)[32]>(&s[0]));
                tmem_ld32(tB + 32, *reinterpret_cast<uint32_t(*)[32]>(&s[32]));
                tmem_wait_ld();
                stamp(s_cnt - 1, 2);
                if ((n0 + TC_BN > u.Lk) || (n0 + TC_BN - 1 > r_begin + offset)) {      // tail or diagonal tile
                    const int lim = min(u.Lk - 1, row + offset) - n0 - 64 * h;         // last visible column of my half
#pragma unroll
                    for (int i = 0; i < 64; ++i)
                        if (i > lim) s[i] = 0xff800000u;                                // -inf
                }
                float mxc[4];
#pragma unroll
                for (int c = 0; c < 4; ++c) mxc[c] = -INFINITY;
#pragma unroll
                for (int i = 0; i < 64; i += 8)
#pragma unroll
                    for (int c = 0; c < 4; ++c)
                        mxc[c] = max3(mxc[c], __uint_as_float(s[i + 2 * c]), __uint_as_float(s[i + 2 * c + 1]));
                const float mx = max3(mxc[0], mxc[1], fmaxf(mxc[2], mxc[3])) * p.scale_log2;
                // lazy rescale: keep exponentiating against m_used until some row's max outgrows it by 2^8.
                // The barrier tells both halves whether ANY of the 32 rows needs the slow path.
                const bool grow = mx > m_used + TC_RESCALE_THRESHOLD || (m_used == -INFINITY && mx > -INFINITY);
                if (bar_red_or(pair_bar, 64, grow)) {
                    float* slot = xch + xpar * 512;
                    xpar ^= 1u;
                    slot[h * 128] = mx;
                    bar_sync_n(pair_bar, 64);
                    const float m_new = fmaxf(m_used, fmaxf(mx, slot[(h ^ 1) * 128]));
                    const float alpha = fast_exp2(m_used - (m_new == -INFINITY ? 0.f : m_new));   // 0 when m_used = -inf
                    l2 = mul2(l2, pack2(alpha, alpha));
                    m_used = m_new;
                    if (j > 0) {                                    // PV(j-1) completed before s_full(j) fired
                        const uint64_t a2 = pack2(alpha, alpha);
#pragma unroll
                        for (int c = 0; c < 4; ++c) {               // 16 columns at a time: this rare path must not
                            uint32_t o[16];                         // raise the register peak of the common one
                            tmem_ld16(tB + 128 + c * 16, o);
                            tmem_wait_ld();
#pragma unroll
                            for (int i = 0; i < 16; i += 2) {
                                float a, b;
                                unpack2(mul2(pack2(__uint_as_float(o[i]), __uint_as_float(o[i + 1])), a2), a, b);
                                o[i] = __float_as_uint(a); o[i + 1] = __float_as_uint(b);
                            }
                            tmem_st16(tB + 128 + c * 16, o);
                        }
                        // PV(j) K-steps 0..3 start on half 0's arrival alone and accumulate into ALL of O:
                        // neither half may arrive before both have finished rescaling.
                        tmem_wait_st();
                        tc_fence_before();
                        bar_sync_n(pair_bar, 64);
                        tc_fence_after();
                    }
                }
                stamp(s_cnt - 1, 3);
                const float neg_m = m_used == -INFINITY ? 0.f : -m_used;
                const uint64_t negm2 = pack2(neg_m, neg_m);
                if (n0 + TC_BN > u.Lk) {
                    // Rows of the V tile past the context may hold NaN garbage (torch.empty cache):
                    // 0 * NaN would poison O, so zero them in shared memory before PV reads them.
                    const uint32_t it = kv_it + j, st = it % S;
                    mbar_wait(v_full(st), (it / S) & 1u);
                    uint8_t* vt = smem_gen + L::V_OFF + st * TC_TILE_BYTES;
                    const int first = u.Lk - n0;                    // first invalid row of this tile
                    for (int idx = row_in_tile; idx < (TC_BN - first) * 16; idx += 128) {
                        const int r = first + (idx >> 4), c16 = idx & 15;
                        *reinterpret_cast<int4*>(vt + (c16 >> 3) * TC_HALF_BYTES + r * 128 + (c16 & 7) * 16) = make_int4(0, 0, 0, 0);
                    }
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                }
                {
                    constexpr int EXP_LAG = 6;
                    uint64_t acc1 = pack2(0.f, 0.f);
                    uint32_t pk[8];
#pragma unroll
                    for (int i = 0; i < 32 + EXP_LAG; ++i) {
                        if (i < 32) {
                            float x0, x1, p0, p1;
                            unpack2(fma2(pack2(__uint_as_float(s[2 * i]), __uint_as_float(s[2 * i + 1])), scale2, negm2), x0, x1);
                            if ((i & 3) < EMU) exp2_poly2(x0, x1, p0, p1);
                            else { p0 = fast_exp2(x0); p1 = fast_exp2(x1); }
                            s[2 * i] = __float_as_uint(p0); s[2 * i + 1] = __float_as_uint(p1);
                        }
                        if (i >= EXP_LAG) {
                            const int kk = i - EXP_LAG;
                            const float p0 = __uint_as_float(s[2 * kk]), p1 = __uint_as_float(s[2 * kk + 1]);
                            if (kk & 1) acc1 = add2(acc1, pack2(p0, p1));
                            else l2 = add2(l2, pack2(p0, p1));
                            pk[kk & 7] = cvt_bf16x2(p0, p1);
                            if ((kk & 7) == 7) tmem_st8(tB + (kk >> 3) * 8, pk);
                        }
                    }
                    l2 = add2(l2, acc1);
                }
                stamp(s_cnt - 1, 4);
                tmem_wait_st();
                tc_fence_before();
                mbar_arrive(my_p_bar);
                stamp(s_cnt - 1, 5);
            }
            kv_it += u.n_max;

            // ---- epilogue: O / l -> bf16 -> global (rows beyond the sequence are dropped) --------
            if (nt > 0) {
                float la, lb;
                unpack2(l2, la, lb);
                const float l_mine = la + lb;
                float* slot = xch + xpar * 512;
                xpar ^= 1u;
                slot[h * 128] = l_mine;
                bar_sync_n(pair_bar, 64);
                const float l_other = slot[(h ^ 1) * 128];
                const float l_sum = h == 0 ? l_mine + l_other : l_other + l_mine;    // half 0 + half 1 in both
                mbar_wait(o_final(t), o_cnt & 1u);
                ++o_cnt;
                tc_fence_after();
                const float inv = l_sum > 1e-30f ? 1.f / l_sum : 0.f;  // see exp2_poly2 on masked scores
                __nv_bfloat16* orow = p.out + (int64_t)(u.q0 + row) * p.out_token_stride + (int64_t)u.head(t) * B200_HD + 64 * h;
#pragma unroll
                for (int c = 0; c < 2; ++c) {
                    uint32_t o[32];
                    tmem_ld32(tB + 128 + c * 32, o);
                    tmem_wait_ld();
                    if (row < u.Lq) {
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            int4 v;
                            v.x = (int)cvt_bf16x2(__uint_as_float(o[8 * i + 0]) * inv, __uint_as_float(o[8 * i + 1]) * inv);
                            v.y = (int)cvt_bf16x2(__uint_as_float(o[8 * i + 2]) * inv, __uint_as_float(o[8 * i + 3]) * inv);
                            v.z = (int)cvt_bf16x2(__uint_as_float(o[8 * i + 4]) * inv, __uint_as_float(o[8 * i + 5]) * inv);
                            v.w = (int)cvt_bf16x2(__uint_as_float(o[8 * i + 6]) * inv, __uint_as_float(o[8 * i + 7]) * inv);
                            *reinterpret_cast<int4*>(orow + c * 32 + i * 8) = v;
                        }
                    }
                }
                // O(t) is overwritten by the next unit's first PV, which waits on half 0's next arrival; half 0
                // cannot get there before BOTH halves pass the pair barrier of the next unit's first tile,
                // i.e. after both epilogues have read their columns.
                tc_fence_before();
            }
        }
    } else {
        asm volatile("setmaxnreg.inc.sync.aligned.u32 " PS_STR(PS_W1_HIGH) ";");
        // ======================== softmax + correction + epilogue warpgroups =====================
        const int t = (warp - 4) >> 2;                              // query tile of this warpgroup
        const int wq = warp & 3;                                    // TMEM lane quarter this warp may touch
        const int row_in_tile = wq * 32 + lane;
        const uint32_t lane_addr = (uint32_t)(wq * 32) << 16;
        const uint32_t tS = tmem_base + t * 256 + lane_addr;
        const uint32_t tO = tS + 128;
        const uint64_t scale2 = pack2(p.scale_log2, p.scale_log2);
        uint32_t s_cnt = 0, o_cnt = 0, kv_it = 0;
        auto stamp = [&](uint32_t n, int e) {
            if (TRACE && blockIdx.x == 0 && row_in_tile == 0 && (int)n < p.trace_tiles) p.trace[(t * p.trace_tiles + n) * 8 + e] = clock64();
        };
        Unit u;
        UnitWalk walk;
        walk.begin(p);
        while (walk.next(p, u)) {
            if (TRACE && blockIdx.x == 0 && row_in_tile == 0 && s_cnt > 0 && (int)(s_cnt - 1) < p.trace_tiles)   // unit decoded
                p.trace[((4 + t) * p.trace_tiles + (s_cnt - 1)) * 8 + 7] = clock64();
            const int nt = u.nt(t);
            const int r_begin = u.rbegin(t);
            // packed tiles: rows 0..63 belong to head(t), rows 64..127 to head(t) + 1, both at query rows r_begin + 0..63
            const int row = r_begin + (u.pack == 2 ? (row_in_tile & 63) : row_in_tile);   // query index within the sequence
            const int offset = u.Lk - u.Lq;
            float m_used = -INFINITY;                               // running maximum, scaled (log2) domain
            uint64_t l2 = pack2(0.f, 0.f);                          // two partial row sums

            for (int j = 0; j < nt; ++j) {
                const int n0 = j * TC_BN;
                stamp(s_cnt, 0);
                mbar_wait(s_full(t), s_cnt & 1u);
                stamp(s_cnt, 1);
                ++s_cnt;
                tc_fence_after();
                uint32_t s[128];
#pragma unroll
                for (int c = 0; c < 4; ++c) tmem_ld32(tS + c * 32, *reinterpret_cast<uint32_t(*)[32]>(&s[c * 32]));
                tmem_wait_ld();
                stamp(s_cnt - 1, 2);
                if constexpr (TRACE) {                              // stamp 6: all four score loads have really landed
                    uint32_t dep = s[0] ^ s[31] ^ s[32] ^ s[63] ^ s[64] ^ s[95] ^ s[96] ^ s[127];
                    asm volatile("" ::"r"(dep) : "memory");
                    stamp(s_cnt - 1, 6);
                }
                if ((n0 + TC_BN > u.Lk) || (n0 + TC_BN - 1 > r_begin + offset)) {      // tail or diagonal tile
                    const int lim = min(u.Lk - 1, row + offset) - n0;                  // last visible column of this row
#pragma unroll
                    for (int i = 0; i < 128; ++i)
                        if (i > lim) s[i] = 0xff800000u;                                // -inf
                }
                // 8 independent max chains (FMNMX3 latency would serialise 2 chains of 32)
                float mxc[8];
#pragma unroll
                for (int c = 0; c < 8; ++c) mxc[c] = -INFINITY;
#pragma unroll
                for (int i = 0; i < 128; i += 16)
#pragma unroll
                    for (int c = 0; c < 8; ++c)
                        mxc[c] = max3(mxc[c], __uint_as_float(s[i + 2 * c]), __uint_as_float(s[i + 2 * c + 1]));
                const float mx0 = max3(mxc[0], mxc[1], mxc[2]), mx1 = max3(mxc[3], mxc[4], mxc[5]);
                const float mx = max3(mx0, mx1, fmaxf(mxc[6], mxc[7])) * p.scale_log2;
                if constexpr (TRACE) {                              // stamp 7: row maximum known
                    asm volatile("" ::"f"(mx) : "memory");
                    stamp(s_cnt - 1, 7);
                }
                // lazy rescale: keep exponentiating against m_used until some row's max outgrows it by 2^8
                const bool grow = mx > m_used + TC_RESCALE_THRESHOLD || (m_used == -INFINITY && mx > -INFINITY);
                if (__any_sync(0xffffffffu, grow)) {
                    const float m_new = fmaxf(m_used, mx);
                    const float alpha = fast_exp2(m_used - (m_new == -INFINITY ? 0.f : m_new));   // 0 when m_used = -inf
                    l2 = mul2(l2, pack2(alpha, alpha));
                    m_used = m_new;
                    if (j > 0) {                                    // PV(j-1) completed before s_full(j) fired
                        const uint64_t a2 = pack2(alpha, alpha);
#pragma unroll
                        for (int c = 0; c < 4; ++c) {
                            uint32_t o[32];
                            tmem_ld32(tO + c * 32, o);
                            tmem_wait_ld();
#pragma unroll
                            for (int i = 0; i < 32; i += 2) {
                                float a, b;
                                unpack2(mul2(pack2(__uint_as_float(o[i]), __uint_as_float(o[i + 1])), a2), a, b);
                                o[i] = __float_as_uint(a); o[i + 1] = __float_as_uint(b);
                            }
                            tmem_st32(tO + c * 32, o);
                        }
                    }
                }
                stamp(s_cnt - 1, 3);
                const float neg_m = m_used == -INFINITY ? 0.f : -m_used;
                const uint64_t negm2 = pack2(neg_m, neg_m);
                if (n0 + TC_BN > u.Lk) {
                    // Rows of the V tile past the context may hold NaN garbage (torch.empty cache):
                    // 0 * NaN would poison O, so zero them in shared memory before PV reads them.
                    const uint32_t it = kv_it + j, st = it % S;
                    mbar_wait(v_full(st), (it / S) & 1u);
                    uint8_t* vt = smem_gen + L::V_OFF + st * TC_TILE_BYTES;
                    const int first = u.Lk - n0;                    // first invalid row of this tile
                    for (int idx = row_in_tile; idx < (TC_BN - first) * 16; idx += 128) {
                        const int r = first + (idx >> 4), c16 = idx & 15;
                        *reinterpret_cast<int4*>(vt + (c16 >> 3) * TC_HALF_BYTES + r * 128 + (c16 & 7) * 16) = make_int4(0, 0, 0, 0);
                    }
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                }
                // Software-pipelined exponentials: the consumer (row sum, bf16 pack, TMEM store) of pair k
                // runs EXP_LAG pairs behind its MUFU.ex2 issue, so the in-order warp never stalls on MUFU
                // latency; two row-sum accumulators break the FADD2 chain.  The first 64 columns of P (KV
                // rows 0..63 = the first four K-steps of PV) are published early through p_half so the
                // tensor pipe starts PV while the second half is still being exponentiated.
                {
                    constexpr int EXP_LAG = 6, HALF_AT = 32 + EXP_LAG + 6;
                    uint64_t acc1 = pack2(0.f, 0.f);
                    uint32_t pk[16];
#pragma unroll
                    for (int i = 0; i < 64 + EXP_LAG; ++i) {
                        if (i < 64) {
                            float x0, x1, p0, p1;
                            unpack2(fma2(pack2(__uint_as_float(s[2 * i]), __uint_as_float(s[2 * i + 1])), scale2, negm2), x0, x1);
                            if ((i & 3) < EMU) exp2_poly2(x0, x1, p0, p1);
                            else { p0 = fast_exp2(x0); p1 = fast_exp2(x1); }
                            s[2 * i] = __float_as_uint(p0); s[2 * i + 1] = __float_as_uint(p1);
                        }
                        if (i >= EXP_LAG) {
                            const int k = i - EXP_LAG;
                            const float p0 = __uint_as_float(s[2 * k]), p1 = __uint_as_float(s[2 * k + 1]);
                            if (k & 1) acc1 = add2(acc1, pack2(p0, p1));
                            else l2 = add2(l2, pack2(p0, p1));
                            pk[k & 15] = cvt_bf16x2(p0, p1);
                            if ((k & 15) == 15) tmem_st16(tS + (k >> 4) * 16, pk);
                        }
                        if (PS_P_HALF && i == HALF_AT) {                         // stores of columns 0..63 were issued 6 pairs ago
                            tmem_wait_st();
                            tc_fence_before();
                            mbar_arrive(p_half(t));
                        }
                    }
                    l2 = add2(l2, acc1);
                }
                stamp(s_cnt - 1, 4);
                tmem_wait_st();
                tc_fence_before();
                mbar_arrive(p_full(t));
                stamp(s_cnt - 1, 5);
            }
            kv_it += u.n_max;

            // ---- epilogue: O / l -> bf16 -> shared staging -> global, whole 128-byte lines ----------------
            // (a thread owns a ROW of O; storing it straight to global memory is 16 bytes per row per
            // instruction, 32 partial sectors per warp store -- measured ~6,000 cycles per tile, during which
            // this warpgroup cannot start the next unit.)  Two passes of 64 columns through a 16 KB buffer.
            auto estamp = [&](int e) {                              // TRACE rows 4/5: epilogue of the unit's last tile
                if (TRACE && blockIdx.x == 0 && row_in_tile == 0 && s_cnt > 0 && (int)(s_cnt - 1) < p.trace_tiles)
                    p.trace[((4 + t) * p.trace_tiles + (s_cnt - 1)) * 8 + e] = clock64();
            };
            if (nt > 0) {
                estamp(0);
                mbar_wait(o_final(t), o_cnt & 1u);
                estamp(1);
                ++o_cnt;
                tc_fence_after();
                float la, lb;
                unpack2(l2, la, lb);
                const float l_sum = la + lb;
                const float inv = l_sum > 1e-30f ? 1.f / l_sum : 0.f;  // see exp2_poly2 on masked scores
                const uint64_t inv2 = pack2(inv, inv);
                uint8_t* stage = smem_gen + L::STAGE_OFF + t * L::STAGE_BYTES;
                const int tid_wg = (warp & 3) * 32 + lane;          // == row_in_tile
                const int rmask = u.pack == 2 ? 63 : 127, rshift = u.pack == 2 ? 6 : 7;   // tile row -> (head, query row)
                const int rows_valid = min(rmask + 1, u.Lq - r_begin);   // query rows of this tile inside the sequence
                __nv_bfloat16* obase = p.out + (int64_t)(u.q0 + r_begin) * p.out_token_stride + (int64_t)u.head(t) * B200_HD;
#pragma unroll
                for (int pass = 0; pass < 2; ++pass) {
#pragma unroll
                    for (int c = 0; c < 2; ++c) {
                        uint32_t o[32];
                        tmem_ld32(tO + pass * 64 + c * 32, o);
                        tmem_wait_ld();
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            float f[8];
#pragma unroll
                            for (int q = 0; q < 8; q += 2)
                                unpack2(mul2(pack2(__uint_as_float(o[8 * i + q]), __uint_as_float(o[8 * i + q + 1])), inv2), f[q], f[q + 1]);
                            int4 v;
                            v.x = (int)cvt_bf16x2(f[0], f[1]); v.y = (int)cvt_bf16x2(f[2], f[3]);
                            v.z = (int)cvt_bf16x2(f[4], f[5]); v.w = (int)cvt_bf16x2(f[6], f[7]);
                            const int chunk = c * 4 + i;            // 16-byte chunk of this row's 128 staged bytes
                            *reinterpret_cast<int4*>(stage + row_in_tile * 128 + ((chunk ^ (row_in_tile & 7)) << 4)) = v;
                        }
                    }
                    estamp(2 + 2 * pass);
                    bar_sync_n(1 + t, 128);                          // the half tile is staged
#pragma unroll
                    for (int step = 0; step < 8; ++step) {
                        const int r = step * 16 + (tid_wg >> 3), chunk = tid_wg & 7;
                        const int qi = r & rmask, hi = r >> rshift;
                        if (qi < rows_valid) {
                            const int4 v = *reinterpret_cast<const int4*>(stage + r * 128 + ((chunk ^ (r & 7)) << 4));
                            *reinterpret_cast<int4*>(obase + (int64_t)qi * p.out_token_stride + hi * B200_HD + pass * 64 + chunk * 8) = v;
                        }
                    }
                    estamp(3 + 2 * pass);
                    if (pass == 0) bar_sync_n(1 + t, 128);           // staging is free again
                }
                // The next unit's epilogue writes the staging buffer only after ITS o_final, i.e. after every
                // thread of this warpgroup has arrived on p_full at least once more => the reads above are done.
                // O(t) is overwritten by the next unit's first PV, which waits on this warpgroup's
                // next p_full arrival (program order) => the reads above are complete by then.
                tc_fence_before();
                estamp(6);
            }
        }
    }

    // ---- teardown --------------------------------------------------------------------------------
    tc_fence_before();
    __syncthreads();
    if (warp == ALLOC_WARP) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
    }
}

bool cached_map(CUtensorMap* out, const void* base, uint64_t rows, uint64_t heads, uint64_t row_stride_elems, uint32_t box_rows) {
    static std::unordered_map<MapKey, CUtensorMap, MapKeyHash> cache;
    static std::mutex mu;
    static size_t cap = 0;
    std::lock_guard<std::mutex> lock(mu);
    if (cap == 0) {
        const char* env = getenv("B200ATTN_MAP_CACHE_CAP");     // tests shrink it to force evictions inside one launch
        const long v = env ? atol(env) : 0;
        cap = v > 0 ? (size_t)v : 4096;
    }
    const MapKey key{base, rows, heads, row_stride_elems, box_rows, current_device()};
    auto it = cache.find(key);
    if (it != cache.end()) { *out = it->second; return true; }
    // The q/k/v keys carry total_q (the TMA bound that keeps box loads inside the allocation), so a serving process
    // keeps minting new keys; when the table is full it is emptied.  Safe at any point: callers hold copies.
    if (cache.size() >= cap) cache.clear();
    CUtensorMap m;
    if (!make_map(&m, base, rows, heads, row_stride_elems, box_rows)) return false;
    cache.emplace(key, m);
    *out = m;
    return true;
}

static long long* g_trace = nullptr;
static int g_trace_tiles = 0;

// softmax warpgroups per query tile (1 or 2): B200ATTN_PREFILL_WGS, read once
static int prefill_wgs() {
    static int wgs = -1;
    if (wgs < 0) {
        const char* env = getenv("B200ATTN_PREFILL_WGS");
        wgs = (env ? atoi(env) : PREFILL_WGS_DEFAULT) == 2 ? 2 : 1;
    }
    return wgs;
}

// Decided where the step plan is built (b200_attn_plan): query tiles with <= 64 rows inside their sequence carry two q
// heads per 128-row tile when the unit's four heads share one kv head (group divisible by 4) and the kernel variant
// maps rows to (head, query) pairs (WGS = 1).  B200ATTN_PREFILL_NOPACK=1 turns it off (A/B).
bool prefill_pack_ok(int32_t Hq, int32_t Hkv) {
    static int nopack = -1;
    if (nopack < 0) nopack = getenv("B200ATTN_PREFILL_NOPACK") ? 1 : 0;
    const int G = Hq / Hkv;
    return PS_PACK && !nopack && prefill_wgs() == 1 && G % 4 == 0;
}

int launch_attn_prefill_persist(const PrefillArgs& a, const int32_t* plan, int32_t max_items, cudaStream_t stream) {
    if (a.batch == 0 || a.total_q == 0 || a.max_q == 0) return 0;
    if (get_encode_fn() == nullptr) return -6;
    using KernelFn = void (*)(const CUtensorMap, const CUtensorMap, const CUtensorMap, const CUtensorMap, const PersistParams);
    static const KernelFn kernels[2][5] = {
        {prefill_persist_kernel<0, false, 1>, prefill_persist_kernel<1, false, 1>, prefill_persist_kernel<2, false, 1>,
         prefill_persist_kernel<3, false, 1>, prefill_persist_kernel<4, false, 1>},
        {prefill_persist_kernel<0, false, 2>, prefill_persist_kernel<1, false, 2>, prefill_persist_kernel<2, false, 2>,
         prefill_persist_kernel<3, false, 2>, prefill_persist_kernel<4, false, 2>}};
    static int emu = -1;
    const int wgs = prefill_wgs();
    static PerDeviceOnce once;                                   // the shared-memory opt-in is per device
    if (once.need()) {
        for (int w = 0; w < 2; ++w)
            for (int i = 0; i < 5; ++i) {
                cudaError_t e = cudaFuncSetAttribute(kernels[w][i], cudaFuncAttributeMaxDynamicSharedMemorySize, PsSmem::ALLOC);
                if (e != cudaSuccess) return (int)e;
            }
        once.done();
    }
    if (emu < 0) {
        const char* env = getenv("B200ATTN_PREFILL_EMU");        // A/B knob; default chosen by measurement
        int v = env ? atoi(env) : PREFILL_EMU_DEFAULT;
        emu = v < 0 ? 0 : (v > 4 ? 4 : v);
    }
    const bool paged = a.block_tables != nullptr;
    CUtensorMap mq, mq64, mk, mv;                                // copies: a later look-up may empty the cache
    bool ok = cached_map(&mq, a.q, (uint64_t)a.total_q, (uint64_t)a.Hq, (uint64_t)a.q_token_stride) &&
              cached_map(&mq64, a.q, (uint64_t)a.total_q, (uint64_t)a.Hq, (uint64_t)a.q_token_stride, 64);
    if (paged) {
        const uint64_t rows = (uint64_t)a.num_cache_blocks * B200_PAGE;
        ok = ok && cached_map(&mk, a.k, rows, (uint64_t)a.Hkv, (uint64_t)a.Hkv * B200_HD) &&
             cached_map(&mv, a.v, rows, (uint64_t)a.Hkv, (uint64_t)a.Hkv * B200_HD);
    } else {
        ok = ok && cached_map(&mk, a.k, (uint64_t)a.total_q, (uint64_t)a.Hkv, (uint64_t)a.k_token_stride) &&
             cached_map(&mv, a.v, (uint64_t)a.total_q, (uint64_t)a.Hkv, (uint64_t)a.v_token_stride);
    }
    if (!ok) return -6;
    PersistParams p;
    p.cu_q = a.cu_q; p.cu_k = a.cu_k; p.block_tables = a.block_tables; p.bt_row_stride = a.bt_row_stride;
    p.out = (__nv_bfloat16*)a.out; p.out_token_stride = a.out_token_stride;
    p.Hq = a.Hq; p.G = a.Hq / a.Hkv; p.scale_log2 = a.scale * 1.4426950408889634f;
    p.pair_heads = plan_tile_rows(a.Hq, a.Hkv) == 128 ? 1 : 0;
    p.g_magic16 = (uint32_t)((65536 + p.G - 1) / p.G);
    p.plan = plan; p.items = plan + plan_units_offset(a.batch, max_items);      // the unit table (launch.h)
    int64_t units = (int64_t)max_items * (p.pair_heads ? a.Hq / 2 : a.Hq);
    int grid = sm_count();
    if (units < grid) grid = (int)units;
    if (grid < 1) grid = 1;
    p.trace = g_trace; p.trace_tiles = g_trace_tiles;
    if (g_trace != nullptr) {
        static PerDeviceOnce trace_once;
        if (trace_once.need()) {
            cudaFuncSetAttribute(prefill_persist_kernel<PREFILL_EMU_DEFAULT, true, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, PsSmem::ALLOC);
            cudaFuncSetAttribute(prefill_persist_kernel<PREFILL_EMU_DEFAULT, true, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, PsSmem::ALLOC);
            trace_once.done();
        }
        if (wgs == 2) prefill_persist_kernel<PREFILL_EMU_DEFAULT, true, 2><<<grid, ps_threads(2), PsSmem::ALLOC, stream>>>(mq, mk, mv, mq64, p);
        else prefill_persist_kernel<PREFILL_EMU_DEFAULT, true, 1><<<grid, ps_threads(1), PsSmem::ALLOC, stream>>>(mq, mk, mv, mq64, p);
    } else {
        kernels[wgs - 1][emu]<<<grid, ps_threads(wgs), PsSmem::ALLOC, stream>>>(mq, mk, mv, mq64, p);
    }
    return (int)cudaGetLastError();
}

}  // namespace b200

// Debug only: when `dev_buf` (int64[4][max_tiles][8], zeroed by the caller) is non-null, the next
// (int64[6][max_tiles][8] since the epilogue rows were added) persistent prefill launches run the TRACE build and CTA 0 stamps clock64 at the phase boundaries of
// its first `max_tiles` tiles (rows 0/1: softmax warpgroups, rows 2/3: the MMA issuer's view of them).
extern "C" int b200_debug_prefill_trace(void* dev_buf, int max_tiles) {
    b200::g_trace = (long long*)dev_buf;
    b200::g_trace_tiles = dev_buf ? max_tiles : 0;
    return 0;
}
