// HBM-bound byte movers of the path: KV scatter (piece 1), fused per-head RMSNorm+RoPE
// (piece 4a), hidden-size (add+)RMSNorm (4a'), SiluAndMul (4b).  All use 16-byte coalesced
// accesses; none is reshaped into a GEMM.
#include "common.cuh"
#include "launch.h"

namespace b200 {

// ------------------------------------------------------------------------------------------
// (1) KV scatter.  One thread moves SC_UNROLL 16-byte chunks of K and the matching chunks of V per
//     step: all slot lookups, then all 2 x SC_UNROLL loads, then the stores -- 128 bytes in flight per
//     thread, so a short launch (T x 4 x Hkv x 256 B: 134 MB at T = 16384, 8 kv heads) is not bound by
//     the slot -> load -> store latency chain of one chunk.  Consecutive threads walk consecutive
//     chunks of one token row => every warp issues full 512-byte coalesced loads and stores.
//     POW2 = chunks per row is a power of two (1/2/4/8 kv heads); else a 32-bit divide (5 heads).
//     Reference: attention.py:74-90.
// ------------------------------------------------------------------------------------------
constexpr int SC_UNROLL = 4;

template <bool POW2>
__global__ void __launch_bounds__(256) kv_scatter_kernel(
    const uint8_t* __restrict__ key, const uint8_t* __restrict__ value,
    int64_t key_row_bytes, int64_t value_row_bytes,
    uint8_t* __restrict__ k_cache, uint8_t* __restrict__ v_cache,
    const int32_t* __restrict__ slot_mapping, uint32_t total_chunks, uint32_t chunks_per_row, uint32_t lg) {
    const uint32_t step = gridDim.x * (256u * SC_UNROLL);
    for (uint32_t base = blockIdx.x * (256u * SC_UNROLL) + threadIdx.x; base < total_chunks; base += step) {
        int4 kv[SC_UNROLL], vv[SC_UNROLL];
        int64_t dst[SC_UNROLL];
        bool ok[SC_UNROLL];
#pragma unroll
        for (int u = 0; u < SC_UNROLL; ++u) {
            const uint32_t i = base + u * 256u;
            ok[u] = false;
            if (i < total_chunks) {
                const uint32_t tok = POW2 ? i >> lg : i / chunks_per_row;
                const uint32_t c = POW2 ? i & (chunks_per_row - 1) : i - tok * chunks_per_row;
                const int32_t slot = __ldg(slot_mapping + tok);
                if (slot >= 0) {                                              // attention.py:79-80
                    ok[u] = true;
                    kv[u] = ld_nc_v4(key + (int64_t)tok * key_row_bytes + c * 16);
                    vv[u] = ld_nc_v4(value + (int64_t)tok * value_row_bytes + c * 16);
                    dst[u] = ((int64_t)slot * chunks_per_row + c) * 16;       // 64-bit: SURVEY 7.3(9)
                }
            }
        }
#pragma unroll
        for (int u = 0; u < SC_UNROLL; ++u)
            if (ok[u]) {
                st_v4(k_cache + dst[u], kv[u]);
                st_v4(v_cache + dst[u], vv[u]);
            }
    }
}

// ------------------------------------------------------------------------------------------
// (4a) q_norm/k_norm + RoPE.  Eight lanes per (token, head) row of 128 bf16, four rows per warp.
//      Lane l owns elements [8l, 8l+8) of the low half and [64+8l, 64+8l+8) of the high half,
//      i.e. both members of eight NeoX rotation pairs => two 16-byte loads + two 16-byte stores
//      per lane, no cross-lane traffic for the rotation, a 3-step butterfly for the sum of squares.
//      A warp walks the heads of one (token, head segment) unit four at a time with the loads of the next two
//      head-quads in flight (2 KB per warp; 4 CTAs per SM => 64 KB per SM, what ~800 ns of HBM latency needs);
//      the grid is exactly one wave of resident CTAs (a 4/3-wave grid cost 8 %).  Narrowing conversions are
//      packed F2FP (ALU pipe), never F2F (quarter-rate XU pipe).
//      Roundings follow layernorm.py:85-93 and rotary_embedding.py:60-65 exactly:
//      bf16(x*rstd), bf16(.*w), fp32 rotate (un-fused mul/sub), bf16.
// ------------------------------------------------------------------------------------------
// STORE = true additionally performs the KV scatter of piece (1) in the same launch:
// rows are then (token, head) over Hq + 2*Hkv heads; a k row is written to k_out (if given) AND to
// k_cache[slot], a v row is copied to v_cache[slot]; slot -1 skips the cache write (attention.py:79).
struct KvStoreArgs {
    const __nv_bfloat16* v; int64_t v_token_stride;
    __nv_bfloat16* k_cache; __nv_bfloat16* v_cache; const int32_t* slot_mapping;
};

// bf16 pair <-> two floats without F2F (which runs on the quarter-rate XU pipe): widening is a shift / mask,
// narrowing is ONE packed F2FP (round-to-nearest-even, the same rounding as __float2bfloat16_rn).
__device__ __forceinline__ float bf_lo(uint32_t u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float bf_hi(uint32_t u) { return __uint_as_float(u & 0xffff0000u); }

// Work decomposition (round 2): a warp takes a (token, head segment) unit and walks the segment's heads four at
// a time (one head per 8-lane group).  Everything that depends on the token only -- position, the 16 cos/sin
// values a lane needs, the cache slot, the three row base pointers -- is fetched ONCE per unit instead of once
// per row (the old row-major walk paid a division by H, a position load and four 16-byte cos/sin loads per
// (token, head) row: about as many instructions as the arithmetic itself), and the row type (q / k / v) is two
// compares on the head index.  Within a unit the loads of the next four heads are in flight while the current
// four are normalised, rotated and stored.  `nseg` segments per token keep the machine full at decode sizes.
#ifndef NR_BLOCKS
#define NR_BLOCKS 4
#endif
struct RowData { int4 lo, hi; };                   // elements [8s, 8s+8) and [64+8s, 64+8s+8) of a row

template <bool STORE>
__global__ void __launch_bounds__(256, NR_BLOCKS) qk_norm_rope_kernel(
    const __nv_bfloat16* __restrict__ q, int64_t q_token_stride,
    const __nv_bfloat16* __restrict__ k, int64_t k_token_stride,
    const int64_t* __restrict__ positions, const float* __restrict__ cos_sin,
    const __nv_bfloat16* __restrict__ q_weight, const __nv_bfloat16* __restrict__ k_weight, float eps,
    __nv_bfloat16* __restrict__ q_out, __nv_bfloat16* __restrict__ k_out,
    uint32_t T, uint32_t Hq, uint32_t Hkv, uint32_t nseg, uint32_t hps, const KvStoreArgs st) {
    const int sub = threadIdx.x & 7, grp = (threadIdx.x >> 3) & 3;
    const uint32_t H = STORE ? Hq + 2 * Hkv : Hq + Hkv;
    const bool has_w = q_weight != nullptr;                    // both weights or neither (checked at the ABI)
    const uint32_t warp_global = (blockIdx.x * 256u + threadIdx.x) >> 5, n_warps = gridDim.x * 8u;
    const uint32_t units = T * nseg;

    for (uint32_t unit = warp_global; unit < units; unit += n_warps) {
        const uint32_t tok = unit / nseg, seg = unit - tok * nseg;
        const uint32_t h_begin = seg * hps, h_end = min(H, h_begin + hps);
        // ---- per-token state ----------------------------------------------------------------------
        const float* cs = cos_sin + __ldg(positions + tok) * B200_HD + 8 * sub;          // [cos(64) | sin(64)]
        const float4 c0 = __ldg(reinterpret_cast<const float4*>(cs)), c1 = __ldg(reinterpret_cast<const float4*>(cs + 4));
        const float4 s0 = __ldg(reinterpret_cast<const float4*>(cs + 64)), s1 = __ldg(reinterpret_cast<const float4*>(cs + 68));
        const float c[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
        const float s[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
        const __nv_bfloat16* q_tok = q + (int64_t)tok * q_token_stride + 8 * sub;
        const __nv_bfloat16* k_tok = k + (int64_t)tok * k_token_stride + 8 * sub;
        const __nv_bfloat16* v_tok = STORE ? st.v + (int64_t)tok * st.v_token_stride + 8 * sub : nullptr;
        __nv_bfloat16* qo_tok = q_out + (int64_t)tok * Hq * B200_HD + 8 * sub;
        __nv_bfloat16* ko_tok = k_out != nullptr ? k_out + (int64_t)tok * Hkv * B200_HD + 8 * sub : nullptr;
        int64_t cache_off = -1;                                     // element offset of this token's slot, -1: no cache write
        if (STORE) {
            const int32_t slot = __ldg(st.slot_mapping + tok);
            if (slot >= 0) cache_off = (int64_t)slot * Hkv * B200_HD + 8 * sub;
        }
        auto src_of = [&](uint32_t h) -> const __nv_bfloat16* {
            return h < Hq ? q_tok + h * B200_HD : (STORE && h >= Hq + Hkv ? v_tok + (h - Hq - Hkv) * B200_HD : k_tok + (h - Hq) * B200_HD);
        };
        auto load = [&](uint32_t h) {
            RowData r;
            r.lo = r.hi = make_int4(0, 0, 0, 0);
            if (h < h_end) { const __nv_bfloat16* src = src_of(h); r.lo = ld_nc_v4(src); r.hi = ld_nc_v4(src + 64); }
            return r;
        };
        RowData cur = load(h_begin + grp), nxt = load(h_begin + grp + 4);
        for (uint32_t hb = h_begin; hb < h_end; hb += 4) {           // uniform trip count for the whole warp
            const uint32_t h = hb + grp;
            const RowData nxt2 = load(h + 8);                        // two head-quads (2 KB per warp) in flight
            const bool live = h < h_end, is_q = h < Hq, is_v = STORE && h >= Hq + Hkv;
            if (STORE && is_v && live && cache_off >= 0) {           // v is copied untouched
                __nv_bfloat16* dst = st.v_cache + cache_off + (int64_t)(h - Hq - Hkv) * B200_HD;
                st_v4(dst, cur.lo);
                st_v4(dst + 64, cur.hi);
            }
            const uint32_t lo_u[4] = {(uint32_t)cur.lo.x, (uint32_t)cur.lo.y, (uint32_t)cur.lo.z, (uint32_t)cur.lo.w};
            const uint32_t hi_u[4] = {(uint32_t)cur.hi.x, (uint32_t)cur.hi.y, (uint32_t)cur.hi.z, (uint32_t)cur.hi.w};
            float x1[8], x2[8];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                x1[2 * j] = bf_lo(lo_u[j]); x1[2 * j + 1] = bf_hi(lo_u[j]);
                x2[2 * j] = bf_lo(hi_u[j]); x2[2 * j + 1] = bf_hi(hi_u[j]);
            }
            if (has_w) {                                           // uniform: nullptr => rotation only (RotaryEmbedding.forward)
                float ss = 0.f;
#pragma unroll
                for (int j = 0; j < 8; ++j) ss += x1[j] * x1[j] + x2[j] * x2[j];
#pragma unroll
                for (int o = 4; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
                const float rstd = rsqrtf(__fadd_rn(ss * (1.0f / B200_HD), eps));
                // bf16(x * rstd) then bf16(that * w): two roundings, as in the reference.  The first is one packed F2FP
                // per pair; the second is ONE packed bf16 multiply: the product of two bf16 values is exact in fp32, so
                // HMUL2.BF16 (fp32 product, one round-to-nearest-even) returns the bits of bf16(float(y) * float(w)).
                const __nv_bfloat16* w = (is_q ? q_weight : k_weight) + 8 * sub;        // 2 x 16 bytes, L1-resident
                const int4 wl4 = __ldg(reinterpret_cast<const int4*>(w)), wh4 = __ldg(reinterpret_cast<const int4*>(w + 64));
                const uint32_t wl_u[4] = {(uint32_t)wl4.x, (uint32_t)wl4.y, (uint32_t)wl4.z, (uint32_t)wl4.w};
                const uint32_t wh_u[4] = {(uint32_t)wh4.x, (uint32_t)wh4.y, (uint32_t)wh4.z, (uint32_t)wh4.w};
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const uint32_t wl = wl_u[j], wh = wh_u[j];
                    const uint32_t a = pack_bf16x2(__fmul_rn(x1[2 * j], rstd), __fmul_rn(x1[2 * j + 1], rstd));
                    const uint32_t b = pack_bf16x2(__fmul_rn(x2[2 * j], rstd), __fmul_rn(x2[2 * j + 1], rstd));
                    const __nv_bfloat162 c2 = __hmul2(*reinterpret_cast<const __nv_bfloat162*>(&a), *reinterpret_cast<const __nv_bfloat162*>(&wl));
                    const __nv_bfloat162 d2 = __hmul2(*reinterpret_cast<const __nv_bfloat162*>(&b), *reinterpret_cast<const __nv_bfloat162*>(&wh));
                    const uint32_t cc = *reinterpret_cast<const uint32_t*>(&c2), dd = *reinterpret_cast<const uint32_t*>(&d2);
                    x1[2 * j] = bf_lo(cc); x1[2 * j + 1] = bf_hi(cc);
                    x2[2 * j] = bf_lo(dd); x2[2 * j + 1] = bf_hi(dd);
                }
            }
            if (live && !is_v) {
                uint32_t y1[4], y2[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    // y1 = x1*cos - x2*sin ; y2 = x2*cos + x1*sin  (un-fused, matching eager fp32)
                    const int a = 2 * j, b = 2 * j + 1;
                    y1[j] = pack_bf16x2(__fsub_rn(__fmul_rn(x1[a], c[a]), __fmul_rn(x2[a], s[a])),
                                        __fsub_rn(__fmul_rn(x1[b], c[b]), __fmul_rn(x2[b], s[b])));
                    y2[j] = pack_bf16x2(__fadd_rn(__fmul_rn(x2[a], c[a]), __fmul_rn(x1[a], s[a])),
                                        __fadd_rn(__fmul_rn(x2[b], c[b]), __fmul_rn(x1[b], s[b])));
                }
                const int4 o_lo = make_int4(y1[0], y1[1], y1[2], y1[3]), o_hi = make_int4(y2[0], y2[1], y2[2], y2[3]);
                __nv_bfloat16* dst = is_q ? qo_tok + h * B200_HD : (ko_tok != nullptr ? ko_tok + (h - Hq) * B200_HD : nullptr);
                if (dst != nullptr) {
                    st_v4(dst, o_lo);
                    st_v4(dst + 64, o_hi);
                }
                if (STORE && !is_q && cache_off >= 0) {
                    __nv_bfloat16* cd = st.k_cache + cache_off + (int64_t)(h - Hq) * B200_HD;
                    st_v4(cd, o_lo);
                    st_v4(cd + 64, o_hi);
                }
            }
            cur = nxt; nxt = nxt2;
        }
    }
}

// ------------------------------------------------------------------------------------------
// (4a') hidden-size RMSNorm with optional residual add.  One CTA per token, 16-byte loads,
//       values stay in registers between the reduction and the scaling pass.
// ------------------------------------------------------------------------------------------
template <int MAX_CHUNKS>
__global__ void __launch_bounds__(256) rmsnorm_kernel(
    const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ residual,
    const __nv_bfloat16* __restrict__ weight, float eps, __nv_bfloat16* __restrict__ out, int32_t hidden) {
    __shared__ float red[8];
    const int64_t base = (int64_t)blockIdx.x * hidden;
    const int nchunks = hidden >> 3;
    float v[MAX_CHUNKS][8];
    float ss = 0.f;
#pragma unroll
    for (int it = 0; it < MAX_CHUNKS; ++it) {
        const int c = threadIdx.x + it * 256;
        if (c < nchunks) {
            const int4 xi = *reinterpret_cast<const int4*>(x + base + c * 8);
            const uint32_t xs[4] = {(uint32_t)xi.x, (uint32_t)xi.y, (uint32_t)xi.z, (uint32_t)xi.w};
            if (residual != nullptr) {
                const int4 ri = *reinterpret_cast<const int4*>(residual + base + c * 8);
                const uint32_t rs[4] = {(uint32_t)ri.x, (uint32_t)ri.y, (uint32_t)ri.z, (uint32_t)ri.w};
                uint32_t ro[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float2 a = unpack_bf16x2(xs[j]), b = unpack_bf16x2(rs[j]);
                    v[it][2 * j] = __fadd_rn(a.x, b.x);
                    v[it][2 * j + 1] = __fadd_rn(a.y, b.y);
                    ro[j] = pack_bf16x2(v[it][2 * j], v[it][2 * j + 1]);
                }
                *reinterpret_cast<int4*>(residual + base + c * 8) = make_int4(ro[0], ro[1], ro[2], ro[3]);
            } else {
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float2 a = unpack_bf16x2(xs[j]);
                    v[it][2 * j] = a.x;
                    v[it][2 * j + 1] = a.y;
                }
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) ss += v[it][j] * v[it][j];
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
    __syncthreads();
    float tot = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) tot += red[i];
    const float rstd = rsqrtf(__fadd_rn(tot / (float)hidden, eps));
#pragma unroll
    for (int it = 0; it < MAX_CHUNKS; ++it) {
        const int c = threadIdx.x + it * 256;
        if (c < nchunks) {
            const int4 wi = *reinterpret_cast<const int4*>(weight + c * 8);
            const uint32_t ws[4] = {(uint32_t)wi.x, (uint32_t)wi.y, (uint32_t)wi.z, (uint32_t)wi.w};
            uint32_t o[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float2 w2 = unpack_bf16x2(ws[j]);
                const float a = __bfloat162float(__float2bfloat16_rn(__fmul_rn(v[it][2 * j], rstd)));
                const float b = __bfloat162float(__float2bfloat16_rn(__fmul_rn(v[it][2 * j + 1], rstd)));
                o[j] = pack_bf16x2(__fmul_rn(a, w2.x), __fmul_rn(b, w2.y));
            }
            *reinterpret_cast<int4*>(out + base + c * 8) = make_int4(o[0], o[1], o[2], o[3]);
        }
    }
}

// ------------------------------------------------------------------------------------------
// (4b) SiluAndMul: out[t, i] = silu(x[t, i]) * x[t, I + i], fp32, one rounding.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) silu_and_mul_kernel(
    const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ out, int64_t total_chunks, int32_t chunks_per_row) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total_chunks; i += stride) {
        const int64_t t = i / chunks_per_row;
        const int64_t c = i - t * chunks_per_row;
        const __nv_bfloat16* row = x + t * (int64_t)chunks_per_row * 16;
        const int4 gi = ld_nc_v4(row + c * 8);
        const int4 ui = ld_nc_v4(row + (int64_t)chunks_per_row * 8 + c * 8);
        const uint32_t gs[4] = {(uint32_t)gi.x, (uint32_t)gi.y, (uint32_t)gi.z, (uint32_t)gi.w};
        const uint32_t us[4] = {(uint32_t)ui.x, (uint32_t)ui.y, (uint32_t)ui.z, (uint32_t)ui.w};
        uint32_t o[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float2 g = unpack_bf16x2(gs[j]), u = unpack_bf16x2(us[j]);
            const float a = __fdividef(g.x, 1.0f + __expf(-g.x)) * u.x;
            const float b = __fdividef(g.y, 1.0f + __expf(-g.y)) * u.y;
            o[j] = pack_bf16x2(a, b);
        }
        st_na_v4(out + i * 8, make_int4(o[0], o[1], o[2], o[3]));
    }
}

// ------------------------------------------------------------------------------------------
// host launchers
// ------------------------------------------------------------------------------------------
static inline int grid_for(int64_t threads_needed, int block, int waves_cap = 8) {
    int64_t blocks = (threads_needed + block - 1) / block;
    const int64_t cap = (int64_t)sm_count() * waves_cap;       // multiples of the SM count; grid-stride beyond
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return (int)blocks;
}

int launch_kv_scatter(const void* key, const void* value, int64_t key_row_stride, int64_t value_row_stride,
                      void* k_cache, void* v_cache, const int32_t* slot_mapping,
                      int32_t T, int32_t Hkv, int32_t D, cudaStream_t stream) {
    const uint32_t chunks = (uint32_t)(Hkv * D * 2 / 16);
    uint32_t lg = 0;
    while ((1u << lg) < chunks) ++lg;
    const int64_t total = (int64_t)T * chunks;
    if (total >= (int64_t)1 << 32) return -5;                     // 64 GB of K rows in one call: not a real shape
    const int grid = grid_for((total + SC_UNROLL - 1) / SC_UNROLL, 256);
    if ((1u << lg) == chunks)
        kv_scatter_kernel<true><<<grid, 256, 0, stream>>>(
            (const uint8_t*)key, (const uint8_t*)value, key_row_stride * 2, value_row_stride * 2,
            (uint8_t*)k_cache, (uint8_t*)v_cache, slot_mapping, (uint32_t)total, chunks, lg);
    else                                                          // non power-of-two rows (e.g. 5 heads)
        kv_scatter_kernel<false><<<grid, 256, 0, stream>>>(
            (const uint8_t*)key, (const uint8_t*)value, key_row_stride * 2, value_row_stride * 2,
            (uint8_t*)k_cache, (uint8_t*)v_cache, slot_mapping, (uint32_t)total, chunks, lg);
    return (int)cudaGetLastError();
}

int launch_kv_scatter_generic(const void* key, const void* value, int64_t key_row_stride, int64_t value_row_stride,
                              void* k_cache, void* v_cache, const int32_t* slot_mapping,
                              int32_t T, int32_t Hkv, int32_t D, cudaStream_t stream) {
    return launch_kv_scatter(key, value, key_row_stride, value_row_stride, k_cache, v_cache, slot_mapping, T, Hkv, D, stream);
}

// Units = (token, head segment).  One segment per token when there are enough tokens to fill the machine (the whole
// cos/sin row, position and slot are then fetched once per token); at decode sizes the heads of a token are split into
// up to ceil(H/4) segments so that ~8 warps per SM sub-partition still have work.
struct NormRopeShape { uint32_t nseg, hps; int grid; };
template <bool STORE>
static inline int norm_rope_resident_blocks() {                  // CTAs of this kernel an SM holds (per device, cached)
    static int cached[64] = {0};
    const int dev = current_device();
    if (dev < 0 || dev >= 64) return 2;
    if (cached[dev] == 0) {
        int n = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, qk_norm_rope_kernel<STORE>, 256, 0) != cudaSuccess || n < 1) n = 2;
        cached[dev] = n;
    }
    return cached[dev];
}
static inline NormRopeShape norm_rope_shape(int64_t T, int H, int resident) {
    const int64_t want_units = (int64_t)sm_count() * 32 * 2;      // two units per resident warp
    int64_t nseg = T > 0 ? (want_units + T - 1) / T : 1;
    const int max_seg = (H + 3) / 4;
    if (nseg < 1) nseg = 1;
    if (nseg > max_seg) nseg = max_seg;
    int hps = (int)(((H + nseg - 1) / nseg + 3) / 4 * 4);
    nseg = (H + hps - 1) / hps;
    const int64_t units = T * nseg;
    int64_t blocks = (units + 7) / 8;                              // 8 warps per CTA, one unit per warp per round
    const int64_t cap = (int64_t)sm_count() * resident;           // exactly one wave: the grid-stride loop takes the rest
    if (blocks > cap) blocks = cap;
    return {(uint32_t)nseg, (uint32_t)hps, (int)(blocks < 1 ? 1 : blocks)};
}

int launch_qk_norm_rope(const void* q, int64_t q_token_stride, const void* k, int64_t k_token_stride,
                        const int64_t* positions, const float* cos_sin, const void* qw, const void* kw, float eps,
                        void* q_out, void* k_out, int32_t T, int32_t Hq, int32_t Hkv, cudaStream_t stream) {
    const int64_t rows = (int64_t)T * (Hq + Hkv);
    if (rows >= (int64_t)1 << 31) return -5;
    if (T == 0) return 0;
    const NormRopeShape sh = norm_rope_shape(T, Hq + Hkv, norm_rope_resident_blocks<false>());
    qk_norm_rope_kernel<false><<<sh.grid, 256, 0, stream>>>(
        (const __nv_bfloat16*)q, q_token_stride, (const __nv_bfloat16*)k, k_token_stride, positions, cos_sin,
        (const __nv_bfloat16*)qw, (const __nv_bfloat16*)kw, eps, (__nv_bfloat16*)q_out, (__nv_bfloat16*)k_out,
        (uint32_t)T, (uint32_t)Hq, (uint32_t)Hkv, sh.nseg, sh.hps, KvStoreArgs{});
    return (int)cudaGetLastError();
}

int launch_qk_norm_rope_store(const void* q, int64_t q_token_stride, const void* k, int64_t k_token_stride,
                              const void* v, int64_t v_token_stride,
                              const int64_t* positions, const float* cos_sin, const void* qw, const void* kw, float eps,
                              void* q_out, void* k_out, void* k_cache, void* v_cache, const int32_t* slot_mapping,
                              int32_t T, int32_t Hq, int32_t Hkv, cudaStream_t stream) {
    const int64_t rows = (int64_t)T * (Hq + 2 * Hkv);
    if (rows >= (int64_t)1 << 31) return -5;
    if (T == 0) return 0;
    KvStoreArgs st;
    st.v = (const __nv_bfloat16*)v; st.v_token_stride = v_token_stride;
    st.k_cache = (__nv_bfloat16*)k_cache; st.v_cache = (__nv_bfloat16*)v_cache; st.slot_mapping = slot_mapping;
    const NormRopeShape sh = norm_rope_shape(T, Hq + 2 * Hkv, norm_rope_resident_blocks<true>());
    qk_norm_rope_kernel<true><<<sh.grid, 256, 0, stream>>>(
        (const __nv_bfloat16*)q, q_token_stride, (const __nv_bfloat16*)k, k_token_stride, positions, cos_sin,
        (const __nv_bfloat16*)qw, (const __nv_bfloat16*)kw, eps, (__nv_bfloat16*)q_out, (__nv_bfloat16*)k_out,
        (uint32_t)T, (uint32_t)Hq, (uint32_t)Hkv, sh.nseg, sh.hps, st);
    return (int)cudaGetLastError();
}

int launch_rmsnorm(const void* x, void* residual, const void* weight, float eps, void* out,
                   int32_t T, int32_t hidden, cudaStream_t stream) {
    const int chunks = hidden / 8;
    if (chunks <= 256 * 2)
        rmsnorm_kernel<2><<<T, 256, 0, stream>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)residual,
                                                 (const __nv_bfloat16*)weight, eps, (__nv_bfloat16*)out, hidden);
    else
        rmsnorm_kernel<8><<<T, 256, 0, stream>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)residual,
                                                 (const __nv_bfloat16*)weight, eps, (__nv_bfloat16*)out, hidden);
    return (int)cudaGetLastError();
}

int launch_silu_and_mul(const void* x, void* out, int32_t T, int32_t inter, cudaStream_t stream) {
    const int32_t chunks_per_row = inter / 8;
    const int64_t total = (int64_t)T * chunks_per_row;
    silu_and_mul_kernel<<<grid_for(total, 256, 16), 256, 0, stream>>>(
        (const __nv_bfloat16*)x, (__nv_bfloat16*)out, total, chunks_per_row);
    return (int)cudaGetLastError();
}

}  // namespace b200
