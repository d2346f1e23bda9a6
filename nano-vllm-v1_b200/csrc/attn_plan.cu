// Step plan for varlen attention calls (mixed prefill + decode batches).
//
// One tiny kernel, run ONCE per engine step (every layer of the step reuses its output), turns
// cu_seqlens_q/k into
//   * the UNIT TABLE of the persistent tcgen05 prefill kernel: one 32-byte record per work unit (two query tiles
//     sharing a K/V stream; launch.h), for every query tile (128 rows when the kernel pairs q heads of a GQA group,
//     else 256 rows) of a sequence with q_len > 1 (or every sequence when routing is off), in descending order of
//     KV tiles (longest-processing-time-first for the static snake over the persistent CTAs).  An item
//     contributes Hq/2 units, Hq/4 when its tile is packed (<= 64 valid rows: two heads per tile), Hq in row-pair
//     mode, so the counting sort runs on unit counts: weighted histogram -> scan -> every item claims its run of
//     unit slots -> all threads expand items into unit records;
//   * the list of rows with exactly one query token, which go to the split-K decode kernel
//     (a 128-row UMMA tile would be <1 % utilised by such a row).
// Everything stays on the device: no host read of lengths, so the call is graph-capturable.
//
// The reference has no counterpart: flash-attn schedules inside its own launch
// (attention.py:242-248 passes only cu_seqlens and max_seqlen).
#include "common.cuh"
#include "launch.h"

namespace b200 {

constexpr int PLAN_THREADS = 1024;
constexpr int PLAN_MAX_KEY = PLAN_THREADS - 1;          // one histogram bin per thread (40,960-token context = 320 tiles)

__global__ void __launch_bounds__(PLAN_THREADS) attn_plan_kernel(
    const int32_t* __restrict__ cu_q, const int32_t* __restrict__ cu_k, int32_t batch, int32_t route_decode,
    int32_t tile_rows, int32_t units_per_item, int32_t pack_ok, int32_t* __restrict__ plan, int32_t max_items) {
    __shared__ int32_t hist[PLAN_MAX_KEY + 1];
    __shared__ int32_t n_decode, n_items;
    __shared__ int32_t warp_tot[PLAN_THREADS / 32];
    for (int i = threadIdx.x; i <= PLAN_MAX_KEY; i += PLAN_THREADS) hist[i] = 0;
    if (threadIdx.x == 0) { n_decode = 0; n_items = 0; }
    __syncthreads();
    int32_t* decode_rows = plan + PLAN_HEADER_WORDS;
    int32_t* items = plan + plan_items_offset(batch);                       // scratch: {sequence, tile, first unit, units}
    int32_t* units = plan + plan_units_offset(batch, max_items);

    auto key_of = [tile_rows](int mt, int Lq, int Lk) {
        const int kv_end = min(Lk, (mt + 1) * tile_rows + (Lk - Lq));
        const int k = (kv_end + 127) / 128;
        return min(max(k, 0), PLAN_MAX_KEY);
    };
    // units of one item: a packed tile (pair mode, <= 64 rows of the tile inside the sequence) carries two q heads
    // per 128-row tile, so it needs half as many units
    auto packed = [tile_rows, pack_ok](int mt, int Lq) { return pack_ok && tile_rows == 128 && Lq - mt * 128 <= 64; };
    auto mult_of = [&](int mt, int Lq) { return packed(mt, Lq) ? units_per_item / 2 : units_per_item; };
    // pass 1: histogram of UNIT counts per tile cost; single-token rows go to the decode list
    for (int b = threadIdx.x; b < batch; b += PLAN_THREADS) {
        const int Lq = cu_q[b + 1] - cu_q[b], Lk = cu_k[b + 1] - cu_k[b];
        if (Lq <= 0) continue;
        if (route_decode && Lq == 1) {
            decode_rows[atomicAdd(&n_decode, 1)] = b;
            continue;
        }
        const int tiles = (Lq + tile_rows - 1) / tile_rows;
        for (int mt = 0; mt < tiles; ++mt) atomicAdd(&hist[key_of(mt, Lq, Lk)], mult_of(mt, Lq));
    }
    __syncthreads();
    // descending exclusive scan over the bins: thread i owns bin (MAX_KEY - i)
    {
        const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
        const int mine = hist[PLAN_MAX_KEY - threadIdx.x];
        int inc = mine;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += v;
        }
        if (lane == 31) warp_tot[wid] = inc;
        __syncthreads();
        if (wid == 0) {
            int w = warp_tot[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int v = __shfl_up_sync(0xffffffffu, w, o);
                if (lane >= o) w += v;
            }
            warp_tot[lane] = w;                       // inclusive totals of warps 0..lane
        }
        __syncthreads();
        const int before = (wid > 0 ? warp_tot[wid - 1] : 0) + inc - mine;
        hist[PLAN_MAX_KEY - threadIdx.x] = before;    // first unit of this key (heavier keys come first)
        if (threadIdx.x == PLAN_THREADS - 1) {
            plan[0] = before + mine;                  // number of prefill work units
            plan[1] = n_decode;                       // number of single-token rows
            plan[2] = tile_rows;
        }
    }
    __syncthreads();
    // pass 2: every item claims its run of unit slots inside its key's range
    for (int b = threadIdx.x; b < batch; b += PLAN_THREADS) {
        const int Lq = cu_q[b + 1] - cu_q[b], Lk = cu_k[b + 1] - cu_k[b];
        if (Lq <= 0 || (route_decode && Lq == 1)) continue;
        const int tiles = (Lq + tile_rows - 1) / tile_rows;
        for (int mt = tiles - 1; mt >= 0; --mt) {
            const int m = mult_of(mt, Lq);
            const int pos = atomicAdd(&hist[key_of(mt, Lq, Lk)], m);
            const int it = atomicAdd(&n_items, 1);
            if (it < max_items) *reinterpret_cast<int4*>(items + PLAN_ITEM_WORDS * it) = make_int4(b, mt, pos, m);
        }
    }
    __syncthreads();
    const int ni = min(n_items, max_items);
    if (threadIdx.x == 0) plan[3] = ni;
    // pass 3: expand the items into unit records, all threads, one (item, sub) pair at a time
    for (int idx = threadIdx.x; idx < ni * units_per_item; idx += PLAN_THREADS) {
        const int it = idx / units_per_item, sub = idx - it * units_per_item;
        const int4 rec = *reinterpret_cast<const int4*>(items + PLAN_ITEM_WORDS * it);
        if (sub >= rec.w) continue;
        const int b = rec.x, mt = rec.y;
        const int q0 = cu_q[b], k0 = cu_k[b];
        int4* u = reinterpret_cast<int4*>(units + (int64_t)PLAN_UNIT_WORDS * (rec.z + sub));
        u[0] = make_int4(b, mt, q0, cu_q[b + 1] - q0);
        u[1] = make_int4(k0, cu_k[b + 1] - k0, sub, rec.w < units_per_item ? 2 : 1);
    }
}

int launch_attn_plan(const int32_t* cu_q, const int32_t* cu_k, int32_t batch, int32_t route_decode, int32_t tile_rows,
                     int32_t units_per_item, int32_t pack_ok, int32_t* plan, int32_t max_items, cudaStream_t stream) {
    attn_plan_kernel<<<1, PLAN_THREADS, 0, stream>>>(cu_q, cu_k, batch, route_decode, tile_rows, units_per_item, pack_ok, plan, max_items);
    return (int)cudaGetLastError();
}

}  // namespace b200
