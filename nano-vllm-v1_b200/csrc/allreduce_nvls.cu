// (f3) Tensor-parallel all-reduce FUSED with the residual add + RMSNorm that follows it, on NVLS multicast
// memory.
//
// Every RowParallelLinear of the model (o_proj, down_proj: linear.py:407-410) is followed by
// RMSNorm.add_rms_forward (layernorm.py:95-141, called at qwen3.py:370,374,440).  The reference runs
// NCCL all-reduce -> inductor add+RMSNorm: two launches and one extra [T, H] round trip, 73 times per decode
// step of Qwen3-8B.  Here ONE launch per rank does both, and the reduction itself happens inside the NVSwitch:
//
//   buffer (symmetric: same layout on every rank, mapped twice -- unicast `local`, multicast `mc`):
//       [ slot 0 | slot 1 | flag A | flag B ]
//   copy-in      every block stores its share of the rank's partial sums x into the local slot (slots
//                alternate with a device-resident epoch, so CUDA-graph replays stay consistent and a rank
//                that runs ahead never overwrites what a slower peer still reads);
//   barrier A    blocks of a rank meet on a local counter, block 0 alone signals the group with ONE
//                multimem.red.add on the multicast flag (it lands in every rank's flag word) and polls its
//                own local flag until all `world` increments arrived -- no per-peer stores, no per-peer polls;
//   two-shot     rank r owns 1/world of the message: multimem.ld_reduce (the switch adds the `world` bf16x2
//                inputs in fp32 and returns one rounded result) -> multimem.st of the sum into the same place
//                of EVERY rank's slot; barrier B; then each rank reads the full sum locally;
//   one-shot     (small messages) every rank multimem.ld_reduces every row itself: no second barrier;
//   epilogue     per row: v = float(sum) + float(residual); residual <- bf16(v); out <- bf16(bf16(v * rstd) * w)
//                with rstd from the unrounded fp32 v -- the roundings of add_rms_forward.
//
// All ranks run the same instruction sequence on bit-identical inputs to the epilogue (one reducer per element in
// the two-shot form; the same switch reduction in the one-shot form), so hidden states stay identical across ranks.
// weight == NULL turns the epilogue off: out <- sum (a plain all-reduce for callers not followed by a norm).
// MC = false (world == 1, single-GPU tests) replaces the multimem instructions by plain local accesses.
#include <cstring>

#include "../../include/b200attn.h"
#include "launch.h"
#include "common.cuh"

namespace b200 {

constexpr int NV_MAX_WORLD = 8;
constexpr int NV_BLOCKS = 296;                   // two CTAs per SM: a CTA owns one row at a time
constexpr int NV_THREADS = 512;
constexpr int NV_MAX_CHUNKS = 4;                // 16-byte chunks of a row per thread: hidden <= 512 * 4 * 8 = 16384

struct NvlsComm {
    uint8_t* local; uint8_t* mc;                // unicast / multicast mapping of the symmetric buffer
    uint8_t* peer[NV_MAX_WORLD];                // unicast mappings of every rank's buffer (peer[rank] == local); LL two-shot
    int32_t rank, world;
    size_t slot_bytes;
    size_t ll_max_bytes;                        // largest message of the flag-in-data (LL) forms
    size_t ll_slot_bytes;                       // one LL slot = [ contributions: 2 x ll_max + 1 KB | results: 2 x ll_max ]
    uint32_t* state;                            // device: [0] epoch, [1] blocks done, [2],[3] arrive/go A, [4],[5] arrive/go B, [7] sticky error
    uint64_t timeout_ns;
};

struct NvlsHandle {
    NvlsComm dev;
    int32_t max_blocks;                         // co-residency caps (cooperative launch): barrier forms / flag-in-data forms
    int32_t max_blocks_ll;
    bool has_peers;                             // unicast peer mappings given: the LL two-shot form is available
};

__device__ __forceinline__ uint32_t nv_ld_acquire_gpu(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void nv_st_release_gpu(uint32_t* p, uint32_t v) {
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t nv_ld_relaxed_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
// in-switch reduction of one 16-byte chunk (8 bf16) over all ranks, fp32 accumulation, one rounding
__device__ __forceinline__ uint4 multimem_ld_reduce_bf16x8(const void* mc_addr) {
    uint4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(mc_addr) : "memory");
    return v;
}
// one 16-byte store replicated by the switch into every rank's copy
__device__ __forceinline__ void multimem_st_v4(void* mc_addr, const uint4& v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
                 ::"l"(mc_addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void multimem_red_add_u32(void* mc_addr, uint32_t v) {
    asm volatile("multimem.red.release.sys.global.add.u32 [%0], %1;" ::"l"(mc_addr), "r"(v) : "memory");
}

// Bounded wait.  The error word is sticky: once any wait of this communicator timed out, later waits give up at
// their first check, so a broken group costs ONE timeout, not one per barrier of every queued launch.
template <typename F>
__device__ __forceinline__ void nv_spin(const NvlsComm& c, F done) {
    uint64_t t0 = 0;
    for (uint32_t spins = 0; !done(); ++spins) {
        if ((spins & 1023u) == 1023u) {
            if (*reinterpret_cast<volatile uint32_t*>(c.state + 7) != 0u) return;
            uint64_t now;
            asm volatile("mov.u64 %0, %globaltimer;" : "=l"(now));
            if (t0 == 0) t0 = now;
            else if (now - t0 > c.timeout_ns) { atomicExch(c.state + 7, 1u); return; }
        }
    }
}

// All blocks of all ranks meet here.  Flags are monotonic counters: after call number `epoch` every rank's
// flag word holds epoch * world (mod 2^32; compared by signed difference).
template <bool MC>
__device__ __forceinline__ void nv_barrier(const NvlsComm& c, int which, uint32_t epoch) {
    uint32_t* arrive = c.state + 2 + 2 * which;
    uint32_t* go = arrive + 1;
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();                        // this block's stores before its arrival
        atomicAdd(arrive, 1u);
        if (blockIdx.x == 0) {
            nv_spin(c, [&] { return nv_ld_acquire_gpu(arrive) >= gridDim.x; });
            if (MC) {
                const size_t off = 2 * c.slot_bytes + (size_t)which * 64;
                __threadfence_system();         // everything this rank wrote, before the group sees the signal
                multimem_red_add_u32(c.mc + off, 1u);
                const uint32_t* mine = reinterpret_cast<const uint32_t*>(c.local + off);
                const uint32_t target = epoch * (uint32_t)c.world;
                nv_spin(c, [&] { return (int32_t)(nv_ld_relaxed_sys(mine) - target) >= 0; });
                __threadfence_system();         // acquire: peers' stores are visible to what follows
            }
            nv_st_release_gpu(go, epoch);
        }
        nv_spin(c, [&] { return (int32_t)(nv_ld_acquire_gpu(go) - epoch) >= 0; });
    }
    __syncthreads();
}

__device__ __forceinline__ float block_sum_512(float v, float* red) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();                            // `red` may still be read from the previous row
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    float tot = 0.f;
#pragma unroll
    for (int i = 0; i < NV_THREADS / 32; ++i) tot += red[i];
    return tot;
}

// Epilogue of one row held in registers (`sum` = the all-reduced bf16 values of this thread's chunks):
// plain all-reduce (weight NULL) or v = float(sum) + float(residual); residual <- bf16(v); out <- bf16(bf16(v*rstd)*w).
__device__ __forceinline__ void row_epilogue(const uint4 (&sum)[NV_MAX_CHUNKS], int64_t base, int32_t cpr, uint4* residual,
                                             const uint4* __restrict__ weight, float eps, uint4* out, float* red) {
        if (weight == nullptr) {                                // plain all-reduce
#pragma unroll
            for (int kk = 0; kk < NV_MAX_CHUNKS; ++kk) {
                const int cidx = threadIdx.x + kk * NV_THREADS;
                if (cidx < cpr) out[base + cidx] = sum[kk];
            }
            return;
        }
        float v[NV_MAX_CHUNKS][8];
        float ss = 0.f;
#pragma unroll
        for (int kk = 0; kk < NV_MAX_CHUNKS; ++kk) {
            const int cidx = threadIdx.x + kk * NV_THREADS;
            if (cidx < cpr) {
                const uint32_t xs[4] = {sum[kk].x, sum[kk].y, sum[kk].z, sum[kk].w};
                if (residual != nullptr) {
                    const uint4 ri = residual[base + cidx];
                    const uint32_t rs[4] = {ri.x, ri.y, ri.z, ri.w};
                    uint32_t ro[4];
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const float2 a = unpack_bf16x2(xs[j]), b = unpack_bf16x2(rs[j]);
                        v[kk][2 * j] = __fadd_rn(a.x, b.x);
                        v[kk][2 * j + 1] = __fadd_rn(a.y, b.y);
                        ro[j] = pack_bf16x2(v[kk][2 * j], v[kk][2 * j + 1]);
                    }
                    residual[base + cidx] = make_uint4(ro[0], ro[1], ro[2], ro[3]);
                } else {
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const float2 a = unpack_bf16x2(xs[j]);
                        v[kk][2 * j] = a.x; v[kk][2 * j + 1] = a.y;
                    }
                }
#pragma unroll
                for (int j = 0; j < 8; ++j) ss += v[kk][j] * v[kk][j];
            }
        }
        const float tot = block_sum_512(ss, red);
        const float rstd = rsqrtf(__fadd_rn(tot / (float)(cpr * 8), eps));
#pragma unroll
        for (int kk = 0; kk < NV_MAX_CHUNKS; ++kk) {
            const int cidx = threadIdx.x + kk * NV_THREADS;
            if (cidx < cpr) {
                const uint4 wi = weight[cidx];
                const uint32_t ws[4] = {wi.x, wi.y, wi.z, wi.w};
                uint32_t o[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float2 w2 = unpack_bf16x2(ws[j]);
                    const float a = __bfloat162float(__float2bfloat16_rn(__fmul_rn(v[kk][2 * j], rstd)));
                    const float b = __bfloat162float(__float2bfloat16_rn(__fmul_rn(v[kk][2 * j + 1], rstd)));
                    o[j] = pack_bf16x2(__fmul_rn(a, w2.x), __fmul_rn(b, w2.y));
                }
                out[base + cidx] = make_uint4(o[0], o[1], o[2], o[3]);
            }
        }
}

// x: this rank's partial sums [T, hidden] bf16 (16-byte chunks); residual (in place) / weight may be NULL together.
template <bool MC, bool ONE_SHOT>
__global__ void __launch_bounds__(NV_THREADS) nvls_ar_norm_kernel(
    const NvlsComm c, const uint4* __restrict__ x, uint4* residual, const uint4* __restrict__ weight, float eps,
    uint4* out, int32_t T, int32_t cpr /* 16-byte chunks per row */) {
    __shared__ float red[NV_THREADS / 32];
    const uint32_t epoch = c.state[0] + 1;
    const size_t slot_off = (epoch & 1u) * c.slot_bytes;
    uint4* slot = reinterpret_cast<uint4*>(c.local + slot_off);
    uint4* mc_slot = reinterpret_cast<uint4*>(c.mc + slot_off);
    const int64_t n16 = (int64_t)T * cpr;
    const int64_t stride = (int64_t)gridDim.x * NV_THREADS;
    const int64_t gtid = (int64_t)blockIdx.x * NV_THREADS + threadIdx.x;

    for (int64_t i = gtid; i < n16; i += stride) slot[i] = x[i];
    nv_barrier<MC>(c, 0, epoch);

    if (ONE_SHOT) {
        // Flag B counts CALLS like flag A (both hold epoch * world after every call, whatever mix of forms ran):
        // the one-shot form has no second barrier to wait at, but it still posts its increment.
        if (MC && blockIdx.x == 0 && threadIdx.x == 0) multimem_red_add_u32(c.mc + 2 * c.slot_bytes + 64, 1u);
    } else {
        const int64_t per = (n16 + c.world - 1) / c.world;
        const int64_t lo = (int64_t)c.rank * per, hi = min(lo + per, n16);
        int64_t i = lo + gtid;
        for (; i + stride < hi; i += 2 * stride) {              // two independent switch round trips in flight
            const uint4 a = MC ? multimem_ld_reduce_bf16x8(mc_slot + i) : slot[i];
            const uint4 b = MC ? multimem_ld_reduce_bf16x8(mc_slot + i + stride) : slot[i + stride];
            if (MC) { multimem_st_v4(mc_slot + i, a); multimem_st_v4(mc_slot + i + stride, b); }
        }
        for (; i < hi; i += stride) {
            const uint4 a = MC ? multimem_ld_reduce_bf16x8(mc_slot + i) : slot[i];
            if (MC) multimem_st_v4(mc_slot + i, a);
        }
        nv_barrier<MC>(c, 1, epoch);
    }

    // one row per CTA at a time; the next row's (switch) loads are issued before this row's epilogue
    auto load_row = [&](int32_t row, uint4 (&dst)[NV_MAX_CHUNKS]) {
        const int64_t base = (int64_t)row * cpr;
#pragma unroll
        for (int kk = 0; kk < NV_MAX_CHUNKS; ++kk) {
            const int cidx = threadIdx.x + kk * NV_THREADS;
            if (cidx < cpr) dst[kk] = (ONE_SHOT && MC) ? multimem_ld_reduce_bf16x8(mc_slot + base + cidx) : __ldcg(slot + base + cidx);
        }
    };
    uint4 cur[NV_MAX_CHUNKS], nxt[NV_MAX_CHUNKS];
    if ((int32_t)blockIdx.x < T) load_row(blockIdx.x, cur);
    for (int32_t row = blockIdx.x; row < T; row += gridDim.x) {
        const int32_t ahead = row + gridDim.x;
        if (ahead < T) load_row(ahead, nxt);
        row_epilogue(cur, (int64_t)row * cpr, cpr, residual, weight, eps, out, red);
#pragma unroll
        for (int kk = 0; kk < NV_MAX_CHUNKS; ++kk) cur[kk] = nxt[kk];
    }

    // last block out advances the epoch for the next launch (graph replays read it from device memory)
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        if (atomicAdd(&c.state[1], 1u) == gridDim.x - 1) {
            c.state[1] = 0;
            c.state[2] = 0;
            c.state[4] = 0;
            __threadfence();
            c.state[0] = epoch;
        }
    }
}

// ---------------------------------------------------------------------------------------------------------
// Flag-in-data ("LL") forms: no barrier at all.  A 16-byte chunk of 8 bf16 travels as two 16-byte packets
// {d0, e, d1, e}: the epoch e of the call rides in the same store as the data it guards, so a reader that sees
// both flags of a packet sees its data -- nothing else orders the two ranks (8-byte store atomicity suffices;
// NCCL's LL protocol rests on the same property).  Slots alternate with the epoch and are never reset: a stale
// packet carries an older epoch.  Twice the bytes on the wire, one NVLink hop of latency per phase:
//   LL one-shot  every rank multicasts its packets into region [rank] of every rank (one multimem.st per packet,
//                replicated by the switch), then sums the `world` regions of each row in rank order in fp32.
//   LL two-shot  reduce-scatter by peer stores of packets (part p of the message into rank p's contribution
//                area), the owner sums its part in rank order and multicasts the result packets into the result
//                area of every rank; every rank then reads whole rows.
// Sums are taken in rank order in fp32 and rounded once: bit-identical on every rank and equal to the peer-store
// forms of allreduce.cu.  Blocks of a rank never wait for each other; they do wait for packets that other
// ranks' blocks produce, hence the co-resident (cooperative) launch.
__device__ __forceinline__ void ll_store(uint4* dst, uint32_t d0, uint32_t d1, uint32_t e) {
    asm volatile("st.relaxed.sys.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(dst), "r"(d0), "r"(e), "r"(d1), "r"(e) : "memory");
}
__device__ __forceinline__ void ll_store_mc(uint4* mc_dst, uint32_t d0, uint32_t d1, uint32_t e) {
    multimem_st_v4(mc_dst, make_uint4(d0, e, d1, e));
}
__device__ __forceinline__ uint4 ll_load(const uint4* p) {
    uint4 v;
    asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}
// the two packets of one chunk -> its 16 data bytes (spins until both carry epoch e)
__device__ __forceinline__ uint4 ll_wait_chunk(const NvlsComm& c, const uint4* pk, uint32_t e) {
    uint4 a = ll_load(pk), b = ll_load(pk + 1);
    if (!(a.y == e && a.w == e && b.y == e && b.w == e))
        nv_spin(c, [&] { a = ll_load(pk); b = ll_load(pk + 1); return a.y == e && a.w == e && b.y == e && b.w == e; });
    return make_uint4(a.x, a.z, b.x, b.z);
}
__device__ __forceinline__ void nv_acc_bf16x8(float (&acc)[8], const uint4& v) {
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const float2 f = unpack_bf16x2(w[i]);
        acc[2 * i] += f.x; acc[2 * i + 1] += f.y;
    }
}
// fp32 sum, in rank order, of the chunk every rank sent: the packet loads of four ranks are in flight before the
// first flag is looked at (one or two L2 round trips when the packets are already there, not `world` of them)
__device__ __forceinline__ void ll_sum_ranks(const NvlsComm& c, const uint4* pk, int64_t rank_stride, uint32_t e, float (&acc)[8]) {
#pragma unroll
    for (int r0 = 0; r0 < NV_MAX_WORLD; r0 += 4) {
        if (r0 >= c.world) break;
        uint4 a[4], b[4];
#pragma unroll
        for (int i = 0; i < 4; ++i)
            if (r0 + i < c.world) { a[i] = ll_load(pk + (r0 + i) * rank_stride); b[i] = ll_load(pk + (r0 + i) * rank_stride + 1); }
#pragma unroll
        for (int i = 0; i < 4; ++i)
            if (r0 + i < c.world) {
                if (!(a[i].y == e && a[i].w == e && b[i].y == e && b[i].w == e)) {
                    const uint4 v = ll_wait_chunk(c, pk + (r0 + i) * rank_stride, e);
                    a[i].x = v.x; a[i].z = v.y; b[i].x = v.z; b[i].z = v.w;
                }
                nv_acc_bf16x8(acc, make_uint4(a[i].x, a[i].z, b[i].x, b[i].z));
            }
    }
}
__device__ __forceinline__ uint4 nv_pack_bf16x8(const float (&a)[8]) {
    return make_uint4(pack_bf16x2(a[0], a[1]), pack_bf16x2(a[2], a[3]), pack_bf16x2(a[4], a[5]), pack_bf16x2(a[6], a[7]));
}

template <bool MC, bool TWO>
__global__ void __launch_bounds__(NV_THREADS) nvls_ll_norm_kernel(
    const NvlsComm c, const uint4* __restrict__ x, uint4* residual, const uint4* __restrict__ weight, float eps,
    uint4* out, int32_t T, int32_t cpr) {
    __shared__ float red[NV_THREADS / 32];
    const uint32_t e = c.state[0] + 1;
    const size_t ll_off = 2 * c.slot_bytes + 128 + (size_t)(e & 1u) * c.ll_slot_bytes;
    const size_t res_off = ll_off + 2 * c.ll_max_bytes + 1024;                 // result area of the two-shot form
    const int64_t n16 = (int64_t)T * cpr;
    const int64_t stride = (int64_t)gridDim.x * NV_THREADS;
    const int64_t gtid = (int64_t)blockIdx.x * NV_THREADS + threadIdx.x;
    const int64_t per = (n16 + c.world - 1) / c.world;

    // keep the barrier flags of the other forms in step (both hold epoch * world after EVERY call)
    if (MC && blockIdx.x == 0 && threadIdx.x == 0) {
        multimem_red_add_u32(c.mc + 2 * c.slot_bytes, 1u);
        multimem_red_add_u32(c.mc + 2 * c.slot_bytes + 64, 1u);
    }

    if (!TWO) {
        uint4* dst = reinterpret_cast<uint4*>((MC ? c.mc : c.local) + ll_off) + (int64_t)c.rank * n16 * 2;
        for (int64_t i = gtid; i < n16; i += stride) {
            const uint4 v = x[i];
            if (MC) { ll_store_mc(dst + 2 * i, v.x, v.y, e); ll_store_mc(dst + 2 * i + 1, v.z, v.w, e); }
            else { ll_store(dst + 2 * i, v.x, v.y, e); ll_store(dst + 2 * i + 1, v.z, v.w, e); }
        }
    } else {
        // reduce-scatter: chunk j of part p goes to rank p, contribution slot [rank]; loads first, then the stores
        for (int64_t j = gtid; j < per; j += stride) {
            uint4 v[NV_MAX_WORLD];
#pragma unroll
            for (int p = 0; p < NV_MAX_WORLD; ++p)
                if (p < c.world && (int64_t)p * per + j < n16) v[p] = x[(int64_t)p * per + j];
#pragma unroll
            for (int p = 0; p < NV_MAX_WORLD; ++p)
                if (p < c.world && (int64_t)p * per + j < n16) {
                    uint4* dst = reinterpret_cast<uint4*>(c.peer[p] + ll_off) + ((int64_t)c.rank * per + j) * 2;
                    ll_store(dst, v[p].x, v[p].y, e);
                    ll_store(dst + 1, v[p].z, v[p].w, e);
                }
        }
        // own part: sum the contributions in rank order, multicast the result packets
        const int64_t lo = (int64_t)c.rank * per;
        const uint4* contrib = reinterpret_cast<const uint4*>(c.local + ll_off);
        uint4* res_dst = reinterpret_cast<uint4*>((MC ? c.mc : c.local) + res_off);
        for (int64_t j = gtid; j < per && lo + j < n16; j += stride) {
            float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
            ll_sum_ranks(c, contrib + j * 2, per * 2, e, acc);
            const uint4 sv = nv_pack_bf16x8(acc);
            if (MC) { ll_store_mc(res_dst + (lo + j) * 2, sv.x, sv.y, e); ll_store_mc(res_dst + (lo + j) * 2 + 1, sv.z, sv.w, e); }
            else { ll_store(res_dst + (lo + j) * 2, sv.x, sv.y, e); ll_store(res_dst + (lo + j) * 2 + 1, sv.z, sv.w, e); }
        }
    }

    for (int32_t row = blockIdx.x; row < T; row += gridDim.x) {
        const int64_t base = (int64_t)row * cpr;
        uint4 sum[NV_MAX_CHUNKS];
#pragma unroll
        for (int kk = 0; kk < NV_MAX_CHUNKS; ++kk) {
            const int cidx = threadIdx.x + kk * NV_THREADS;
            if (cidx < cpr) {
                if (TWO) {
                    sum[kk] = ll_wait_chunk(c, reinterpret_cast<const uint4*>(c.local + res_off) + (base + cidx) * 2, e);
                } else {
                    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
                    ll_sum_ranks(c, reinterpret_cast<const uint4*>(c.local + ll_off) + (base + cidx) * 2, n16 * 2, e, acc);
                    sum[kk] = nv_pack_bf16x8(acc);
                }
            }
        }
        row_epilogue(sum, base, cpr, residual, weight, eps, out, red);
    }

    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        if (atomicAdd(&c.state[1], 1u) == gridDim.x - 1) {
            c.state[1] = 0;
            __threadfence();
            c.state[0] = e;
        }
    }
}

}  // namespace b200

using namespace b200;

extern "C" {

static size_t nv_ll_max(size_t max_msg_bytes) {
    const size_t slot = (max_msg_bytes + 1023) / 1024 * 1024;
    return slot < ((size_t)4 << 20) ? slot : ((size_t)4 << 20);       // packets double the bytes: LL forms stop at 4 MB
}

size_t b200_nvls_buffer_bytes(size_t max_msg_bytes) {
    const size_t slot = (max_msg_bytes + 1023) / 1024 * 1024;
    return 2 * slot + 128 + 2 * (4 * nv_ll_max(max_msg_bytes) + 1024);
}

int b200_nvls_create(int rank, int world, void* local_buf, void* multicast_buf, void* const* peer_bufs, size_t max_msg_bytes,
                     void** comm_out) {
    if (!local_buf || !comm_out) return B200_ERR_NULL;
    if (world < 1 || world > NV_MAX_WORLD || rank < 0 || rank >= world) return B200_ERR_RANGE;
    if (world > 1 && !multicast_buf) return B200_ERR_NULL;
    if (((uintptr_t)local_buf & 15) || ((uintptr_t)multicast_buf & 15)) return B200_ERR_ALIGN;
    NvlsHandle* h = new NvlsHandle();
    NvlsComm* c = &h->dev;
    c->local = (uint8_t*)local_buf; c->mc = (uint8_t*)(multicast_buf ? multicast_buf : local_buf);
    c->rank = rank; c->world = world;
    for (int r = 0; r < NV_MAX_WORLD; ++r) c->peer[r] = (peer_bufs && r < world) ? (uint8_t*)peer_bufs[r] : nullptr;
    c->peer[rank] = c->local;
    c->slot_bytes = (max_msg_bytes + 1023) / 1024 * 1024;
    c->ll_max_bytes = nv_ll_max(max_msg_bytes);
    c->ll_slot_bytes = 4 * c->ll_max_bytes + 1024;
    c->timeout_ns = ar_timeout_ns();
    c->state = nullptr;
    h->has_peers = peer_bufs != nullptr || world == 1;
    // Flags and packet epochs start at zero: every rank zeroes its OWN buffer here (no peer has written into it
    // yet), before the host-side group barrier that must follow creation on all ranks.
    cudaError_t e = cudaMemset(c->local, 0, b200_nvls_buffer_bytes(max_msg_bytes));
    if (e == cudaSuccess) e = cudaMalloc((void**)&c->state, 8 * sizeof(uint32_t));
    if (e == cudaSuccess) e = cudaMemset(c->state, 0, 8 * sizeof(uint32_t));
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    int per_sm[2] = {1 << 30, 1 << 30};
    if (e == cudaSuccess) {
        const void* ks[8] = {(const void*)nvls_ar_norm_kernel<true, true>, (const void*)nvls_ar_norm_kernel<true, false>,
                             (const void*)nvls_ar_norm_kernel<false, true>, (const void*)nvls_ar_norm_kernel<false, false>,
                             (const void*)nvls_ll_norm_kernel<true, true>, (const void*)nvls_ll_norm_kernel<true, false>,
                             (const void*)nvls_ll_norm_kernel<false, true>, (const void*)nvls_ll_norm_kernel<false, false>};
        for (int i = 0; i < 8 && e == cudaSuccess; ++i) {
            int n = 0;
            e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, ks[i], NV_THREADS, 0);
            if (n < per_sm[i / 4]) per_sm[i / 4] = n;
        }
    }
    if (e != cudaSuccess) { if (c->state) cudaFree(c->state); delete h; return (int)e; }
    auto cap = [](int per) { const int64_t n = (int64_t)per * sm_count(); return (int32_t)(n < 1 ? 1 : (n > NV_BLOCKS ? NV_BLOCKS : n)); };
    h->max_blocks = cap(per_sm[0]);
    h->max_blocks_ll = cap(per_sm[1]);
    *comm_out = h;
    return B200_OK;
}

int b200_nvls_destroy(void* comm) {
    if (!comm) return B200_ERR_NULL;
    NvlsHandle* h = (NvlsHandle*)comm;
    int rc = B200_OK;
    if (h->dev.state) { cudaError_t e = cudaFree(h->dev.state); if (e != cudaSuccess) rc = (int)e; }
    delete h;                                   // the symmetric buffer belongs to whoever mapped it (the caller)
    return rc;
}

int b200_nvls_status(void* comm, int32_t* status_out) {
    if (!comm || !status_out) return B200_ERR_NULL;
    const NvlsComm* c = &((const NvlsHandle*)comm)->dev;
    uint32_t v = 0;
    cudaError_t e = cudaMemcpy(&v, c->state + 7, sizeof(v), cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) return (int)e;
    *status_out = (int32_t)v;
    return B200_OK;
}

int b200_nvls_allreduce_add_rmsnorm(void* comm, const void* x, void* residual, const void* weight, float eps, void* out,
                                    int32_t num_tokens, int32_t hidden, int32_t form, void* stream) {
    if (!comm || !x || !out) return B200_ERR_NULL;
    if (residual && !weight) return B200_ERR_NULL;               // residual add exists only in front of the norm
    const NvlsHandle* h = (const NvlsHandle*)comm;
    const NvlsComm* c = &h->dev;
    if (num_tokens < 0 || hidden <= 0 || form < 0 || form > 3) return B200_ERR_RANGE;
    if (num_tokens == 0) return B200_OK;
    if (hidden % 8 != 0 || hidden > NV_THREADS * NV_MAX_CHUNKS * 8) return B200_ERR_SHAPE;
    const size_t bytes = (size_t)num_tokens * hidden * 2;
    if (bytes > c->slot_bytes) return B200_ERR_RANGE;
    if (((uintptr_t)x & 15) || ((uintptr_t)out & 15) || ((uintptr_t)residual & 15) || ((uintptr_t)weight & 15)) return B200_ERR_ALIGN;
    const int32_t cpr = hidden / 8;
    const int64_t n16 = (int64_t)num_tokens * cpr;
    if (form == 2 && bytes * 2 * c->world > c->ll_slot_bytes) return B200_ERR_RANGE;      // world regions of packets per slot
    if (form == 3 && (bytes > c->ll_max_bytes || !h->has_peers)) return B200_ERR_RANGE;
    // one CTA per row in the epilogue; the reduce phases want ~1 chunk per thread.  Same on every rank.
    const bool whole = form == 1 || form == 2;                   // every rank touches every chunk itself
    int64_t grid = num_tokens;
    const int64_t by_chunks = ((whole ? n16 : (n16 + c->world - 1) / c->world) + NV_THREADS - 1) / NV_THREADS;
    if (!whole && by_chunks > grid) grid = by_chunks;
    const int64_t cap = form >= 2 ? h->max_blocks_ll : h->max_blocks;
    grid = grid < 1 ? 1 : (grid > cap ? cap : grid);
    cudaStream_t st = (cudaStream_t)stream;
    const uint4* xp = (const uint4*)x; uint4* rp = (uint4*)residual; const uint4* wp = (const uint4*)weight; uint4* op = (uint4*)out;
    const bool mc = c->world > 1;
    const int g = (int)grid;
#define B200_NV_LAUNCH(K) launch_coop(K, g, NV_THREADS, st, *c, xp, rp, wp, eps, op, num_tokens, cpr)
    switch (form) {
        case 0: return mc ? B200_NV_LAUNCH((nvls_ar_norm_kernel<true, false>)) : B200_NV_LAUNCH((nvls_ar_norm_kernel<false, false>));
        case 1: return mc ? B200_NV_LAUNCH((nvls_ar_norm_kernel<true, true>)) : B200_NV_LAUNCH((nvls_ar_norm_kernel<false, true>));
        case 2: return mc ? B200_NV_LAUNCH((nvls_ll_norm_kernel<true, false>)) : B200_NV_LAUNCH((nvls_ll_norm_kernel<false, false>));
        default: return mc ? B200_NV_LAUNCH((nvls_ll_norm_kernel<true, true>)) : B200_NV_LAUNCH((nvls_ll_norm_kernel<false, true>));
    }
#undef B200_NV_LAUNCH
}

}  // extern "C"
