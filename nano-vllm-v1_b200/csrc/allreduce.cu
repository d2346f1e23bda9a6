// (f3) Small-message tensor-parallel all-reduce over NVLink peer memory.
//
// Replaces dist.all_reduce(y) after o_proj / down_proj (linear.py:407-410) for decode-sized messages:
// the bf16 partial sums of N ranks ([tokens, hidden], <= B200_AR_MAX_BYTES) are summed in ONE kernel
// launch per rank, no NCCL call, CUDA-graph capturable.  Every rank owns one cudaMalloc'ed buffer that
// its peers map through CUDA IPC (NVSwitch gives every pair full bandwidth, so all traffic is direct
// peer loads):
//
//   [ slot 0 | slot 1 | arrival flags A[world] | arrival flags B[world] ]
//
// One-shot (world == 2 or small messages): every block copies its share of the input into the local slot, all
// blocks of all ranks meet at a barrier, then each block sums its share over all ranks' slots in rank order in
// fp32 and writes bf16 back in place.  Two-shot (larger messages, world > 2): after the same first barrier a
// rank reduces only ITS 1/world of the message, writes the result into its own slot, a second barrier (flags B)
// publishes it, and every rank gathers the reduced parts.  Either way all ranks add in the same order, so they
// hold bit-identical results (every rank must feed identical activations to the next layer).
// The barrier is hierarchical (ar_barrier): blocks of a rank meet on a local counter, block 0 alone exchanges
// flags with the peers, then releases its siblings -- two system-scope fences per barrier per rank.
// Slots alternate with the epoch, so a rank that runs ahead into the next all-reduce never overwrites
// data a slower peer is still reading: it cannot pass that call's barrier before the peer got there.
// The epoch lives in device memory and is advanced by the kernel itself (last block out), so a captured
// graph replays correctly.  Waits are bounded: a protocol bug traps instead of hanging the GPU.
#include <cstring>

#include "../../include/b200attn.h"
#include "launch.h"
#include "common.cuh"

namespace b200 {

constexpr int AR_MAX_WORLD = 8;
constexpr int AR_BLOCKS = 128;                  // most blocks per launch (all must be co-resident: they spin on each other)
constexpr int AR_THREADS = 512;

struct ArComm {
    uint8_t* peer[AR_MAX_WORLD];                // IPC-mapped buffers, peer[rank] is the local one
    int32_t rank, world;
    size_t slot_bytes;
    uint32_t* epoch;                            // device: [0] = completed all-reduces, [1] = blocks done in this launch,
                                                // [2],[3] = arrive counter / go flag of barrier A, [4],[5] = of barrier B,
                                                // [7] = sticky error (a bounded wait timed out)
    uint64_t timeout_ns;                        // bound of every wait (B200ATTN_AR_TIMEOUT_S, default 30 s)
};

// Host-side handle behind the opaque `comm` pointer of the C ABI.
struct ArHandle {
    ArComm dev;                                 // passed to the kernels by value
    int32_t max_blocks;                         // co-resident blocks on this device (<= AR_BLOCKS): all blocks spin on each other
};

// The flag itself travels with relaxed system-scope accesses; ordering comes from ONE system fence before
// the store (release) and ONE after the poll succeeds (acquire).  Acquire loads inside the spin loop would
// pay a fence per iteration.
__device__ __forceinline__ void st_relaxed_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_relaxed_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ uint32_t ld_acquire_gpu(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_gpu(uint32_t* p, uint32_t v) {
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// All blocks of all ranks meet here.  Hierarchical on purpose: a system-scope fence is a GPU-wide operation,
// and one per block (48 blocks: +9 us measured) serialises; so the blocks of a rank first meet on a local
// counter (gpu scope), block 0 alone talks to the peers (two system fences per barrier per rank), then
// releases its siblings through a local flag.  `which` = 0 / 1 selects flag set A / B.
// Bounded wait helper: polls `done()`; after c.timeout_ns of wall clock it raises the communicator's sticky error
// word (epoch[7], read by b200_ar_status) and gives up, so a stalled or dead peer costs a wrong result that the
// host can detect -- not a trapped, unusable CUDA context.
template <typename F>
__device__ __forceinline__ void ar_spin(const ArComm& c, F done) {
    uint64_t t0 = 0;
    for (uint32_t spins = 0; !done(); ++spins) {
        if ((spins & 1023u) == 1023u) {
            if (*reinterpret_cast<volatile uint32_t*>(c.epoch + 7) != 0u) return;    // sticky: one timeout per communicator
            uint64_t now;
            asm volatile("mov.u64 %0, %globaltimer;" : "=l"(now));
            if (t0 == 0) t0 = now;
            else if (now - t0 > c.timeout_ns) { atomicExch(c.epoch + 7, 1u); return; }
        }
    }
}

__device__ __forceinline__ void ar_barrier(const ArComm& c, int which, size_t flags_off, uint32_t epoch) {
    uint32_t* arrive = c.epoch + 2 + 2 * which;
    uint32_t* go = arrive + 1;
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();                        // this block's slot writes before its arrival
        atomicAdd(arrive, 1u);
    }
    if (blockIdx.x == 0) {
        if (threadIdx.x < 32) {
            if (threadIdx.x == 0) {
                ar_spin(c, [&] { return ld_acquire_gpu(arrive) >= gridDim.x; });
                __threadfence_system();         // everything this rank wrote is visible before the flags are:
                for (int r = 0; r < c.world; ++r)   // the SAME thread fences and stores (release pattern)
                    st_relaxed_sys(reinterpret_cast<uint32_t*>(c.peer[r] + flags_off) + c.rank, epoch);
            }
            __syncwarp();
            if ((int)threadIdx.x < c.world) {
                const uint32_t* mine = reinterpret_cast<const uint32_t*>(c.peer[c.rank] + flags_off) + threadIdx.x;
                ar_spin(c, [&] { return (int32_t)(ld_relaxed_sys(mine) - epoch) >= 0; });
                __threadfence_system();         // acquire side, in the lane that observed the flag
            }
            __syncwarp();
            if (threadIdx.x == 0) {
                __threadfence_system();
                st_release_gpu(go, epoch);
            }
        }
    }
    if (threadIdx.x == 0) ar_spin(c, [&] { return (int32_t)(ld_acquire_gpu(go) - epoch) >= 0; });
    __syncthreads();
}

__device__ __forceinline__ uint4 ld_cg_u4(const void* p) { return __ldcg(reinterpret_cast<const uint4*>(p)); }

__device__ __forceinline__ void acc_bf16x8(float (&a)[8], const uint4& v) {
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const float2 f = unpack_bf16x2(w[i]);
        a[2 * i] += f.x; a[2 * i + 1] += f.y;
    }
}
__device__ __forceinline__ uint4 pack_bf16x8(const float (&a)[8]) {
    return make_uint4(pack_bf16x2(a[0], a[1]), pack_bf16x2(a[2], a[3]), pack_bf16x2(a[4], a[5]), pack_bf16x2(a[6], a[7]));
}

// n16 = message size in 16-byte chunks (8 bf16).  In place on `data`.
template <bool TWO_SHOT>
__global__ void __launch_bounds__(AR_THREADS) allreduce_kernel(const ArComm c, uint4* data, int64_t n16) {
    const uint32_t epoch = c.epoch[0] + 1;
    const size_t slot_off = (epoch & 1u) * c.slot_bytes;
    const size_t flags_a = 2 * c.slot_bytes, flags_b = flags_a + sizeof(uint32_t) * AR_MAX_WORLD;
    uint4* my_slot = reinterpret_cast<uint4*>(c.peer[c.rank] + slot_off);
    const int64_t stride = (int64_t)gridDim.x * AR_THREADS;
    const int64_t tid = (int64_t)blockIdx.x * AR_THREADS + threadIdx.x;

    if (!TWO_SHOT) {
        for (int64_t i = tid; i < n16; i += stride) my_slot[i] = data[i];
        ar_barrier(c, 0, flags_a, epoch);
        int64_t i = tid;
        for (; i + 3 * stride < n16; i += 4 * stride) {         // 4 x world independent 16-byte peer loads in flight
            uint4 v[4][AR_MAX_WORLD];
#pragma unroll
            for (int r = 0; r < AR_MAX_WORLD; ++r)
                if (r < c.world) {
                    const uint4* src = reinterpret_cast<const uint4*>(c.peer[r] + slot_off) + i;
#pragma unroll
                    for (int k = 0; k < 4; ++k) v[k][r] = ld_cg_u4(src + k * stride);
                }
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                float a[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
                for (int r = 0; r < AR_MAX_WORLD; ++r)
                    if (r < c.world) acc_bf16x8(a, v[k][r]);
                data[i + k * stride] = pack_bf16x8(a);
            }
        }
        for (; i < n16; i += stride) {
            float a[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
            for (int r = 0; r < c.world; ++r) acc_bf16x8(a, ld_cg_u4(reinterpret_cast<const uint4*>(c.peer[r] + slot_off) + i));
            data[i] = pack_bf16x8(a);
        }
    } else {
        // part p of the message = chunks [p*per, min((p+1)*per, n16)); block b owns {tid + k*stride} inside every part
        const int64_t per = (n16 + c.world - 1) / c.world;
        for (int p = 0; p < c.world; ++p) {
            const int64_t plo = (int64_t)p * per, phi = min(plo + per, n16);
            for (int64_t i = plo + tid; i < phi; i += stride) my_slot[i] = data[i];
        }
        ar_barrier(c, 0, flags_a, epoch);
        const int64_t lo = (int64_t)c.rank * per, hi = min(lo + per, n16);
        for (int64_t i = lo + tid; i < hi; i += stride) {
            float a[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
            for (int r = 0; r < c.world; ++r) acc_bf16x8(a, ld_cg_u4(reinterpret_cast<const uint4*>(c.peer[r] + slot_off) + i));
            const uint4 v = pack_bf16x8(a);
            my_slot[i] = v;                     // peers read the reduced part from here after barrier B
            data[i] = v;
        }
        ar_barrier(c, 1, flags_b, epoch);
        // gather: index i of my own part maps to i - lo + p*per in part p; all peers' loads of one step are
        // issued before the first store (world - 1 independent round trips in flight per thread)
        for (int64_t j = tid; j < per; j += stride) {
            uint4 v[AR_MAX_WORLD];
#pragma unroll
            for (int p = 0; p < AR_MAX_WORLD; ++p)
                if (p < c.world && p != c.rank && (int64_t)p * per + j < n16)
                    v[p] = ld_cg_u4(reinterpret_cast<const uint4*>(c.peer[p] + slot_off) + (int64_t)p * per + j);
#pragma unroll
            for (int p = 0; p < AR_MAX_WORLD; ++p)
                if (p < c.world && p != c.rank && (int64_t)p * per + j < n16) data[(int64_t)p * per + j] = v[p];
        }
    }
    // last block out advances the epoch for the next launch (graph replays read it from device memory)
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        if (atomicAdd(&c.epoch[1], 1u) == gridDim.x - 1) {
            c.epoch[1] = 0;
            c.epoch[2] = 0;                     // arrive counters of both barriers, for the next launch
            c.epoch[4] = 0;
            __threadfence();
            c.epoch[0] = epoch;
        }
    }
}

// Two-shot, PUSH form (two_shot == 2).  Remote stores are posted, remote loads are round trips: instead of
// parking the input locally and letting the peers pull, a rank scatters part p of its input straight into
// rank p's slot, reduces its own part from LOCAL memory after barrier A, broadcasts the reduced part into every
// rank's result region, and after barrier B copies the result region out locally.
// slot = [ contributions: world x per chunks | results: n16 chunks ]   (slot_bytes = 2 x the largest message)
__global__ void __launch_bounds__(AR_THREADS) allreduce_push_kernel(const ArComm c, uint4* data, int64_t n16) {
    const uint32_t epoch = c.epoch[0] + 1;
    const size_t slot_off = (epoch & 1u) * c.slot_bytes;
    const size_t flags_a = 2 * c.slot_bytes, flags_b = flags_a + sizeof(uint32_t) * AR_MAX_WORLD;
    const int64_t stride = (int64_t)gridDim.x * AR_THREADS;
    const int64_t tid = (int64_t)blockIdx.x * AR_THREADS + threadIdx.x;
    const int64_t per = (n16 + c.world - 1) / c.world;
    const size_t res_off = slot_off + c.slot_bytes / 2;

    for (int64_t j = tid; j < per; j += stride) {
        uint4 v[AR_MAX_WORLD];
#pragma unroll
        for (int p = 0; p < AR_MAX_WORLD; ++p)
            if (p < c.world && (int64_t)p * per + j < n16) v[p] = data[(int64_t)p * per + j];
#pragma unroll
        for (int p = 0; p < AR_MAX_WORLD; ++p)
            if (p < c.world && (int64_t)p * per + j < n16)
                reinterpret_cast<uint4*>(c.peer[p] + slot_off)[(int64_t)c.rank * per + j] = v[p];
    }
    ar_barrier(c, 0, flags_a, epoch);
    const uint4* contrib = reinterpret_cast<const uint4*>(c.peer[c.rank] + slot_off);
    for (int64_t j = tid; j < per && (int64_t)c.rank * per + j < n16; j += stride) {
        float a[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int r = 0; r < AR_MAX_WORLD; ++r)
            if (r < c.world) acc_bf16x8(a, ld_cg_u4(contrib + (int64_t)r * per + j));
        const uint4 v = pack_bf16x8(a);
#pragma unroll
        for (int q = 0; q < AR_MAX_WORLD; ++q)
            if (q < c.world) reinterpret_cast<uint4*>(c.peer[q] + res_off)[(int64_t)c.rank * per + j] = v;
    }
    ar_barrier(c, 1, flags_b, epoch);
    const uint4* res = reinterpret_cast<const uint4*>(c.peer[c.rank] + res_off);
    for (int64_t i = tid; i < n16; i += stride) data[i] = ld_cg_u4(res + i);
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        if (atomicAdd(&c.epoch[1], 1u) == gridDim.x - 1) {
            c.epoch[1] = 0;
            c.epoch[2] = 0;
            c.epoch[4] = 0;
            __threadfence();
            c.epoch[0] = epoch;
        }
    }
}

// One-shot, PUSH form (two_shot == 3): every rank stores its whole message into slot region [rank] of every
// peer (posted stores), ONE barrier, then each rank sums its `world` local copies.  No remote loads at all;
// needs world x message <= slot.
__global__ void __launch_bounds__(AR_THREADS) allreduce_push1_kernel(const ArComm c, uint4* data, int64_t n16) {
    const uint32_t epoch = c.epoch[0] + 1;
    const size_t slot_off = (epoch & 1u) * c.slot_bytes;
    const size_t flags_a = 2 * c.slot_bytes;
    const int64_t stride = (int64_t)gridDim.x * AR_THREADS;
    const int64_t tid = (int64_t)blockIdx.x * AR_THREADS + threadIdx.x;
    for (int64_t i = tid; i < n16; i += stride) {
        const uint4 v = data[i];
#pragma unroll
        for (int q = 0; q < AR_MAX_WORLD; ++q)
            if (q < c.world) reinterpret_cast<uint4*>(c.peer[q] + slot_off)[(int64_t)c.rank * n16 + i] = v;
    }
    ar_barrier(c, 0, flags_a, epoch);
    const uint4* mine = reinterpret_cast<const uint4*>(c.peer[c.rank] + slot_off);
    for (int64_t i = tid; i < n16; i += stride) {
        float a[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int r = 0; r < AR_MAX_WORLD; ++r)
            if (r < c.world) acc_bf16x8(a, ld_cg_u4(mine + (int64_t)r * n16 + i));
        data[i] = pack_bf16x8(a);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        if (atomicAdd(&c.epoch[1], 1u) == gridDim.x - 1) {
            c.epoch[1] = 0;
            c.epoch[2] = 0;
            c.epoch[4] = 0;
            __threadfence();
            c.epoch[0] = epoch;
        }
    }
}

uint64_t ar_timeout_ns() {
    const char* env = getenv("B200ATTN_AR_TIMEOUT_S");
    double sec = env ? atof(env) : 30.0;
    if (!(sec > 0.0)) sec = 30.0;
    return (uint64_t)(sec * 1e9);
}

}  // namespace b200

using namespace b200;

extern "C" {

size_t b200_ar_buffer_bytes(size_t max_msg_bytes) {
    const size_t slot = 2 * ((max_msg_bytes + 1023) / 1024 * 1024);   // contributions | results (push form)
    return 2 * slot + 2 * sizeof(uint32_t) * AR_MAX_WORLD;
}

int b200_ar_alloc(size_t bytes, void** dev_ptr, void* ipc_handle_out64) {
    if (!dev_ptr || !ipc_handle_out64) return B200_ERR_NULL;
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle is 64 bytes");
    void* p = nullptr;
    cudaError_t e = cudaMalloc(&p, bytes);
    if (e != cudaSuccess) return (int)e;
    e = cudaMemset(p, 0, bytes);
    if (e != cudaSuccess) return (int)e;
    e = cudaDeviceSynchronize();
    if (e != cudaSuccess) return (int)e;
    cudaIpcMemHandle_t h;
    e = cudaIpcGetMemHandle(&h, p);
    if (e != cudaSuccess) return (int)e;
    memcpy(ipc_handle_out64, &h, 64);
    *dev_ptr = p;
    return B200_OK;
}

int b200_ar_open(const void* ipc_handle64, void** peer_ptr) {
    if (!ipc_handle64 || !peer_ptr) return B200_ERR_NULL;
    cudaIpcMemHandle_t h;
    memcpy(&h, ipc_handle64, 64);
    return (int)cudaIpcOpenMemHandle(peer_ptr, h, cudaIpcMemLazyEnablePeerAccess);
}

int b200_ar_create(int rank, int world, void* const* peer_ptrs, size_t max_msg_bytes, void** comm_out) {
    if (!peer_ptrs || !comm_out) return B200_ERR_NULL;
    if (world < 1 || world > AR_MAX_WORLD || rank < 0 || rank >= world) return B200_ERR_RANGE;
    ArHandle* h = new ArHandle();
    ArComm* c = &h->dev;
    for (int r = 0; r < AR_MAX_WORLD; ++r) c->peer[r] = r < world ? (uint8_t*)peer_ptrs[r] : nullptr;
    c->rank = rank; c->world = world;
    c->slot_bytes = 2 * ((max_msg_bytes + 1023) / 1024 * 1024);
    c->timeout_ns = ar_timeout_ns();
    cudaError_t e = cudaMalloc((void**)&c->epoch, 8 * sizeof(uint32_t));
    if (e == cudaSuccess) e = cudaMemset(c->epoch, 0, 8 * sizeof(uint32_t));
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    // co-residency bound: the smallest occupancy of the four kernels x SM count
    int per_sm = 1 << 30;
    if (e == cudaSuccess) {
        const void* ks[4] = {(const void*)allreduce_kernel<false>, (const void*)allreduce_kernel<true>,
                             (const void*)allreduce_push_kernel, (const void*)allreduce_push1_kernel};
        for (int i = 0; i < 4 && e == cudaSuccess; ++i) {
            int n = 0;
            e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, ks[i], AR_THREADS, 0);
            if (n < per_sm) per_sm = n;
        }
    }
    if (e != cudaSuccess) { if (c->epoch) cudaFree(c->epoch); delete h; return (int)e; }
    int64_t cap = (int64_t)per_sm * sm_count();
    h->max_blocks = (int32_t)(cap < 1 ? 1 : (cap > AR_BLOCKS ? AR_BLOCKS : cap));
    *comm_out = h;
    return B200_OK;
}

// Tear-down (the reference destroys its process group at exit, model_runner.py:187-193): closes the IPC
// mappings of the peers' buffers, frees the local buffer (it must come from b200_ar_alloc) and the
// communicator.  The caller synchronises the device and the group first (no rank may still be inside a call).
int b200_ar_destroy(void* comm) {
    if (!comm) return B200_ERR_NULL;
    ArHandle* h = (ArHandle*)comm;
    ArComm* c = &h->dev;
    int rc = B200_OK;
    for (int r = 0; r < c->world; ++r) {
        if (!c->peer[r]) continue;
        cudaError_t e = (r == c->rank) ? cudaFree(c->peer[r]) : cudaIpcCloseMemHandle(c->peer[r]);
        if (e != cudaSuccess && rc == B200_OK) rc = (int)e;
    }
    if (c->epoch) { cudaError_t e = cudaFree(c->epoch); if (e != cudaSuccess && rc == B200_OK) rc = (int)e; }
    delete h;
    return rc;
}

// 0 = healthy; 1 = some wait inside a kernel exceeded B200ATTN_AR_TIMEOUT_S since the communicator was made
// (results after that point are not trustworthy).  Synchronises the device: for tests, tools and shutdown.
int b200_ar_status(void* comm, int32_t* status_out) {
    if (!comm || !status_out) return B200_ERR_NULL;
    const ArComm* c = &((const ArHandle*)comm)->dev;
    uint32_t v = 0;
    cudaError_t e = cudaMemcpy(&v, c->epoch + 7, sizeof(v), cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) return (int)e;
    *status_out = (int32_t)v;
    return B200_OK;
}

int b200_allreduce_bf16(void* comm, void* data, int64_t n_elems, int two_shot, void* stream) {
    if (!comm || !data) return B200_ERR_NULL;
    const ArHandle* h = (const ArHandle*)comm;
    const ArComm* c = &h->dev;
    if (n_elems < 0 || (n_elems & 7) || (size_t)n_elems * 2 > c->slot_bytes / 2) return B200_ERR_RANGE;
    if (((uintptr_t)data & 15) != 0) return B200_ERR_ALIGN;
    if (n_elems == 0) return B200_OK;
    // About four 16-byte chunks per thread in the phase that reads the peers (memory-level parallelism comes
    // from the unrolled loads, not from long loops); the grid depends only on n_elems and the group size, so it
    // is the same on every rank.  All blocks spin on each other: the grid is capped at the co-resident count and
    // launched cooperatively.
    const bool two = (two_shot == 1 || two_shot == 2) && c->world > 2;
    // push form: world x ceil(n16 / world) chunks of contributions must stay inside the first half of a slot
    // (only the largest message with a group size that does not divide it can miss by world - 1 chunks)
    if (two && two_shot == 2 && (size_t)((n_elems / 8 + c->world - 1) / c->world) * c->world * 16 > c->slot_bytes / 2) two_shot = 1;
    const int64_t work = two ? (n_elems / 8 + c->world - 1) / c->world : (n_elems / 8 + 3) / 4;   // chunks per thread-step
    int64_t grid = (work + AR_THREADS - 1) / AR_THREADS;
    grid = grid < 1 ? 1 : (grid > h->max_blocks ? h->max_blocks : grid);
    cudaStream_t st = (cudaStream_t)stream;
    uint4* d = (uint4*)data;
    const int64_t n16 = n_elems / 8;
    if (two_shot == 3) {
        if ((size_t)n_elems * 2 * c->world > c->slot_bytes) return B200_ERR_RANGE;
        return launch_coop(allreduce_push1_kernel, (int)grid, AR_THREADS, st, *c, d, n16);
    }
    if (two && two_shot == 2) return launch_coop(allreduce_push_kernel, (int)grid, AR_THREADS, st, *c, d, n16);
    if (two) return launch_coop(allreduce_kernel<true>, (int)grid, AR_THREADS, st, *c, d, n16);
    return launch_coop(allreduce_kernel<false>, (int)grid, AR_THREADS, st, *c, d, n16);
}

}  // extern "C"
