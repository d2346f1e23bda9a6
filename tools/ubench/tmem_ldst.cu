// Micro-benchmark: latency / throughput of tcgen05.ld and tcgen05.st (32x32b) as one warp per SM sub-partition
// (128 threads) or two (256 threads: warps w and w+4 touch the same TMEM lanes, different columns) see them.
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tmem_ldst.bin tmem_ldst.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define LD32(addr, r) asm volatile( \
    "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];" \
    : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), \
      "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31]) : "r"(addr))
#define ST32(addr, r) asm volatile( \
    "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};" \
    :: "r"(addr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), \
      "r"(r[16]), "r"(r[17]), "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31]) : "memory")

// MODE 0: NLD x (ld.x32) then wait::ld, per iteration.  MODE 1: NLD x st.x32 then wait::st.  MODE 2: ld, +1, st back.
template <int MODE, int NLD>
__global__ void k(long long* cycles, uint32_t* sink, int iters) {
    __shared__ uint32_t tmem_ptr;
    const int warp = threadIdx.x >> 5;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"((uint32_t)__cvta_generic_to_shared(&tmem_ptr)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t base = tmem_ptr + ((uint32_t)((warp & 3) * 32) << 16) + (warp >> 2) * 128;
    uint32_t r[4][32];
#pragma unroll
    for (int c = 0; c < 4; ++c)
#pragma unroll
        for (int i = 0; i < 32; ++i) r[c][i] = threadIdx.x + i + c;
#pragma unroll
    for (int c = 0; c < 4; ++c) ST32(base + c * 32, r[c]);
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    uint32_t acc = 0;
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
        if (MODE == 0 || MODE == 2) {
#pragma unroll
            for (int c = 0; c < NLD; ++c) LD32(base + c * 32, r[c]);
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
            for (int c = 0; c < NLD; ++c) acc += r[c][0] ^ r[c][31];        // dependent uses of the first and last register of each load
        }
        if (MODE == 3) {                                          // the prefill softmax's load + row-maximum phase
#pragma unroll
            for (int c = 0; c < 4; ++c) LD32(base + c * 32, r[c]);
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            float mc[8];
#pragma unroll
            for (int c = 0; c < 8; ++c) mc[c] = __uint_as_float(acc);
#pragma unroll
            for (int c = 0; c < 4; ++c)
#pragma unroll
                for (int i = 0; i < 32; i += 16)
#pragma unroll
                    for (int q = 0; q < 8; ++q)
                        asm("max.f32 %0, %0, %1, %2;" : "+f"(mc[q]) : "f"(__uint_as_float(r[c][i + 2 * q])), "f"(__uint_as_float(r[c][i + 2 * q + 1])));
            float m;
            asm("max.f32 %0, %1, %2, %3;" : "=f"(m) : "f"(mc[0]), "f"(mc[1]), "f"(mc[2]));
            asm("max.f32 %0, %0, %1, %2;" : "+f"(m) : "f"(mc[3]), "f"(mc[4]));
            asm("max.f32 %0, %0, %1, %2;" : "+f"(m) : "f"(mc[5]), "f"(mc[6]));
            acc = __float_as_uint(fmaxf(m, mc[7])) & 0x7fffffu;
        }
        if (MODE == 1 || MODE == 2) {
#pragma unroll
            for (int c = 0; c < NLD; ++c) { r[c][0] += acc; ST32(base + c * 32, r[c]); }
            asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
        }
    }
    const long long t1 = clock64();
    sink[blockIdx.x * blockDim.x + threadIdx.x] = acc + r[0][1];
    if (threadIdx.x == 0 && blockIdx.x == 0) cycles[0] = t1 - t0;
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem_ptr) : "memory");
}

template <int MODE, int NLD>
void run(const char* name, int threads) {
    long long* cyc; uint32_t* sink;
    cudaMalloc(&cyc, 8); cudaMalloc(&sink, 148 * 256 * 4);
    const int iters = 1000;
    k<MODE, NLD><<<148, threads>>>(cyc, sink, iters);
    k<MODE, NLD><<<148, threads>>>(cyc, sink, iters);
    cudaDeviceSynchronize();
    long long h = 0; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    const double per = (double)h / iters;
    printf("%-28s x%d (%5d B/warp) warps/SMSP %d: %7.1f cycles per round, %6.1f B/clk per SMSP  (%s)\n", name, NLD, NLD * 4096,
           threads / 128, per, (double)NLD * 4096 * (threads / 128) * (MODE == 2 ? 2 : 1) / per, cudaGetErrorString(cudaGetLastError()));
    cudaFree(cyc); cudaFree(sink);
}

int main() {
    for (int threads : {128, 256}) {
        run<0, 1>("ld.x32 + wait", threads);
        run<0, 2>("ld.x32 + wait", threads);
        run<0, 4>("ld.x32 + wait", threads);
        run<1, 1>("st.x32 + wait", threads);
        run<1, 2>("st.x32 + wait", threads);
        run<1, 4>("st.x32 + wait", threads);
        run<2, 4>("ld + st round trip", threads);
        run<3, 4>("ld x4 + wait + 67 max3", threads);
    }
    return 0;
}
