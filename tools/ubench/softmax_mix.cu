// Micro-benchmark: how many cycles does ONE warp (or two warps sharing an SM sub-partition) need for the
// per-tile softmax instruction mix of the prefill kernel (128 scores per thread)?  Separates the cost of
// MUFU.EX2, FMNMX3, F2FP and FADD2/FFMA2.  Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint64_t pack2(float lo, float hi) { uint64_t r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void unpack2(uint64_t v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ uint64_t fma2(uint64_t a, uint64_t b, uint64_t c) { uint64_t d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ uint64_t add2(uint64_t a, uint64_t b) { uint64_t d; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ float max3(float a, float b, float c) { float d; asm("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c)); return d; }
__device__ __forceinline__ float ex2(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ uint32_t cvt2(float lo, float hi) { uint32_t r; asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo)); return r; }

// MODE bits: 1 = max3 pass, 2 = fma2, 4 = ex2, 8 = add2 row sum, 16 = cvt pack, 32 = plain 2-input max pass
template <int MODE>
__global__ void mix(float* out, long long* cycles, int iters, float seed) {
    float s[128];
#pragma unroll
    for (int i = 0; i < 128; ++i) s[i] = seed * (float)(i + threadIdx.x) * 1e-3f;
    uint64_t l2 = pack2(0.f, 0.f), l3 = pack2(0.f, 0.f);
    uint32_t acc = 0;
    float m = -1e30f;
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
        if (MODE & 1) {
            float mc[8];
#pragma unroll
            for (int c = 0; c < 8; ++c) mc[c] = m;
#pragma unroll
            for (int i = 0; i < 128; i += 16)
#pragma unroll
                for (int c = 0; c < 8; ++c) mc[c] = max3(mc[c], s[i + 2 * c], s[i + 2 * c + 1]);
            m = max3(max3(mc[0], mc[1], mc[2]), max3(mc[3], mc[4], mc[5]), fmaxf(mc[6], mc[7]));
        }
        if (MODE & 32) {
            float mc[8];
#pragma unroll
            for (int c = 0; c < 8; ++c) mc[c] = m;
#pragma unroll
            for (int i = 0; i < 128; i += 8)
#pragma unroll
                for (int c = 0; c < 8; ++c) mc[c] = fmaxf(mc[c], s[i + c]);
            m = fmaxf(fmaxf(fmaxf(mc[0], mc[1]), fmaxf(mc[2], mc[3])), fmaxf(fmaxf(mc[4], mc[5]), fmaxf(mc[6], mc[7])));
        }
        const uint64_t sc = pack2(0.999f, 0.999f), nm = pack2(-m * 1e-9f, -m * 1e-9f);
#pragma unroll
        for (int i = 0; i < 64; ++i) {
            float x0 = s[2 * i], x1 = s[2 * i + 1];
            if (MODE & 2) unpack2(fma2(pack2(x0, x1), sc, nm), x0, x1);
            if (MODE & 4) { x0 = ex2(x0); x1 = ex2(x1); }
            if (MODE & 8) { if (i & 1) l3 = add2(l3, pack2(x0, x1)); else l2 = add2(l2, pack2(x0, x1)); }
            if (MODE & 16) acc ^= cvt2(x0, x1);
            s[2 * i] = x0; s[2 * i + 1] = x1;
        }
    }
    const long long t1 = clock64();
    float a, b, c, d;
    unpack2(l2, a, b); unpack2(l3, c, d);
    float r = a + b + c + d + m + __uint_as_float(acc);
#pragma unroll
    for (int i = 0; i < 128; ++i) r += s[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
    if (threadIdx.x == 0 && blockIdx.x == 0) cycles[0] = t1 - t0;
}

template <int MODE>
void run(const char* name, int threads) {
    float* out; long long* cyc;
    cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 8);
    const int iters = 200;
    mix<MODE><<<148, threads>>>(out, cyc, iters, 1.0f);
    mix<MODE><<<148, threads>>>(out, cyc, iters, 1.0f);
    cudaDeviceSynchronize();
    long long h = 0; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    printf("%-34s warps/SMSP %d: %7.1f cycles per 128-score row-tile  (%s)\n", name, threads / 128, (double)h / iters, cudaGetErrorString(cudaGetLastError()));
    cudaFree(out); cudaFree(cyc);
}

int main() {
    for (int threads : {128, 256}) {
        run<4>("ex2 only", threads);
        run<2 | 4>("fma2+ex2", threads);
        run<2 | 4 | 8>("fma2+ex2+add2", threads);
        run<2 | 4 | 16>("fma2+ex2+cvt", threads);
        run<2 | 4 | 8 | 16>("fma2+ex2+add2+cvt", threads);
        run<1>("max3 only", threads);
        run<32>("max (2-input) only", threads);
        run<1 | 2 | 4 | 8 | 16>("full: max3+fma2+ex2+add2+cvt", threads);
        run<32 | 2 | 4 | 8 | 16>("full with 2-input max", threads);
        run<2 | 8 | 16>("fma2+add2+cvt (no ex2)", threads);
    }
    return 0;
}
