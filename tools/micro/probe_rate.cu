// Micro-benchmark 2: what bounds scattered reads on B200 -- bytes, DRAM row activations or address translation?
//   (a) random 16-byte reads (.L2::64B) over regions of 64 MB .. 9.6 GB: a TLB-bound rate falls with the region size, an
//       activation-bound rate does not (beyond L2);
//   (b) 1, 2 or 4 independent reads in flight per thread;
//   (c) a SECOND read next to each random one, at +64 B .. +64 KB: a cheap neighbour means row-buffer / page locality pays.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/micro/probe_rate.bin tools/micro/probe_rate.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
typedef long long i64;
typedef unsigned long long u64;

__device__ __forceinline__ u64 mix(u64 x) {
    x ^= x >> 33; x *= 0xff51afd7ed558ccdull; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ull; x ^= x >> 33;
    return x;
}
__device__ __forceinline__ i64 ld(const char *p) {
    i64 a, b;
    asm volatile("ld.global.nc.L2::64B.v2.s64 {%0,%1}, [%2];" : "=l"(a), "=l"(b) : "l"(p));
    return a + b;
}

// UNROLL independent random reads per thread and iteration; PAIR > 0 adds a neighbour read at +PAIR bytes
template <int UNROLL>
__global__ void __launch_bounds__(256) probe(const char *__restrict__ base, u64 n_cells, u64 m, i64 pair, i64 *out) {
    i64 acc = 0;
    const u64 stride = (u64)gridDim.x * blockDim.x;
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < m; i += stride * UNROLL) {
        i64 v[UNROLL], w[UNROLL];
#pragma unroll
        for (int k = 0; k < UNROLL; ++k) {
            const u64 j = i + k * stride;
            const char *p = base + (mix(j) % n_cells) * 64;
            v[k] = j < m ? ld(p) : 0;
            w[k] = (pair && j < m) ? ld(p + pair) : 0;
        }
#pragma unroll
        for (int k = 0; k < UNROLL; ++k) acc += v[k] + w[k];
    }
    if (acc == 0x1234567) out[0] = acc;
}

template <int UNROLL>
static float run(const char *base, u64 n_cells, u64 m, i64 pair, i64 *out, int ctas_per_sm) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    probe<UNROLL><<<148 * ctas_per_sm, 256>>>(base, n_cells, m, pair, out);
    cudaEventRecord(e0);
    probe<UNROLL><<<148 * ctas_per_sm, 256>>>(base, n_cells, m, pair, out);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    return ms * 1e3f;
}

int main() {
    const u64 m = 25000000ull;
    const size_t max_bytes = 9600ull << 20;
    char *base; i64 *out;
    if (cudaMalloc(&base, max_bytes + (1 << 20)) != cudaSuccess) { printf("cudaMalloc failed\n"); return 1; }
    cudaMemset(base, 1, max_bytes + (1 << 20));
    cudaMalloc(&out, 64);
    printf("(a) region size, 1 read in flight per thread, 8 CTAs/SM, %llu reads\n", m);
    const size_t mbs[] = {64, 200, 600, 2400, 9600};
    for (size_t mb : mbs) {
        const u64 cells = (mb << 20) / 64;
        const float us = run<1>(base, cells, m, 0, out, 8);
        printf("  region %5zu MB: %8.1f us  %6.1f G reads/s\n", mb, us, m / us / 1e3);
    }
    const u64 cells = (600ull << 20) / 64;
    printf("(b) reads in flight per thread, 600 MB region\n");
    for (int c : {2, 4, 8}) {
        printf("  %d CTAs/SM: x1 %6.1f  x2 %6.1f  x4 %6.1f G reads/s\n", c, m / run<1>(base, cells, m, 0, out, c) / 1e3,
               m / run<2>(base, cells, m, 0, out, c) / 1e3, m / run<4>(base, cells, m, 0, out, c) / 1e3);
    }
    printf("(c) a neighbour read beside each random one (x2 unroll, 8 CTAs/SM, 600 MB): time relative to no neighbour\n");
    const float t0 = run<2>(base, cells, m, 0, out, 8);
    for (i64 pair : {16ll, 64ll, 128ll, 256ll, 512ll, 1024ll, 2048ll, 4096ll, 65536ll}) {
        const float t = run<2>(base, cells, m, pair, out, 8);
        printf("  +%6lld B: %8.1f us = %.2f x\n", pair, t, t / t0);
    }
    cudaDeviceSynchronize();
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
