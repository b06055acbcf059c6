// Micro-benchmark: how many DRAM bytes does ONE scattered 16/32-byte read cost on B200, by load flavour?
// 12.5 M random 32-byte-aligned reads over a 600 MB array (nothing is reused, nothing fits L2).  Run under
//   ncu --metrics dram__bytes_read.sum,lts__t_sectors_srcunit_tex_op_read.sum,gpu__time_duration.sum
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/micro/probe_gran.bin tools/micro/probe_gran.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
typedef long long i64;
typedef unsigned long long u64;

__device__ __forceinline__ u64 mix(u64 x) {
    x ^= x >> 33; x *= 0xff51afd7ed558ccdull; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ull; x ^= x >> 33;
    return x;
}

template <int MODE>
__global__ void __launch_bounds__(256) probe(const char *__restrict__ base, u64 n_cells, u64 m, i64 *out) {
    __shared__ __align__(16) i64 sh[256][2];
    i64 acc = 0;
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < m; i += (u64)gridDim.x * blockDim.x) {
        const char *p = base + (mix(i) % n_cells) * 32;
        i64 a = 0, b = 0, c = 0, d = 0;
        if (MODE == 0) asm volatile("ld.global.nc.v2.s64 {%0,%1}, [%2];" : "=l"(a), "=l"(b) : "l"(p));
        if (MODE == 1) asm volatile("ld.global.v2.s64 {%0,%1}, [%2];" : "=l"(a), "=l"(b) : "l"(p));
        if (MODE == 2) asm volatile("ld.global.cg.v2.s64 {%0,%1}, [%2];" : "=l"(a), "=l"(b) : "l"(p));
        if (MODE == 3) asm volatile("ld.global.nc.L1::no_allocate.v2.s64 {%0,%1}, [%2];" : "=l"(a), "=l"(b) : "l"(p));
        if (MODE == 4) asm volatile("ld.global.nc.v4.b64 {%0,%1,%2,%3}, [%4];" : "=l"(a), "=l"(b), "=l"(c), "=l"(d) : "l"(p));
        if (MODE == 5) asm volatile("ld.global.cv.v2.s64 {%0,%1}, [%2];" : "=l"(a), "=l"(b) : "l"(p));
        if (MODE == 6) asm volatile("ld.global.L1::no_allocate.L2::evict_first.v4.b64 {%0,%1,%2,%3}, [%4];" : "=l"(a), "=l"(b), "=l"(c), "=l"(d) : "l"(p));
        if (MODE == 7) {
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((unsigned)__cvta_generic_to_shared(&sh[threadIdx.x][0])), "l"(p) : "memory");
            asm volatile("cp.async.commit_group;" ::: "memory");
            asm volatile("cp.async.wait_group 0;" ::: "memory");
            a = sh[threadIdx.x][0]; b = sh[threadIdx.x][1];
        }
        if (MODE == 8) asm volatile("ld.global.nc.s64 %0, [%1];" : "=l"(a) : "l"(p));
        if (MODE == 9) asm volatile("ld.global.nc.L2::64B.v2.s64 {%0,%1}, [%2];" : "=l"(a), "=l"(b) : "l"(p));
        acc += a + b + c + d;
    }
    if (acc == 0x1234567) out[0] = acc;
}

template <int MODE>
static void run(const char *name, const char *base, u64 n_cells, u64 m, i64 *out) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    probe<MODE><<<148 * 8, 256>>>(base, n_cells, m, out);
    cudaEventRecord(e0);
    probe<MODE><<<148 * 8, 256>>>(base, n_cells, m, out);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    printf("mode %d %-44s %8.1f us  %s\n", MODE, name, ms * 1e3, cudaGetErrorString(cudaGetLastError()));
}

int main(int argc, char **argv) {
    const int gran = argc > 1 ? atoi(argv[1]) : 0;
    if (gran > 0) printf("cudaLimitMaxL2FetchGranularity=%d -> %s\n", gran, cudaGetErrorString(cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, gran)));
    size_t g = 0;
    cudaDeviceGetLimit(&g, cudaLimitMaxL2FetchGranularity);
    printf("L2 fetch granularity limit: %zu\n", g);
    const u64 n_cells = 18750000ull, m = 12500000ull;
    char *base; i64 *out;
    cudaMalloc(&base, n_cells * 32);
    cudaMemset(base, 1, n_cells * 32);
    cudaMalloc(&out, 64);
    run<0>("ld.global.nc.v2.s64 (16 B)", base, n_cells, m, out);
    run<1>("ld.global.v2.s64", base, n_cells, m, out);
    run<2>("ld.global.cg.v2.s64", base, n_cells, m, out);
    run<3>("ld.global.nc.L1::no_allocate.v2.s64", base, n_cells, m, out);
    run<4>("ld.global.nc.v4.b64 (32 B, LDG.256)", base, n_cells, m, out);
    run<5>("ld.global.cv.v2.s64", base, n_cells, m, out);
    run<6>("ld.global.L1::no_allocate.L2::evict_first.v4.b64", base, n_cells, m, out);
    run<7>("cp.async.cg 16 B -> shared", base, n_cells, m, out);
    run<8>("ld.global.nc.s64 (8 B)", base, n_cells, m, out);
    run<9>("ld.global.nc.L2::64B.v2.s64", base, n_cells, m, out);
    cudaDeviceSynchronize();
    return 0;
}
