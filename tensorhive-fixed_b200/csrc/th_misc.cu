// Stand-alone K1 entry point and the device-side synthetic workload generators used by bench.py and the
// full-size property tests.  Nothing here is on a reference call path.
#include "th_common.cuh"

namespace {

struct PairStats { i64 min_lo, max_lo, min_hi, max_hi; };

__global__ void __launch_bounds__(256) sp_stats_kernel(const i64 *__restrict__ hi, const i64 *__restrict__ lo, i64 n,
                                                       PairStats *out) {
    i64 mnl = TH_I64_MAX, mxl = TH_I64_MIN, mnh = TH_I64_MAX, mxh = TH_I64_MIN;
    for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (i64)gridDim.x * blockDim.x) {
        const i64 l = lo[i], h = hi[i];
        mnl = l < mnl ? l : mnl; mxl = l > mxl ? l : mxl;
        mnh = h < mnh ? h : mnh; mxh = h > mxh ? h : mxh;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        i64 t;
        t = __shfl_xor_sync(0xffffffffu, mnl, o); mnl = t < mnl ? t : mnl;
        t = __shfl_xor_sync(0xffffffffu, mxl, o); mxl = t > mxl ? t : mxl;
        t = __shfl_xor_sync(0xffffffffu, mnh, o); mnh = t < mnh ? t : mnh;
        t = __shfl_xor_sync(0xffffffffu, mxh, o); mxh = t > mxh ? t : mxh;
    }
    if ((threadIdx.x & 31) == 0) {
        atomicMin(&out->min_lo, mnl); atomicMax(&out->max_lo, mxl);
        atomicMin(&out->min_hi, mnh); atomicMax(&out->max_hi, mxh);
    }
}

__global__ void __launch_bounds__(256) sp_rec_kernel(const i64 *__restrict__ hi, const i64 *__restrict__ lo, i64 n,
                                                     i64 min_lo, i64 min_hi, ThRec *rec) {
    for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (i64)gridDim.x * blockDim.x) {
        ThRec r;
        r.k0 = (u64)lo[i] - (u64)min_lo; r.k1 = (u32)((u64)hi[i] - (u64)min_hi); r.idx = (u32)i;
        rec[i] = r;
    }
}

__global__ void __launch_bounds__(256) sp_perm_kernel(const ThRec *__restrict__ rec, i64 n, i64 *perm) {
    for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (i64)gridDim.x * blockDim.x)
        perm[i] = rec[i].idx;
}

// splitmix64: counter-based, so every element is a pure function of (seed, index)
__device__ __forceinline__ u64 sx_mix(u64 x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}

__global__ void __launch_bounds__(256) sx_samples_kernel(i64 n_uuids, i64 spu, i64 t0, u64 seed, i64 *off, i64 *ts,
                                                         i64 *gpu, i64 *mem) {
    const i64 total = n_uuids * spu;
    const i64 stride = (i64)gridDim.x * blockDim.x;
    for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        const i64 k = i % spu;
        const u64 r = sx_mix(seed ^ (u64)i * 0xD1342543DE82EF95ull);
        ts[i] = t0 + k * 1000000;                       // 1 s cadence
        const u32 a = (u32)(r & 0xffff), b = (u32)((r >> 16) & 0xffffff), c = (u32)(r >> 40);
        gpu[i] = (a < 19661) ? 0 : (i64)(b % 101);      // 0 w.p. 0.3, else U{0..100}
        mem[i] = (i64)(c % 101);
    }
    for (i64 u = (i64)blockIdx.x * blockDim.x + threadIdx.x; u <= n_uuids; u += stride) off[u] = u * spu;
}

__global__ void __launch_bounds__(256) sx_windows_kernel(i64 n_uuids, i64 spu, i64 t0, u64 seed, i64 w, i64 *wu,
                                                         i64 *ws, i64 *we) {
    const i64 stride = (i64)gridDim.x * blockDim.x;
    for (i64 j = (i64)blockIdx.x * blockDim.x + threadIdx.x; j < w; j += stride) {
        // windows are dealt to uuids in contiguous runs; window r of a uuid starts r * pitch seconds in
        const i64 u = (i64)(((__int128)j * n_uuids) / w);
        const i64 first = (i64)((((__int128)u * w) + n_uuids - 1) / n_uuids);
        const i64 next = (i64)((((__int128)(u + 1) * w) + n_uuids - 1) / n_uuids);
        const i64 per = next - first > 0 ? next - first : 1;
        const i64 r = j - first;
        const i64 pitch = spu / per > 0 ? spu / per : 1;   // seconds between window starts
        const u64 h = sx_mix(seed ^ (u64)j * 0xA24BAED4963EE407ull);
        const u32 cls = (u32)(h % 100);
        i64 start = t0 + r * pitch * 1000000;
        i64 len_us;
        if (cls < 1) {                    // 1 %: window that ended before logging began -> avg = -1
            start = t0 - 3600ll * 1000000 - r * 1000000;
            len_us = 30ll * 1000000;
        } else if (cls < 11) {            // 10 %: ends exactly where the next window starts (shared sample)
            len_us = pitch * 1000000;
        } else {                          // U[20 s, 180 s] with microsecond jitter on half of them
            len_us = (20 + (i64)((h >> 8) % 161)) * 1000000;
            if ((h >> 40) & 1) { start += (i64)((h >> 41) % 1000000); len_us += (i64)((h >> 20) % 1000000); }
        }
        wu[j] = u; ws[j] = start; we[j] = start + len_us;
    }
}

}  // namespace

extern "C" int th_sort_pairs(int device, int64_t n, const int64_t *k_hi, const int64_t *k_lo, int64_t *out_perm,
                             int mem, void *stream) {
    TH_REQUIRE(n >= 0 && n < (1ll << 30), "th_sort_pairs: n=%lld outside [0, 2^30)", (long long)n);
    TH_REQUIRE(mem == TH_MEM_HOST || mem == TH_MEM_DEVICE, "th_sort_pairs: bad mem flag %d", mem);
    if (n == 0) return TH_OK;
    TH_REQUIRE(k_hi && k_lo && out_perm, "th_sort_pairs: NULL array");
    int prev = 0;
    TH_CUDA(cudaGetDevice(&prev));
    if (prev != device) TH_CUDA(cudaSetDevice(device));
    cudaStream_t st = th_call_stream(stream, mem, device);
    int rc = TH_OK;
    {
        ThScratch sc(st);
        const size_t zn = (size_t)n;
        do {
#define SP_TRY(x) if ((rc = (x)) != TH_OK) break;
#define SP_CUDA(x) { cudaError_t e__ = (x); if (e__ != cudaSuccess) { th_set_error("th_sort_pairs: %s -> %s", #x, cudaGetErrorString(e__)); rc = TH_ECUDA; break; } }
            const void *d_hi, *d_lo;
            void *d_perm;
            SP_TRY(sc.in(k_hi, 8 * zn, mem, &d_hi));
            SP_TRY(sc.in(k_lo, 8 * zn, mem, &d_lo));
            SP_TRY(sc.out(out_perm, 8 * zn, mem, &d_perm));
            PairStats *d_stats;
            SP_TRY(sc.alloc((void **)&d_stats, sizeof(PairStats)));
            PairStats init;
            init.min_lo = TH_I64_MAX; init.max_lo = TH_I64_MIN; init.min_hi = TH_I64_MAX; init.max_hi = TH_I64_MIN;
            SP_CUDA(cudaMemcpyAsync(d_stats, &init, sizeof(init), cudaMemcpyHostToDevice, st));
            const int grid = th_grid_for(n, 256 * 4, 8);
            sp_stats_kernel<<<grid, 256, 0, st>>>((const i64 *)d_hi, (const i64 *)d_lo, n, d_stats);
            PairStats stats;
            SP_CUDA(cudaMemcpyAsync(&stats, d_stats, sizeof(stats), cudaMemcpyDeviceToHost, st));
            SP_CUDA(cudaStreamSynchronize(st));
            if (((u64)stats.max_hi - (u64)stats.min_hi) >= (1ull << 32)) {
                th_set_error("th_sort_pairs: k_hi spread does not fit 32 bits");
                rc = TH_EINVAL;
                break;
            }
            ThRec *ra, *rb, *sorted;
            SP_TRY(sc.alloc((void **)&ra, sizeof(ThRec) * zn));
            SP_TRY(sc.alloc((void **)&rb, sizeof(ThRec) * zn));
            sp_rec_kernel<<<grid, 256, 0, st>>>((const i64 *)d_hi, (const i64 *)d_lo, n, stats.min_lo, stats.min_hi, ra);
            int passes = 0;
            SP_TRY(th_radix_sort(ra, rb, n, th_bits_for(((u64)stats.max_lo - (u64)stats.min_lo)),
                                 th_bits_for(((u64)stats.max_hi - (u64)stats.min_hi)), &sorted, &passes, st));
            sp_perm_kernel<<<grid, 256, 0, st>>>(sorted, n, (i64 *)d_perm);
            SP_CUDA(cudaGetLastError());
            SP_TRY(th_copy_back(out_perm, d_perm, 8 * zn, mem, st));
            SP_CUDA(cudaStreamSynchronize(st));
#undef SP_TRY
#undef SP_CUDA
        } while (0);
    }
    if (prev != device) cudaSetDevice(prev);
    return rc;
}

extern "C" int th_synth_samples(int device, int64_t n_uuids, int64_t samples_per_uuid, int64_t t0_us, uint64_t seed,
                                int64_t *uuid_offsets, int64_t *ts_us, int64_t *gpu_util, int64_t *mem_util,
                                void *stream) {
    TH_REQUIRE(n_uuids > 0 && samples_per_uuid > 0, "th_synth_samples: sizes must be positive");
    TH_REQUIRE(uuid_offsets && ts_us && gpu_util && mem_util, "th_synth_samples: NULL array");
    int prev = 0;
    TH_CUDA(cudaGetDevice(&prev));
    if (prev != device) TH_CUDA(cudaSetDevice(device));
    sx_samples_kernel<<<TH_NUM_SMS * 16, 256, 0, (cudaStream_t)stream>>>(n_uuids, samples_per_uuid, t0_us, seed,
                                                                         (i64 *)uuid_offsets, (i64 *)ts_us, (i64 *)gpu_util, (i64 *)mem_util);
    cudaError_t e = cudaGetLastError();
    if (prev != device) cudaSetDevice(prev);
    if (e != cudaSuccess) { th_set_error("th_synth_samples: launch -> %s", cudaGetErrorString(e)); return TH_ECUDA; }
    return TH_OK;
}

extern "C" int th_synth_windows(int device, int64_t n_uuids, int64_t samples_per_uuid, int64_t t0_us, uint64_t seed,
                                int64_t w, int64_t *win_uuid, int64_t *win_start, int64_t *win_end, void *stream) {
    TH_REQUIRE(n_uuids > 0 && samples_per_uuid > 0 && w > 0, "th_synth_windows: sizes must be positive");
    TH_REQUIRE(win_uuid && win_start && win_end, "th_synth_windows: NULL array");
    int prev = 0;
    TH_CUDA(cudaGetDevice(&prev));
    if (prev != device) TH_CUDA(cudaSetDevice(device));
    sx_windows_kernel<<<TH_NUM_SMS * 8, 256, 0, (cudaStream_t)stream>>>(n_uuids, samples_per_uuid, t0_us, seed, w, (i64 *)win_uuid,
                                                                        (i64 *)win_start, (i64 *)win_end);
    cudaError_t e = cudaGetLastError();
    if (prev != device) cudaSetDevice(prev);
    if (e != cudaSuccess) { th_set_error("th_synth_windows: launch -> %s", cudaGetErrorString(e)); return TH_ECUDA; }
    return TH_OK;
}
