// K4 -- segmented reduction of usage samples into reservation windows (sm_100a).
//
// Replaces the reference's usage averager: per reservation a JSON file that is re-read and re-written on
// every tick (tensorhive/core/services/UsageLoggingService.py:85-116,159-171) and, on expiry,
// `avg(values) = sum(values) // len(values)` or -1 (:25-30, :218-219).  A sample belongs to a window iff
// it was logged while the reservation was in current_events(): start <= ts <= end, both inclusive
// (tensorhive/models/Reservation.py:96-97), so touching windows share their boundary sample.
//
// One warp per window.  Samples of a uuid are ts-sorted, so the window is a contiguous index range
// [a, b): each end is located by an interpolation guess followed by ONE coalesced 32-wide probe (exact
// for regular logging cadence, which is what the service produces), with a warp-uniform binary search
// only when the guess misses.  The range is then summed with 16-byte loads of gpu_util / mem_util.  The ts
// column is only probed, never streamed.
#include "th_common.cuh"

namespace {

constexpr int US_THREADS = 256;
constexpr int US_WARPS = US_THREADS / 32;

template <bool GE>
__device__ __forceinline__ bool us_pred(i64 x, i64 key) { return GE ? (x >= key) : (x > key); }

// first index in [lo, hi) whose ts satisfies pred (GE: ts >= key, else ts > key); hi if none.
// t_lo = ts[lo], t_hi = ts[hi - 1] (ignored when lo >= hi).  Warp-uniform arguments and result.
template <bool GE>
__device__ __forceinline__ i64 us_search(const i64 *__restrict__ ts, i64 lo, i64 hi, i64 key, i64 t_lo, i64 t_hi,
                                         int lane) {
    if (lo >= hi) return lo;
    if (us_pred<GE>(t_lo, key)) return lo;
    if (!us_pred<GE>(t_hi, key)) return hi;
    // invariant: pred(ts[lo]) false, pred(ts[hi - 1]) true, hence t_lo < t_hi
    const double frac = (double)(key - t_lo) / (double)(t_hi - t_lo);
    i64 g = lo + (i64)(frac * (double)(hi - 1 - lo));
    if (g < lo) g = lo;
    if (g > hi - 1) g = hi - 1;
    i64 w0 = g - 15;
    if (w0 + 32 > hi) w0 = hi - 32;
    if (w0 < lo) w0 = lo;
    const i64 idx = w0 + lane;
    const i64 v = idx < hi ? __ldg(&ts[idx]) : TH_I64_MAX;
    const u32 bal = __ballot_sync(0xffffffffu, idx < hi ? us_pred<GE>(v, key) : true);
    i64 l, r;   // answer = first true in [l, r], pred(ts[r]) known true
    if (bal == 0u) { l = w0 + 32; r = hi - 1; }
    else if (bal == 0xffffffffu && w0 > lo) { l = lo + 1; r = w0; }
    else return w0 + (__ffs(bal) - 1);
    while (l < r) {
        const i64 mid = l + ((r - l) >> 1);
        if (us_pred<GE>(__ldg(&ts[mid]), key)) r = mid; else l = mid + 1;
    }
    return l;
}

__device__ __forceinline__ i64 us_floor_div(i64 a, i64 b) {   // b > 0; Python's //
    i64 q = a / b;
    if ((a % b != 0) && (a < 0)) --q;
    return q;
}

template <bool VEC>
__global__ void __launch_bounds__(US_THREADS) k4_usage_kernel(i64 s_total, i64 n_uuids, const i64 *__restrict__ off,
                                                              const i64 *__restrict__ ts, const i64 *__restrict__ gpu,
                                                              const i64 *__restrict__ mem, i64 w,
                                                              const i64 *__restrict__ win_uuid,
                                                              const i64 *__restrict__ win_start,
                                                              const i64 *__restrict__ win_end, i64 *__restrict__ o_sg,
                                                              i64 *__restrict__ o_sm, i64 *__restrict__ o_cnt,
                                                              i64 *__restrict__ o_ag, i64 *__restrict__ o_am) {
    const int lane = threadIdx.x & 31;
    const i64 wstride = (i64)gridDim.x * US_WARPS;
    for (i64 q = (i64)blockIdx.x * US_WARPS + (threadIdx.x >> 5); q < w; q += wstride) {
        const i64 u = win_uuid[q];
        i64 a = 0, b = 0;
        if (u >= 0 && u < n_uuids) {
            const i64 lo = off[u], hi = off[u + 1];
            if (lo < hi) {
                const i64 ws = win_start[q], we = win_end[q];
                const i64 t_lo = __ldg(&ts[lo]), t_hi = __ldg(&ts[hi - 1]);
                a = us_search<true>(ts, lo, hi, ws, t_lo, t_hi, lane);
                b = a;
                if (a < hi) b = us_search<false>(ts, a, hi, we, __ldg(&ts[a]), t_hi, lane);
            }
        }
        i64 sg = 0, sm = 0;
        if (VEC) {
            // element pairs (2p, 2p + 1): 16-byte aligned because the column bases are
            const i64 p1 = (b + 1) >> 1;
            for (i64 p = (a >> 1) + lane; p < p1; p += 32) {
                const i64 i0 = p << 1;
                i64 g0, g1 = 0, m0, m1 = 0;
                if (i0 + 1 < s_total) {
                    const longlong2 g2 = th_ld16(gpu + i0);
                    const longlong2 m2 = th_ld16(mem + i0);
                    g0 = g2.x; g1 = g2.y; m0 = m2.x; m1 = m2.y;
                } else {
                    g0 = __ldg(&gpu[i0]); m0 = __ldg(&mem[i0]);
                }
                if (i0 >= a && i0 < b) { sg += g0; sm += m0; }
                if (i0 + 1 >= a && i0 + 1 < b) { sg += g1; sm += m1; }
            }
        } else {
            for (i64 i = a + lane; i < b; i += 32) { sg += __ldg(&gpu[i]); sm += __ldg(&mem[i]); }
        }
        sg = th_warp_sum(sg);
        sm = th_warp_sum(sm);
        if (lane == 0) {
            const i64 c = b - a;
            if (o_sg) o_sg[q] = sg;
            if (o_sm) o_sm[q] = sm;
            if (o_cnt) o_cnt[q] = c;
            if (o_ag) o_ag[q] = c > 0 ? us_floor_div(sg, c) : -1;
            if (o_am) o_am[q] = c > 0 ? us_floor_div(sm, c) : -1;
        }
    }
}

}  // namespace

extern "C" int th_usage_reduce(int device, int64_t s, int64_t n_uuids, const int64_t *uuid_offsets,
                               const int64_t *ts_us, const int64_t *gpu_util, const int64_t *mem_util, int64_t w,
                               const int64_t *win_uuid, const int64_t *win_start, const int64_t *win_end,
                               int64_t *out_sum_gpu, int64_t *out_sum_mem, int64_t *out_count, int64_t *out_avg_gpu,
                               int64_t *out_avg_mem, int mem, void *stream) {
    TH_REQUIRE(s >= 0 && w >= 0 && n_uuids >= 0, "th_usage_reduce: negative size");
    TH_REQUIRE(mem == TH_MEM_HOST || mem == TH_MEM_DEVICE, "th_usage_reduce: bad mem flag %d", mem);
    if (w == 0) return TH_OK;
    TH_REQUIRE(uuid_offsets && win_uuid && win_start && win_end, "th_usage_reduce: NULL array");
    TH_REQUIRE(s == 0 || (ts_us && gpu_util && mem_util), "th_usage_reduce: NULL sample column");
    int prev = 0;
    TH_CUDA(cudaGetDevice(&prev));
    if (mem == TH_MEM_HOST && prev != device) TH_CUDA(cudaSetDevice(device));
    cudaStream_t st = (cudaStream_t)stream;
    int rc = TH_OK;
    {
        ThScratch sc(st);
        const size_t zs = (size_t)s, zw = (size_t)w, zu = (size_t)n_uuids + 1;
        const void *d_off, *d_ts, *d_g, *d_m, *d_wu, *d_ws, *d_we;
        void *d_sg, *d_sm, *d_c, *d_ag, *d_am;
        do {
#define US_TRY(x) if ((rc = (x)) != TH_OK) break;
            US_TRY(sc.in(uuid_offsets, 8 * zu, mem, &d_off));
            US_TRY(sc.in(ts_us, 8 * zs, mem, &d_ts));
            US_TRY(sc.in(gpu_util, 8 * zs, mem, &d_g));
            US_TRY(sc.in(mem_util, 8 * zs, mem, &d_m));
            US_TRY(sc.in(win_uuid, 8 * zw, mem, &d_wu));
            US_TRY(sc.in(win_start, 8 * zw, mem, &d_ws));
            US_TRY(sc.in(win_end, 8 * zw, mem, &d_we));
            US_TRY(sc.out(out_sum_gpu, 8 * zw, mem, &d_sg));
            US_TRY(sc.out(out_sum_mem, 8 * zw, mem, &d_sm));
            US_TRY(sc.out(out_count, 8 * zw, mem, &d_c));
            US_TRY(sc.out(out_avg_gpu, 8 * zw, mem, &d_ag));
            US_TRY(sc.out(out_avg_mem, 8 * zw, mem, &d_am));
            const bool vec = (((uintptr_t)d_g | (uintptr_t)d_m) & 15) == 0;
            const int grid = th_grid_for(w, US_WARPS, 64);
            if (vec)
                k4_usage_kernel<true><<<grid, US_THREADS, 0, st>>>(s, n_uuids, (const i64 *)d_off, (const i64 *)d_ts,
                                                                   (const i64 *)d_g, (const i64 *)d_m, w,
                                                                   (const i64 *)d_wu, (const i64 *)d_ws,
                                                                   (const i64 *)d_we, (i64 *)d_sg, (i64 *)d_sm,
                                                                   (i64 *)d_c, (i64 *)d_ag, (i64 *)d_am);
            else
                k4_usage_kernel<false><<<grid, US_THREADS, 0, st>>>(s, n_uuids, (const i64 *)d_off, (const i64 *)d_ts,
                                                                    (const i64 *)d_g, (const i64 *)d_m, w,
                                                                    (const i64 *)d_wu, (const i64 *)d_ws,
                                                                    (const i64 *)d_we, (i64 *)d_sg, (i64 *)d_sm,
                                                                    (i64 *)d_c, (i64 *)d_ag, (i64 *)d_am);
            cudaError_t e = cudaGetLastError();
            if (e != cudaSuccess) { th_set_error("th_usage_reduce: launch -> %s", cudaGetErrorString(e)); rc = TH_ECUDA; break; }
            US_TRY(th_copy_back(out_sum_gpu, d_sg, 8 * zw, mem, st));
            US_TRY(th_copy_back(out_sum_mem, d_sm, 8 * zw, mem, st));
            US_TRY(th_copy_back(out_count, d_c, 8 * zw, mem, st));
            US_TRY(th_copy_back(out_avg_gpu, d_ag, 8 * zw, mem, st));
            US_TRY(th_copy_back(out_avg_mem, d_am, 8 * zw, mem, st));
#undef US_TRY
            if (mem == TH_MEM_HOST) {
                e = cudaStreamSynchronize(st);
                if (e != cudaSuccess) { th_set_error("th_usage_reduce: sync -> %s", cudaGetErrorString(e)); rc = TH_ECUDA; }
            }
        } while (0);
    }
    if (mem == TH_MEM_HOST && prev != device) cudaSetDevice(prev);
    return rc;
}
