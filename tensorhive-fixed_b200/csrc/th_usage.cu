// K4 -- segmented reduction of usage samples into reservation windows (sm_100a).
//
// Replaces the reference's usage averager: per reservation a JSON file that is re-read and re-written on
// every tick (tensorhive/core/services/UsageLoggingService.py:85-116,159-171) and, on expiry,
// `avg(values) = sum(values) // len(values)` or -1 (:25-30, :218-219).  A sample belongs to a window iff
// it was logged while the reservation was in current_events(): start <= ts <= end, both inclusive
// (tensorhive/models/Reservation.py:96-97), so touching windows share their boundary sample.
//
// Eight lanes per window, four windows per warp: the range search is the same instruction stream for every
// lane of a group, so sharing a warp between four windows cuts the issue slots per window by 4 (the first
// version, one warp per window, was instruction-issue bound at 70 % SM throughput, 36 % DRAM).  Samples of a
// uuid are ts-sorted, so the window is a contiguous index range [a, b): both ends are located by an
// interpolation guess followed by ONE coalesced 16-wide probe per end (exact for a regular logging cadence,
// which is what the service produces; both probes are independent and issued back to back), with a
// group-uniform binary search only when a guess misses.  The range is then summed with 16-byte loads of
// gpu_util / mem_util.  The ts column is only probed, never streamed.
#include "th_common.cuh"

namespace {

constexpr int US_THREADS = 256;
constexpr int US_GROUP = 8;                              // lanes per window
constexpr int US_GROUPS_PER_BLOCK = US_THREADS / US_GROUP;

template <bool GE>
__device__ __forceinline__ bool us_pred(i64 x, i64 key) { return GE ? (x >= key) : (x > key); }

// first index in [lo, hi) whose ts satisfies pred (GE: ts >= key, else ts > key); hi if none.
// t_lo = ts[lo], t_hi = ts[hi - 1] (ignored when lo >= hi).  Arguments and result are uniform within the
// 8-lane group `gmask`; gl = lane index inside the group.
template <bool GE>
__device__ __forceinline__ i64 us_search(const i64 *__restrict__ ts, i64 lo, i64 hi, i64 key, i64 t_lo, i64 t_hi,
                                         int gl, u32 gmask) {
    if (lo >= hi) return lo;
    if (us_pred<GE>(t_lo, key)) return lo;
    if (!us_pred<GE>(t_hi, key)) return hi;
    // invariant: pred(ts[lo]) false, pred(ts[hi - 1]) true, hence t_lo < t_hi
    const double frac = (double)(key - t_lo) / (double)(t_hi - t_lo);
    i64 g = lo + (i64)(frac * (double)(hi - 1 - lo));
    if (g < lo) g = lo;
    if (g > hi - 1) g = hi - 1;
    i64 w0 = g - 7;
    if (w0 + 16 > hi) w0 = hi - 16;
    if (w0 < lo) w0 = lo;
    const i64 i0 = w0 + gl, i1 = w0 + 8 + gl;
    const bool f0 = i0 < hi ? !us_pred<GE>(__ldg(&ts[i0]), key) : false;
    const bool f1 = i1 < hi ? !us_pred<GE>(__ldg(&ts[i1]), key) : false;
    const int nfalse = __popc(__ballot_sync(gmask, f0) & gmask) + __popc(__ballot_sync(gmask, f1) & gmask);
    i64 l, r;   // answer = first true in [l, r], pred(ts[r]) known true
    if (nfalse == 0) { l = lo + 1; r = w0; }
    else if (nfalse == 16) { l = w0 + 16; r = hi - 1; }
    else return w0 + nfalse;
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
    const int gl = lane & (US_GROUP - 1);
    const u32 gmask = 0xffu << (lane & ~(US_GROUP - 1));
    const i64 gstride = (i64)gridDim.x * US_GROUPS_PER_BLOCK;
    for (i64 q = (i64)blockIdx.x * US_GROUPS_PER_BLOCK + (threadIdx.x / US_GROUP); q < w; q += gstride) {
        const i64 u = win_uuid[q];
        i64 a = 0, b = 0;
        if (u >= 0 && u < n_uuids) {
            const i64 lo = off[u], hi = off[u + 1];
            if (lo < hi) {
                const i64 ws = win_start[q], we = win_end[q];
                const i64 t_lo = __ldg(&ts[lo]), t_hi = __ldg(&ts[hi - 1]);
                // the two ends are independent searches over the whole segment: their probes overlap
                a = us_search<true>(ts, lo, hi, ws, t_lo, t_hi, gl, gmask);
                b = us_search<false>(ts, lo, hi, we, t_lo, t_hi, gl, gmask);
                if (b < a) b = a;   // end before start: empty window
            }
        }
        i64 sg = 0, sm = 0;
        if (VEC) {
            // element pairs (2p, 2p + 1): 16-byte aligned because the column bases are
            const i64 p1 = (b + 1) >> 1;
            for (i64 p = (a >> 1) + gl; p < p1; p += 2 * US_GROUP) {
                const i64 i0 = p << 1;
                const i64 j0 = (p + US_GROUP) << 1;
                const bool second = p + US_GROUP < p1;
                i64 g0, g1 = 0, m0, m1 = 0, h0 = 0, h1 = 0, n0 = 0, n1 = 0;
                // issue every load of the iteration before consuming any
                if (i0 + 1 < s_total) {
                    const longlong2 g2 = th_ld16(gpu + i0);
                    const longlong2 m2 = th_ld16(mem + i0);
                    g0 = g2.x; g1 = g2.y; m0 = m2.x; m1 = m2.y;
                } else {
                    g0 = __ldg(&gpu[i0]); m0 = __ldg(&mem[i0]);
                }
                if (second) {
                    if (j0 + 1 < s_total) {
                        const longlong2 g2 = th_ld16(gpu + j0);
                        const longlong2 m2 = th_ld16(mem + j0);
                        h0 = g2.x; h1 = g2.y; n0 = m2.x; n1 = m2.y;
                    } else {
                        h0 = __ldg(&gpu[j0]); n0 = __ldg(&mem[j0]);
                    }
                }
                if (i0 >= a && i0 < b) { sg += g0; sm += m0; }
                if (i0 + 1 >= a && i0 + 1 < b) { sg += g1; sm += m1; }
                if (second) {
                    if (j0 < b) { sg += h0; sm += n0; }          // j0 > a always
                    if (j0 + 1 < b) { sg += h1; sm += n1; }
                }
            }
        } else {
            for (i64 i = a + gl; i < b; i += US_GROUP) { sg += __ldg(&gpu[i]); sm += __ldg(&mem[i]); }
        }
#pragma unroll
        for (int o = US_GROUP / 2; o > 0; o >>= 1) {
            sg += __shfl_xor_sync(gmask, sg, o);
            sm += __shfl_xor_sync(gmask, sm, o);
        }
        if (gl == 0) {
            const i64 c = b - a;
            if (o_sg) o_sg[q] = sg;
            if (o_sm) o_sm[q] = sm;
            if (o_cnt) o_cnt[q] = c;
            if (o_ag) o_ag[q] = c > 0 ? us_floor_div(sg, c) : -1;
            if (o_am) o_am[q] = c > 0 ? us_floor_div(sm, c) : -1;
        }
    }
}


// ---- narrow sample format: utilisation as uint8 -------------------------------------
// gpu_util / mem_util are percentages (0..100), so one byte each is lossless; the range sum then reads 2 bytes per
// sample instead of 16.  Same range search as above (ts stays int64: it is probed, never streamed); the sum loads
// 16 samples per lane and load (128-bit), masks the bytes outside [a, b) and adds four bytes per __dp4a.
__device__ __forceinline__ u32 us_mask4(int lo, int hi, int base) {   // 0xff for bytes base..base+3 that lie in [lo, hi)
    u32 m = 0;
#pragma unroll
    for (int t = 0; t < 4; ++t)
        if (base + t >= lo && base + t < hi) m |= 0xffu << (8 * t);
    return m;
}
__device__ __forceinline__ int us_sum16(const uint4 v, int lo, int hi) {
    u32 acc = 0;   // unsigned bytes x 1: at most 16 * 255
    acc = __dp4a(v.x & us_mask4(lo, hi, 0), 0x01010101u, acc);
    acc = __dp4a(v.y & us_mask4(lo, hi, 4), 0x01010101u, acc);
    acc = __dp4a(v.z & us_mask4(lo, hi, 8), 0x01010101u, acc);
    acc = __dp4a(v.w & us_mask4(lo, hi, 12), 0x01010101u, acc);
    return (int)acc;
}

template <bool VEC>
__global__ void __launch_bounds__(US_THREADS) k4_usage_u8_kernel(i64 s_total, i64 n_uuids, const i64 *__restrict__ off,
                                                                 const i64 *__restrict__ ts, const u8 *__restrict__ gpu,
                                                                 const u8 *__restrict__ mem, i64 w,
                                                                 const i64 *__restrict__ win_uuid,
                                                                 const i64 *__restrict__ win_start,
                                                                 const i64 *__restrict__ win_end, i64 *__restrict__ o_sg,
                                                                 i64 *__restrict__ o_sm, i64 *__restrict__ o_cnt,
                                                                 i64 *__restrict__ o_ag, i64 *__restrict__ o_am) {
    const int lane = threadIdx.x & 31;
    const int gl = lane & (US_GROUP - 1);
    const u32 gmask = 0xffu << (lane & ~(US_GROUP - 1));
    const i64 gstride = (i64)gridDim.x * US_GROUPS_PER_BLOCK;
    for (i64 q = (i64)blockIdx.x * US_GROUPS_PER_BLOCK + (threadIdx.x / US_GROUP); q < w; q += gstride) {
        const i64 u = win_uuid[q];
        i64 a = 0, b = 0;
        if (u >= 0 && u < n_uuids) {
            const i64 lo = off[u], hi = off[u + 1];
            if (lo < hi) {
                const i64 ws = win_start[q], we = win_end[q];
                const i64 t_lo = __ldg(&ts[lo]), t_hi = __ldg(&ts[hi - 1]);
                a = us_search<true>(ts, lo, hi, ws, t_lo, t_hi, gl, gmask);
                b = us_search<false>(ts, lo, hi, we, t_lo, t_hi, gl, gmask);
                if (b < a) b = a;
            }
        }
        i64 sg = 0, sm = 0;
        if (VEC) {
            const i64 c1 = (b + 15) >> 4;                       // 16-sample chunks, 16-byte aligned
            for (i64 c = (a >> 4) + gl; c < c1; c += US_GROUP) {
                const i64 i0 = c << 4;
                const int lo = (int)(a > i0 ? a - i0 : 0);
                const int hi = (int)(b < i0 + 16 ? b - i0 : 16);
                if (i0 + 16 <= s_total) {
                    const uint4 g4 = __ldg(reinterpret_cast<const uint4 *>(gpu + i0));
                    const uint4 m4 = __ldg(reinterpret_cast<const uint4 *>(mem + i0));
                    sg += us_sum16(g4, lo, hi);
                    sm += us_sum16(m4, lo, hi);
                } else {
                    for (int t = lo; t < hi; ++t) { sg += gpu[i0 + t]; sm += mem[i0 + t]; }
                }
            }
        } else {
            for (i64 i = a + gl; i < b; i += US_GROUP) { sg += gpu[i]; sm += mem[i]; }
        }
#pragma unroll
        for (int o = US_GROUP / 2; o > 0; o >>= 1) {
            sg += __shfl_xor_sync(gmask, sg, o);
            sm += __shfl_xor_sync(gmask, sm, o);
        }
        if (gl == 0) {
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

static int usage_reduce_impl(int elem, int device, int64_t s, int64_t n_uuids, const int64_t *uuid_offsets,
                             const int64_t *ts_us, const void *gpu_util, const void *mem_util, int64_t w,
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
            US_TRY(sc.in(gpu_util, (size_t)elem * zs, mem, &d_g));
            US_TRY(sc.in(mem_util, (size_t)elem * zs, mem, &d_m));
            US_TRY(sc.in(win_uuid, 8 * zw, mem, &d_wu));
            US_TRY(sc.in(win_start, 8 * zw, mem, &d_ws));
            US_TRY(sc.in(win_end, 8 * zw, mem, &d_we));
            US_TRY(sc.out(out_sum_gpu, 8 * zw, mem, &d_sg));
            US_TRY(sc.out(out_sum_mem, 8 * zw, mem, &d_sm));
            US_TRY(sc.out(out_count, 8 * zw, mem, &d_c));
            US_TRY(sc.out(out_avg_gpu, 8 * zw, mem, &d_ag));
            US_TRY(sc.out(out_avg_mem, 8 * zw, mem, &d_am));
            const bool vec = (((uintptr_t)d_g | (uintptr_t)d_m) & 15) == 0;
            const int grid = th_grid_for(w, US_GROUPS_PER_BLOCK, 64);
            if (elem == 1 && vec)
                k4_usage_u8_kernel<true><<<grid, US_THREADS, 0, st>>>(s, n_uuids, (const i64 *)d_off, (const i64 *)d_ts,
                                                                      (const u8 *)d_g, (const u8 *)d_m, w,
                                                                      (const i64 *)d_wu, (const i64 *)d_ws,
                                                                      (const i64 *)d_we, (i64 *)d_sg, (i64 *)d_sm,
                                                                      (i64 *)d_c, (i64 *)d_ag, (i64 *)d_am);
            else if (elem == 1)
                k4_usage_u8_kernel<false><<<grid, US_THREADS, 0, st>>>(s, n_uuids, (const i64 *)d_off, (const i64 *)d_ts,
                                                                       (const u8 *)d_g, (const u8 *)d_m, w,
                                                                       (const i64 *)d_wu, (const i64 *)d_ws,
                                                                       (const i64 *)d_we, (i64 *)d_sg, (i64 *)d_sm,
                                                                       (i64 *)d_c, (i64 *)d_ag, (i64 *)d_am);
            else if (vec)
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


extern "C" int th_usage_reduce(int device, int64_t s, int64_t n_uuids, const int64_t *uuid_offsets,
                               const int64_t *ts_us, const int64_t *gpu_util, const int64_t *mem_util, int64_t w,
                               const int64_t *win_uuid, const int64_t *win_start, const int64_t *win_end,
                               int64_t *out_sum_gpu, int64_t *out_sum_mem, int64_t *out_count, int64_t *out_avg_gpu,
                               int64_t *out_avg_mem, int mem, void *stream) {
    return usage_reduce_impl(8, device, s, n_uuids, uuid_offsets, ts_us, gpu_util, mem_util, w, win_uuid, win_start, win_end,
                             out_sum_gpu, out_sum_mem, out_count, out_avg_gpu, out_avg_mem, mem, stream);
}

extern "C" int th_usage_reduce_u8(int device, int64_t s, int64_t n_uuids, const int64_t *uuid_offsets,
                                  const int64_t *ts_us, const uint8_t *gpu_util_u8, const uint8_t *mem_util_u8, int64_t w,
                                  const int64_t *win_uuid, const int64_t *win_start, const int64_t *win_end,
                                  int64_t *out_sum_gpu, int64_t *out_sum_mem, int64_t *out_count, int64_t *out_avg_gpu,
                                  int64_t *out_avg_mem, int mem, void *stream) {
    return usage_reduce_impl(1, device, s, n_uuids, uuid_offsets, ts_us, gpu_util_u8, mem_util_u8, w, win_uuid, win_start,
                             win_end, out_sum_gpu, out_sum_mem, out_count, out_avg_gpu, out_avg_mem, mem, stream);
}
