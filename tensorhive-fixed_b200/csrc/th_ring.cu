// Device-resident usage-sample ring.
//
// Replaces the usage logger's storage: the reference keeps ONE JSON FILE PER RESERVATION and re-reads and re-writes it
// on every tick (tensorhive/core/services/UsageLoggingService.py:85-116, 159-171), then averages the lists on expiry
// (:196-229).  Here a monitoring tick costs one small host -> device copy (the tick's samples) and one scatter kernel;
// the log itself never leaves the GPU and th_ring_reduce (K4, uint8 path) reads it in place -- no re-upload, no
// concatenation, and the cost of a tick does not depend on how long the log already is.
//
// Layout: one region per GPU uuid inside three device columns (ts int64, gpu_util uint8, mem_util uint8), regions
// addressed by a per-uuid entry {begin, len, t_first, slope}.  A region that fills up is doubled (all regions are laid
// out again by one copy kernel: amortised O(1) per sample).
//
// Timestamps are "delta-encoded" in the one case that matters for traffic: while a uuid's log is an exact arithmetic
// run (ts[k] = t_first + k * cadence -- a monitor that stamps ticks on a fixed grid), its entry carries the cadence
// instead of an interpolation slope and K4 finds both window ends by division, never touching the ts column.  The
// first sample that breaks the run turns the uuid back into a probed one (the ts column is always stored, so nothing
// is lost).  Semantics are those of Log.updated_log: the caller appends a sample only when both metrics are present.
#include "th_common.cuh"
#include <mutex>
#include <vector>

int th_usage_reduce_table_u8(int device, const void *tab, int64_t n_uuids, int64_t cap, const int64_t *ts_us,
                             const uint8_t *gpu_u8, const uint8_t *mem_u8, int64_t w, const int64_t *win_uuid,
                             const int64_t *win_start, const int64_t *win_end, int64_t *out_sum_gpu,
                             int64_t *out_sum_mem, int64_t *out_count, int64_t *out_avg_gpu, int64_t *out_avg_mem,
                             int mem, cudaStream_t st);

namespace {

struct __align__(16) RingSeg { i64 lo; i64 n; i64 t_lo; double inv; };      // th_usage.cu's UsSeg, field for field

struct HostSeg {
    i64 begin, len, cap;
    i64 t_first, t_last;
    i64 cadence;       // > 0: exact run so far; 0: fewer than two samples; -1: irregular
};

__global__ void __launch_bounds__(256) rg_scatter_kernel(i64 k, const i64 *__restrict__ pos, const i64 *__restrict__ ts,
                                                        const u8 *__restrict__ g, const u8 *__restrict__ m, i64 *o_ts,
                                                        u8 *o_g, u8 *o_m) {
    for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < k; i += (i64)gridDim.x * blockDim.x) {
        const i64 p = pos[i];
        o_ts[p] = ts[i]; o_g[p] = g[i]; o_m[p] = m[i];
    }
}

// one block per uuid (grid-stride): copy its samples from the old layout to the new one
__global__ void __launch_bounds__(256) rg_relayout_kernel(i64 n_uuids, const i64 *__restrict__ old_begin,
                                                         const i64 *__restrict__ len, const i64 *__restrict__ new_begin,
                                                         const i64 *__restrict__ ts, const u8 *__restrict__ g,
                                                         const u8 *__restrict__ m, i64 *o_ts, u8 *o_g, u8 *o_m) {
    for (i64 u = blockIdx.x; u < n_uuids; u += gridDim.x) {
        const i64 a = old_begin[u], b = new_begin[u], n = len[u];
        for (i64 i = threadIdx.x; i < n; i += blockDim.x) { o_ts[b + i] = ts[a + i]; o_g[b + i] = g[a + i]; o_m[b + i] = m[a + i]; }
    }
}

}  // namespace

struct th_ring {
    int device;
    std::mutex mu;
    std::vector<HostSeg> seg;
    i64 cap_total, samples;
    i64 first_cap;             // capacity a new uuid starts with
    i64 *ts;
    u8 *gpu, *mem;
    RingSeg *tab;              // device copy of the entries K4 reads
    i64 tab_cap;
    bool tab_dirty;
    i64 relayouts;
};

static void ring_free(th_ring *r) {
    if (r->ts) cudaFree(r->ts);
    if (r->gpu) cudaFree(r->gpu);
    if (r->mem) cudaFree(r->mem);
    if (r->tab) cudaFree(r->tab);
    r->ts = nullptr; r->gpu = nullptr; r->mem = nullptr; r->tab = nullptr;
}

extern "C" int th_ring_create(int device, int64_t samples_per_uuid_hint, th_ring **out) {
    TH_REQUIRE(out != nullptr, "th_ring_create: out is NULL");
    int count = 0;
    TH_CUDA(cudaGetDeviceCount(&count));
    TH_REQUIRE(device >= 0 && device < count, "th_ring_create: device %d not in [0, %d)", device, count);
    th_ring *r = new th_ring();
    r->device = device;
    r->cap_total = 0; r->samples = 0;
    r->first_cap = samples_per_uuid_hint > 16 ? ((samples_per_uuid_hint + 15) / 16) * 16 : 16;
    r->ts = nullptr; r->gpu = nullptr; r->mem = nullptr; r->tab = nullptr;
    r->tab_cap = 0; r->tab_dirty = true; r->relayouts = 0;
    *out = r;
    return TH_OK;
}

extern "C" int th_ring_destroy(th_ring *r) {
    if (!r) return TH_OK;
    int prev = 0;
    cudaGetDevice(&prev);
    cudaSetDevice(r->device);
    cudaDeviceSynchronize();
    ring_free(r);
    cudaSetDevice(prev);
    delete r;
    return TH_OK;
}

extern "C" int th_ring_info(th_ring *r, int64_t *n_uuids, int64_t *samples, int64_t *capacity, int64_t *device_bytes,
                            int64_t *regular_uuids, int64_t *relayouts) {
    TH_REQUIRE(r != nullptr, "th_ring_info: handle is NULL");
    std::lock_guard<std::mutex> lock(r->mu);
    i64 reg = 0;
    for (const HostSeg &s : r->seg) reg += s.cadence > 0 ? 1 : 0;
    if (n_uuids) *n_uuids = (int64_t)r->seg.size();
    if (samples) *samples = r->samples;
    if (capacity) *capacity = r->cap_total;
    if (device_bytes) *device_bytes = r->cap_total * 10 + r->tab_cap * (i64)sizeof(RingSeg);
    if (regular_uuids) *regular_uuids = reg;
    if (relayouts) *relayouts = r->relayouts;
    return TH_OK;
}

// lay every region out again with the capacities in `want_cap` (host, one per uuid) and copy the samples over
static int ring_relayout(th_ring *r, const std::vector<i64> &want_cap, cudaStream_t st) {
    const size_t nu = r->seg.size();
    std::vector<i64> old_begin(nu), len(nu), new_begin(nu);
    i64 total = 0;
    for (size_t u = 0; u < nu; ++u) {
        old_begin[u] = r->seg[u].begin; len[u] = r->seg[u].len;
        new_begin[u] = total;
        total += want_cap[u];                      // capacities are multiples of 16: every region starts 16-byte aligned
    }
    i64 *n_ts = nullptr; u8 *n_g = nullptr, *n_m = nullptr;
    const size_t zt = (size_t)(total > 0 ? total : 16);
    TH_CUDA(cudaMalloc((void **)&n_ts, 8 * zt));
    cudaError_t e = cudaMalloc((void **)&n_g, zt);
    if (e == cudaSuccess) e = cudaMalloc((void **)&n_m, zt);
    if (e != cudaSuccess) {
        cudaFree(n_ts); if (n_g) cudaFree(n_g);
        th_set_error("th_ring_append: cudaMalloc(%zu samples) -> %s", zt, cudaGetErrorString(e));
        return TH_ENOMEM;
    }
    if (r->samples > 0) {
        i64 *d_meta = nullptr;                     // old_begin | len | new_begin
        TH_CUDA(cudaMalloc((void **)&d_meta, 3 * 8 * nu));
        TH_CUDA(cudaMemcpyAsync(d_meta, old_begin.data(), 8 * nu, cudaMemcpyHostToDevice, st));
        TH_CUDA(cudaMemcpyAsync(d_meta + nu, len.data(), 8 * nu, cudaMemcpyHostToDevice, st));
        TH_CUDA(cudaMemcpyAsync(d_meta + 2 * nu, new_begin.data(), 8 * nu, cudaMemcpyHostToDevice, st));
        const int grid = (int)(nu < (size_t)TH_NUM_SMS * 8 ? (nu > 0 ? nu : 1) : (size_t)TH_NUM_SMS * 8);
        rg_relayout_kernel<<<grid, 256, 0, st>>>((i64)nu, d_meta, d_meta + nu, d_meta + 2 * nu, r->ts, r->gpu, r->mem, n_ts,
                                                 n_g, n_m);
        TH_CUDA(cudaGetLastError());
        TH_CUDA(cudaStreamSynchronize(st));        // the host vectors above are pageable: finish before they go away
        cudaFree(d_meta);
    }
    // cudaFree waits for everything in flight on the device, a reduction still reading the old columns included
    if (r->ts) cudaFree(r->ts);
    if (r->gpu) cudaFree(r->gpu);
    if (r->mem) cudaFree(r->mem);
    r->ts = n_ts; r->gpu = n_g; r->mem = n_m;
    r->cap_total = total;
    for (size_t u = 0; u < nu; ++u) { r->seg[u].begin = new_begin[u]; r->seg[u].cap = want_cap[u]; }
    r->relayouts += 1;
    return TH_OK;
}

// One monitoring tick (or any batch): k samples as HOST arrays -- uuid code (>= 0; codes beyond the ring's current uuid
// count extend it), timestamp, the two utilisation percentages.  Timestamps of one uuid must not decrease.  Cost: one
// H2D copy of the batch and one scatter kernel, whatever the log's length; a region that is full doubles first.
extern "C" int th_ring_append(th_ring *r, int64_t k, const int64_t *uuid, const int64_t *ts_us, const uint8_t *gpu_u8,
                              const uint8_t *mem_u8, void *stream) {
    TH_REQUIRE(r != nullptr, "th_ring_append: handle is NULL");
    TH_REQUIRE(k >= 0, "th_ring_append: negative k");
    if (k == 0) return TH_OK;
    TH_REQUIRE(uuid && ts_us && gpu_u8 && mem_u8, "th_ring_append: NULL array");
    std::lock_guard<std::mutex> lock(r->mu);
    DeviceGuard guard(r->device);
    if (!guard.ok) { th_set_error("th_ring_append: cannot select device %d", r->device); return TH_ECUDA; }
    cudaStream_t st = th_call_stream(stream, TH_MEM_HOST, r->device);
    // 1. validate against a scratch copy of the bookkeeping, so that a refused batch changes nothing
    i64 max_u = -1;
    for (i64 i = 0; i < k; ++i) {
        TH_REQUIRE(uuid[i] >= 0 && uuid[i] < (1ll << 24), "th_ring_append: uuid code %lld outside [0, 2^24)", (long long)uuid[i]);
        if (uuid[i] > max_u) max_u = uuid[i];
    }
    std::vector<HostSeg> seg = r->seg;
    if ((i64)seg.size() <= max_u) {
        HostSeg empty; empty.begin = 0; empty.len = 0; empty.cap = 0; empty.t_first = 0; empty.t_last = TH_I64_MIN; empty.cadence = 0;
        seg.resize((size_t)max_u + 1, empty);
    }
    std::vector<i64> pos((size_t)k);
    bool grow = false;
    for (i64 i = 0; i < k; ++i) {
        HostSeg &s = seg[(size_t)uuid[i]];
        const i64 t = ts_us[i];
        TH_REQUIRE(s.len == 0 || t >= s.t_last, "th_ring_append: timestamps of uuid %lld go backwards", (long long)uuid[i]);
        if (s.len == 0) { s.t_first = t; s.cadence = 0; }
        else {
            // t >= t_last: the step as an unsigned difference (stamps at both ends of int64 must not overflow); a step of
            // 2^62 us or more is never a cadence
            const u64 step = (u64)t - (u64)s.t_last;
            if (s.len == 1) s.cadence = (step > 0 && step < (1ull << 62)) ? (i64)step : -1;
            else if (s.cadence > 0 && step != (u64)s.cadence) s.cadence = -1;      // the run is broken: back to probing
        }
        s.t_last = t;
        pos[(size_t)i] = s.len;        // index inside the region; the region's begin is added below (it may still move)
        s.len += 1;
        if (s.len > s.cap) grow = true;
    }
    // 2. make room
    if (grow || r->ts == nullptr) {
        std::vector<i64> want(seg.size());
        for (size_t u = 0; u < seg.size(); ++u) {
            i64 c = seg[u].cap;
            if (seg[u].len > c) { c = c > 0 ? c : r->first_cap; while (c < seg[u].len) c *= 2; }
            want[u] = c;
        }
        const std::vector<HostSeg> keep = r->seg;
        r->seg.resize(seg.size(), HostSeg{0, 0, 0, 0, TH_I64_MIN, 0});
        const int rc = ring_relayout(r, want, st);
        if (rc != TH_OK) { r->seg = keep; return rc; }
        for (size_t u = 0; u < seg.size(); ++u) { seg[u].begin = r->seg[u].begin; seg[u].cap = r->seg[u].cap; }
    }
    for (i64 i = 0; i < k; ++i) pos[(size_t)i] += seg[(size_t)uuid[i]].begin;
    // 3. the tick's samples: one staged copy each, one scatter
    {
        ThScratch sc(st);
        const void *d_pos, *d_ts, *d_g, *d_m;
        TH_TRY(sc.in(pos.data(), 8 * (size_t)k, TH_MEM_HOST, &d_pos));
        TH_TRY(sc.in(ts_us, 8 * (size_t)k, TH_MEM_HOST, &d_ts));
        TH_TRY(sc.in(gpu_u8, (size_t)k, TH_MEM_HOST, &d_g));
        TH_TRY(sc.in(mem_u8, (size_t)k, TH_MEM_HOST, &d_m));
        rg_scatter_kernel<<<th_grid_for(k, 256, 4), 256, 0, st>>>(k, (const i64 *)d_pos, (const i64 *)d_ts, (const u8 *)d_g,
                                                                 (const u8 *)d_m, r->ts, r->gpu, r->mem);
        TH_CUDA(cudaGetLastError());
        TH_CUDA(cudaStreamSynchronize(st));        // `pos` and the caller's arrays are pageable host memory
    }
    r->seg.swap(seg);
    r->samples += k;
    r->tab_dirty = true;
    return TH_OK;
}

// Window sums / averages straight from the ring (K4, uint8 path): UsageLoggingService.handle_expired_logs :196-229 for
// a whole list of expired reservations.  Window arrays and outputs follow `mem`; windows on uuids the ring has never
// seen count as empty (avg = -1).  The only upload is the per-uuid entry table (32 bytes per uuid), and only when
// something was appended since the last reduction.
extern "C" int th_ring_reduce(th_ring *r, int64_t w, const int64_t *win_uuid, const int64_t *win_start,
                              const int64_t *win_end, int64_t *out_sum_gpu, int64_t *out_sum_mem, int64_t *out_count,
                              int64_t *out_avg_gpu, int64_t *out_avg_mem, int mem, void *stream) {
    TH_REQUIRE(r != nullptr, "th_ring_reduce: handle is NULL");
    TH_REQUIRE(w >= 0, "th_ring_reduce: negative w");
    TH_REQUIRE(mem == TH_MEM_HOST || mem == TH_MEM_DEVICE, "th_ring_reduce: bad mem flag %d", mem);
    if (w == 0) return TH_OK;
    std::lock_guard<std::mutex> lock(r->mu);
    DeviceGuard guard(r->device);
    if (!guard.ok) { th_set_error("th_ring_reduce: cannot select device %d", r->device); return TH_ECUDA; }
    cudaStream_t st = th_call_stream(stream, mem, r->device);
    const i64 nu = (i64)r->seg.size();
    if (r->tab_dirty || r->tab == nullptr) {
        if (nu > r->tab_cap || r->tab == nullptr) {
            if (r->tab) cudaFree(r->tab);
            r->tab = nullptr;
            r->tab_cap = nu > 0 ? nu + nu / 2 + 16 : 16;
            TH_CUDA(cudaMalloc((void **)&r->tab, sizeof(RingSeg) * (size_t)r->tab_cap));
        }
        std::vector<RingSeg> host((size_t)(nu > 0 ? nu : 1));
        for (i64 u = 0; u < nu; ++u) {
            const HostSeg &s = r->seg[(size_t)u];
            RingSeg e;
            e.lo = s.begin; e.n = s.len; e.t_lo = s.t_first; e.inv = 0.0;
            if (s.len >= 2 && s.cadence > 0 && s.cadence < (1ll << 52)) e.inv = -(double)s.cadence;      // exact run
            else if (s.len >= 2 && s.t_last > s.t_first) e.inv = (double)(s.len - 1) / ((double)s.t_last - (double)s.t_first);
            host[(size_t)u] = e;
        }
        if (nu > 0) {
            TH_CUDA(cudaMemcpyAsync(r->tab, host.data(), sizeof(RingSeg) * (size_t)nu, cudaMemcpyHostToDevice, st));
            TH_CUDA(cudaStreamSynchronize(st));    // `host` is a local
        }
        r->tab_dirty = false;
    }
    return th_usage_reduce_table_u8(r->device, r->tab, nu, r->cap_total, (const int64_t *)r->ts, r->gpu, r->mem, w, win_uuid, win_start,
                                    win_end, out_sum_gpu, out_sum_mem, out_count, out_avg_gpu, out_avg_mem, mem, st);
}
