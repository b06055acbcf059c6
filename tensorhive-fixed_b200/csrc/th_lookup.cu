// K3 / K5 -- per-uuid point-in-time lookups and the joins the service ticks run on top of them (sm_100a).
//
// Replaces (reference):
//   Reservation.current_events(uuid)[0]           tensorhive/models/Reservation.py:89-104 as used by
//                                                 tensorhive/core/services/ProtectionService.py:90-93
//   Reservation.upcoming_events_for_resource      Reservation.py:106-118 as used by
//                                                 JobSchedulingService.py:94-102,117-132 and scheduling.py:48-53
//   ProtectionService.do_run join                 ProtectionService.py:84-102
//
// One warp per uuid / task: the time grid gives the end of the candidate prefix in two probes, then the
// warp walks backwards 32 rows per step with 16-byte loads until the running maximum of non-cancelled
// `end` (m1) proves nothing earlier can still be current.
#include "th_common.cuh"
#include <mutex>

namespace {

constexpr int LK_THREADS = 256;
constexpr int LK_WARPS = LK_THREADS / 32;

// lowest-id non-cancelled row of segment [lo, hi) with start <= now <= end; returns row index or -1 (all lanes)
__device__ __forceinline__ int lk_current_row(const ThGrid &g, i64 u, const ThP1 *__restrict__ p1,
                                              const ThP3 *__restrict__ p3, int lo, i64 now, int lane,
                                              i64 *best_id_out) {
    int c1 = th_grid_prefix_end<false>(g, p1, u, now);   // rows with start <= now
    i64 best_id = TH_I64_MAX;
    int best_row = -1;
    while (c1 > lo && th_ld8(&p1[c1 - 1].m1) >= now) {
        const int c0 = c1 - 32 > lo ? c1 - 32 : lo;
        const int j = c0 + lane;
        if (j < c1) {
            const longlong2 e = th_ld16(&p3[j]);   // {end, idc}
            if (!th_idc_cancelled(e.y) && e.x >= now && th_idc_id(e.y) < best_id) {
                best_id = th_idc_id(e.y);
                best_row = j;
            }
        }
        c1 = c0;
    }
    const i64 wmin = th_warp_min(best_id);
    const u32 who = __ballot_sync(0xffffffffu, best_id == wmin && best_row >= 0);
    int row = -1;
    if (who) row = __shfl_sync(0xffffffffu, best_row, __ffs(who) - 1);
    *best_id_out = wmin;
    return row;
}

__global__ void __launch_bounds__(LK_THREADS) k3_current_kernel(ThGrid g, const int *__restrict__ seg, const ThP1 *__restrict__ p1,
                                                                const ThP3 *__restrict__ p3, const ThP4 *__restrict__ p4,
                                                                i64 n_uuids, i64 now, i64 *__restrict__ out_res,
                                                                i64 *__restrict__ out_owner) {
    // let a dependent launch (the K5 join, launched with programmatic stream serialisation) start loading its process
    // rows while this grid builds the table; it waits for this grid's completion before it reads the table
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    const int lane = threadIdx.x & 31;
    const i64 wstride = (i64)gridDim.x * LK_WARPS;
    for (i64 u = (i64)blockIdx.x * LK_WARPS + (threadIdx.x >> 5); u < n_uuids; u += wstride) {
        i64 id;
        const int row = lk_current_row(g, u, p1, p3, seg[u] + 1, now, lane, &id);   // seg[u] = sentinel
        if (lane == 0) {
            out_res[u] = row >= 0 ? id : -1;
            out_owner[u] = row >= 0 ? th_ld8(&p4[row].owner) : -1;
        }
    }
}

// Walks the rows of upcoming_events_for_resource(uuid, period) for one warp:
//   (start < now AND end > now) OR (now <= start <= now + period), cancelled rows dropped.
// Calls f(row, owner-needed?) through the accumulators below; returns count and the first row in
// (start, id) order = the lowest row index.
struct UpAcc {
    int count;
    int first_row;
    int owner_hit;   // any matching row whose owner passes the test
};

// G lanes cooperate on one uuid / task (G = 8: four tasks per warp -- the matching rows are few, so 32 lanes per
// task left most issue slots empty; measured 57 us -> see bench service_sweep).  gl = lane index in the group,
// gmask = the group's lanes in the warp.
template <bool WANT_OWNER, int G>
__device__ __forceinline__ UpAcc lk_upcoming(const ThGrid &g, i64 u, const ThP1 *__restrict__ p1,
                                             const ThP3 *__restrict__ p3, const ThP4 *__restrict__ p4, int lo, i64 now,
                                             i64 later, int gl, u32 gmask, int gbase, int test, i64 owner) {
    UpAcc a;
    a.count = 0; a.first_row = 0x7fffffff; a.owner_hit = 0;
    int c1 = th_grid_prefix_end<false>(g, p1, u, later);   // rows with start <= now + period
    while (c1 > lo) {
        const longlong2 top = th_ld16(&p1[c1 - 1]);     // {start, m1}
        if (!(top.x >= now || top.y > now)) break;     // earlier rows start before now and none is still running
        const int c0 = c1 - G > lo ? c1 - G : lo;
        const int j = c0 + gl;
        bool mt = false;
        if (j < c1) {
            const longlong2 e = th_ld16(&p3[j]);        // {end, idc}
            const i64 s = th_ld8(&p1[j].start);
            mt = !th_idc_cancelled(e.y) && (s >= now || e.x > now);
        }
        bool oh = false;
        if (WANT_OWNER && mt) {
            const i64 o = th_ld8(&p4[j].owner);
            oh = test == TH_OWNER_ANY ? true : (test == TH_OWNER_OTHER ? (o != owner) : (o == owner));
        }
        const u32 bal = __ballot_sync(gmask, mt) & gmask;
        if (bal) {
            a.count += __popc(bal);
            a.first_row = c0 + (__ffs(bal) - 1 - gbase);   // chunks go downwards, so the last one seen wins
        }
        if (WANT_OWNER && (__ballot_sync(gmask, oh) & gmask)) a.owner_hit = 1;
        c1 = c0;
    }
    return a;
}

constexpr int LK_GROUP = 4;   // lanes per uuid / task: 250 k job tasks in 22.5 us with 4 lanes, 27.5 us with 8 (chains of dependent probes)
constexpr int LK_GROUPS_PER_BLOCK = LK_THREADS / LK_GROUP;

__global__ void __launch_bounds__(LK_THREADS) k3_upcoming_kernel(ThGrid g, const int *__restrict__ seg, const ThP1 *__restrict__ p1,
                                                                 const ThP3 *__restrict__ p3, const ThP4 *__restrict__ p4,
                                                                 i64 n_uuids, i64 now, i64 later, i64 k,
                                                                 const i64 *__restrict__ uuids, i64 *__restrict__ o_id,
                                                                 i64 *__restrict__ o_start, i64 *__restrict__ o_owner,
                                                                 i64 *__restrict__ o_count) {
    const int lane = threadIdx.x & 31;
    const int gl = lane & (LK_GROUP - 1), gbase = lane & ~(LK_GROUP - 1);
    const u32 gmask = ((1u << LK_GROUP) - 1) << gbase;
    const i64 gstride = (i64)gridDim.x * LK_GROUPS_PER_BLOCK;
    for (i64 q = (i64)blockIdx.x * LK_GROUPS_PER_BLOCK + (threadIdx.x / LK_GROUP); q < k; q += gstride) {
        const i64 u = uuids[q];
        UpAcc a;
        a.count = 0; a.first_row = 0; a.owner_hit = 0;
        if (u >= 0 && u < n_uuids)
            a = lk_upcoming<false, LK_GROUP>(g, u, p1, p3, p4, seg[u] + 1, now, later, gl, gmask, gbase, 0, 0);
        if (gl == 0) {
            if (a.count > 0) {
                const int r = a.first_row;
                o_id[q] = th_idc_id(th_ld8(&p3[r].idc));
                o_start[q] = th_ld8(&p1[r].start);
                o_owner[q] = th_ld8(&p4[r].owner);
            } else {
                o_id[q] = -1; o_start[q] = -1; o_owner[q] = -1;
            }
            o_count[q] = a.count;
        }
    }
}

__global__ void __launch_bounds__(LK_THREADS) k5_owner_flags_kernel(ThGrid g, const int *__restrict__ seg, const ThP1 *__restrict__ p1,
                                                                    const ThP3 *__restrict__ p3,
                                                                    const ThP4 *__restrict__ p4, i64 n_uuids, i64 now,
                                                                    i64 later, i64 t, const i64 *__restrict__ task_uuid,
                                                                    const i64 *__restrict__ task_owner, int test,
                                                                    u8 *__restrict__ out_flag) {
    const int lane = threadIdx.x & 31;
    const int gl = lane & (LK_GROUP - 1), gbase = lane & ~(LK_GROUP - 1);
    const u32 gmask = ((1u << LK_GROUP) - 1) << gbase;
    const i64 gstride = (i64)gridDim.x * LK_GROUPS_PER_BLOCK;
    for (i64 q = (i64)blockIdx.x * LK_GROUPS_PER_BLOCK + (threadIdx.x / LK_GROUP); q < t; q += gstride) {
        const i64 u = task_uuid[q];
        int flag = 0;
        if (u >= 0 && u < n_uuids) {
            const UpAcc a = lk_upcoming<true, LK_GROUP>(g, u, p1, p3, p4, seg[u] + 1, now, later, gl, gmask, gbase, test,
                                                        task_owner[q]);
            flag = a.owner_hit;
        }
        if (gl == 0) out_flag[q] = (u8)flag;
    }
}

// ProtectionService.py:89-102 for one process row, given the per-uuid current reservation table
__global__ void __launch_bounds__(256) k5_violation_kernel(const i64 *__restrict__ cur_res, const i64 *__restrict__ cur_owner,
                                                           i64 n_uuids, int strict, i64 p,
                                                           const i64 *__restrict__ proc_uuid,
                                                           const i64 *__restrict__ proc_owner, u8 *__restrict__ out_flag,
                                                           i64 *__restrict__ out_res) {
    // The process rows do not depend on the table: the first four per thread are read BEFORE waiting for the grid that
    // builds it (no-op unless this kernel was launched with programmatic stream serialisation behind that grid).
    const i64 stride = (i64)gridDim.x * blockDim.x;
    const i64 i0 = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    i64 pu[4], po[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const i64 i = i0 + k * stride;
        pu[k] = -1; po[k] = -1;
        if (i < p) { pu[k] = th_ld_stream(&proc_uuid[i]); po[k] = th_ld_stream(&proc_owner[i]); }
    }
    asm volatile("griddepcontrol.wait;" ::: "memory");
    auto judge = [&](i64 i, i64 u, i64 powner) {
        i64 res = -1, owner = -1;
        if (u >= 0 && u < n_uuids) { res = cur_res[u]; owner = cur_owner[u]; }
        u8 flag = 0;
        i64 rid = -1;
        if (res >= 0) {
            // a reservation is current: `reservation.user is None` (owner < 0) silences the GPU (:94-95)
            if (owner >= 0 && powner != owner) { flag = 1; rid = res; }
        } else if (strict) {
            flag = 1;   // unreserved GPU in strict mode: every process is an intruder, reservation=None (:100-102)
        }
        out_flag[i] = flag;
        out_res[i] = rid;
    };
#pragma unroll
    for (int k = 0; k < 4; ++k)
        if (i0 + k * stride < p) judge(i0 + k * stride, pu[k], po[k]);
    for (i64 i = i0 + 4 * stride; i < p; i += stride) judge(i, proc_uuid[i], proc_owner[i]);
}

}  // namespace

extern "C" int th_current_owner(const th_snapshot *h, int64_t now_us, int64_t *out_res_id, int64_t *out_owner, int mem,
                                void *stream) {
    TH_REQUIRE(h != nullptr && h->loaded, "th_current_owner: snapshot not loaded");
    TH_REQUIRE(mem == TH_MEM_HOST || mem == TH_MEM_DEVICE, "th_current_owner: bad mem flag %d", mem);
    if (h->n_uuids == 0) return TH_OK;
    TH_REQUIRE(out_res_id && out_owner, "th_current_owner: NULL output");
    DeviceGuard guard(h->device);
    if (!guard.ok) { th_set_error("th_current_owner: cannot select device %d", h->device); return TH_ECUDA; }
    cudaStream_t st = th_call_stream(stream, mem, h->device);
    ThScratch sc(st);
    const size_t zu = (size_t)h->n_uuids;
    void *d_res, *d_own;
    TH_TRY(sc.out(out_res_id, 8 * zu, mem, &d_res));
    TH_TRY(sc.out(out_owner, 8 * zu, mem, &d_own));
    k3_current_kernel<<<th_grid_for(h->n_uuids, LK_WARPS, 16), LK_THREADS, 0, st>>>(th_grid_of(h), h->seg, h->p1, h->p3, h->p4, h->n_uuids,
                                                                                    now_us, (i64 *)d_res, (i64 *)d_own);
    TH_CUDA(cudaGetLastError());
    TH_TRY(th_copy_back(out_res_id, d_res, 8 * zu, mem, st));
    TH_TRY(th_copy_back(out_owner, d_own, 8 * zu, mem, st));
    if (mem == TH_MEM_HOST) TH_CUDA(cudaStreamSynchronize(st));
    return TH_OK;
}

// now + period for period >= 0, clamped: a horizon beyond INT64_MAX means "no upper bound" (the reference's datetime
// arithmetic cannot get there; through the C ABI it can, and must not wrap into the past)
static inline int64_t th_later(int64_t now_us, int64_t period_us) {
    return now_us > TH_I64_MAX - period_us ? TH_I64_MAX : now_us + period_us;
}

extern "C" int th_upcoming(const th_snapshot *h, int64_t now_us, int64_t period_us, int64_t k, const int64_t *uuids,
                           int64_t *out_first_id, int64_t *out_first_start, int64_t *out_first_owner,
                           int64_t *out_count, int mem, void *stream) {
    TH_REQUIRE(h != nullptr && h->loaded, "th_upcoming: snapshot not loaded");
    TH_REQUIRE(k >= 0 && period_us >= 0, "th_upcoming: negative k or period");
    TH_REQUIRE(mem == TH_MEM_HOST || mem == TH_MEM_DEVICE, "th_upcoming: bad mem flag %d", mem);
    if (k == 0) return TH_OK;
    TH_REQUIRE(uuids && out_first_id && out_first_start && out_first_owner && out_count, "th_upcoming: NULL array");
    DeviceGuard guard(h->device);
    if (!guard.ok) { th_set_error("th_upcoming: cannot select device %d", h->device); return TH_ECUDA; }
    cudaStream_t st = th_call_stream(stream, mem, h->device);
    ThScratch sc(st);
    const size_t zk = (size_t)k;
    const void *d_u;
    void *d_id, *d_s, *d_o, *d_c;
    TH_TRY(sc.in(uuids, 8 * zk, mem, &d_u));
    TH_TRY(sc.out(out_first_id, 8 * zk, mem, &d_id));
    TH_TRY(sc.out(out_first_start, 8 * zk, mem, &d_s));
    TH_TRY(sc.out(out_first_owner, 8 * zk, mem, &d_o));
    TH_TRY(sc.out(out_count, 8 * zk, mem, &d_c));
    k3_upcoming_kernel<<<th_grid_for(k, LK_GROUPS_PER_BLOCK, 16), LK_THREADS, 0, st>>>(th_grid_of(h), h->seg, h->p1, h->p3, h->p4, h->n_uuids, now_us,
                                                                            th_later(now_us, period_us), k, (const i64 *)d_u,
                                                                            (i64 *)d_id, (i64 *)d_s, (i64 *)d_o, (i64 *)d_c);
    TH_CUDA(cudaGetLastError());
    TH_TRY(th_copy_back(out_first_id, d_id, 8 * zk, mem, st));
    TH_TRY(th_copy_back(out_first_start, d_s, 8 * zk, mem, st));
    TH_TRY(th_copy_back(out_first_owner, d_o, 8 * zk, mem, st));
    TH_TRY(th_copy_back(out_count, d_c, 8 * zk, mem, st));
    if (mem == TH_MEM_HOST) TH_CUDA(cudaStreamSynchronize(st));
    return TH_OK;
}

extern "C" int th_upcoming_owner_flags(const th_snapshot *h, int64_t now_us, int64_t period_us, int64_t t,
                                       const int64_t *task_uuid, const int64_t *task_owner, int test,
                                       uint8_t *out_flag_u8, int mem, void *stream) {
    TH_REQUIRE(h != nullptr && h->loaded, "th_upcoming_owner_flags: snapshot not loaded");
    TH_REQUIRE(t >= 0 && period_us >= 0, "th_upcoming_owner_flags: negative t or period");
    TH_REQUIRE(test >= TH_OWNER_ANY && test <= TH_OWNER_SAME, "th_upcoming_owner_flags: bad test %d", test);
    TH_REQUIRE(mem == TH_MEM_HOST || mem == TH_MEM_DEVICE, "th_upcoming_owner_flags: bad mem flag %d", mem);
    if (t == 0) return TH_OK;
    TH_REQUIRE(task_uuid && task_owner && out_flag_u8, "th_upcoming_owner_flags: NULL array");
    DeviceGuard guard(h->device);
    if (!guard.ok) { th_set_error("th_upcoming_owner_flags: cannot select device %d", h->device); return TH_ECUDA; }
    cudaStream_t st = th_call_stream(stream, mem, h->device);
    ThScratch sc(st);
    const size_t zt = (size_t)t;
    const void *d_u, *d_o;
    void *d_f;
    TH_TRY(sc.in(task_uuid, 8 * zt, mem, &d_u));
    TH_TRY(sc.in(task_owner, 8 * zt, mem, &d_o));
    TH_TRY(sc.out(out_flag_u8, zt, mem, &d_f));
    k5_owner_flags_kernel<<<th_grid_for(t, LK_GROUPS_PER_BLOCK, 16), LK_THREADS, 0, st>>>(th_grid_of(h), h->seg, h->p1, h->p3, h->p4, h->n_uuids, now_us,
                                                                               th_later(now_us, period_us), t, (const i64 *)d_u,
                                                                               (const i64 *)d_o, test, (u8 *)d_f);
    TH_CUDA(cudaGetLastError());
    TH_TRY(th_copy_back(out_flag_u8, d_f, zt, mem, st));
    if (mem == TH_MEM_HOST) TH_CUDA(cudaStreamSynchronize(st));
    return TH_OK;
}

extern "C" int th_violation_join(const th_snapshot *h, int64_t now_us, int strict, int64_t p, const int64_t *proc_uuid,
                                 const int64_t *proc_owner, uint8_t *out_flag_u8, int64_t *out_res_id, int mem,
                                 void *stream) {
    TH_REQUIRE(h != nullptr && h->loaded, "th_violation_join: snapshot not loaded");
    TH_REQUIRE(p >= 0, "th_violation_join: negative p");
    TH_REQUIRE(mem == TH_MEM_HOST || mem == TH_MEM_DEVICE, "th_violation_join: bad mem flag %d", mem);
    if (p == 0) return TH_OK;
    TH_REQUIRE(proc_uuid && proc_owner && out_flag_u8 && out_res_id, "th_violation_join: NULL array");
    DeviceGuard guard(h->device);
    if (!guard.ok) { th_set_error("th_violation_join: cannot select device %d", h->device); return TH_ECUDA; }
    cudaStream_t st = th_call_stream(stream, mem, h->device);
    ThScratch sc(st);
    const size_t zp = (size_t)p, zu = (size_t)h->n_uuids;
    // scratch: the per-uuid "current reservation" table the join probes
    i64 *cur_res, *cur_owner;
    TH_TRY(sc.alloc((void **)&cur_res, 8 * (zu ? zu : 1)));
    TH_TRY(sc.alloc((void **)&cur_owner, 8 * (zu ? zu : 1)));
    if (h->n_uuids > 0) {
        k3_current_kernel<<<th_grid_for(h->n_uuids, LK_WARPS, 16), LK_THREADS, 0, st>>>(th_grid_of(h), h->seg, h->p1, h->p3, h->p4,
                                                                                        h->n_uuids, now_us, cur_res,
                                                                                        cur_owner);
    }
    const void *d_u, *d_o;
    void *d_f, *d_r;
    TH_TRY(sc.in(proc_uuid, 8 * zp, mem, &d_u));
    TH_TRY(sc.in(proc_owner, 8 * zp, mem, &d_o));
    TH_TRY(sc.out(out_flag_u8, zp, mem, &d_f));
    TH_TRY(sc.out(out_res_id, 8 * zp, mem, &d_r));
    {
        // programmatic dependent launch: the join may start (and read its process rows) while K3 is still running
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)th_grid_for(p, 256 * 4, 16));
        cfg.blockDim = dim3(256);
        cfg.stream = st;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = at;
        cfg.numAttrs = 1;
        TH_CUDA(cudaLaunchKernelEx(&cfg, k5_violation_kernel, (const i64 *)cur_res, (const i64 *)cur_owner, (i64)h->n_uuids,
                                   strict ? 1 : 0, (i64)p, (const i64 *)d_u, (const i64 *)d_o, (u8 *)d_f, (i64 *)d_r));
    }
    TH_CUDA(cudaGetLastError());
    TH_TRY(th_copy_back(out_flag_u8, d_f, zp, mem, st));
    TH_TRY(th_copy_back(out_res_id, d_r, 8 * zp, mem, st));
    if (mem == TH_MEM_HOST) TH_CUDA(cudaStreamSynchronize(st));
    return TH_OK;
}


// ---------------------------------------------------------------------------------------------------------
// th_violation_join_static -- the violation join for callers that must not allocate device memory.
//
// Who that is: a protection tick that runs on a GPU whose HBM a reserved user's job has already filled (an
// allocation there can fail, or succeed only by evicting someone), and host processes that promised their
// operator a fixed device footprint after start-up.  th_violation_join needs a per-uuid table of the current
// reservation (id, owner) between its two kernels and takes it from the stream-ordered allocator; this variant
// keeps that table in a fixed array that is part of the library image instead.
//
// The price is the fixed size: TH_STATIC_JOIN_MAX_UUIDS uuids (1 MiB of device memory on every GPU that loads
// the library, held until the process exits).  Larger snapshots are refused with TH_ECAPACITY.
//
// The array is shared by every call on a device and still holds the previous call's table when the next one
// starts.  Nothing in it has to be zeroed: the first kernel rewrites entries [0, n_uuids) and the second
// kernel reads only those.  Two calls whose kernels overlap on the same GPU (different streams, or different
// threads on the same snapshot) would overwrite each other's table -- serialise them.
//
// The device address of the array is looked up the first time a device is used (cudaGetSymbolAddress, with
// that device current).  Make that first call outside any stream capture.
static __device__ __align__(256) i64 g_static_join_table[2 * TH_STATIC_JOIN_MAX_UUIDS];   // [res_id | owner]

namespace {
constexpr int SJ_MAX_DEVICES = 64;
std::mutex g_sj_mutex;
i64 *g_sj_ptr[SJ_MAX_DEVICES] = {nullptr};

// device address of g_static_join_table on the CURRENT device; resolved once per device
int sj_table(int device, i64 **out) {
    if (device < 0 || device >= SJ_MAX_DEVICES) {
        th_set_error("th_violation_join_static: device index %d is beyond the %d this entry point tracks", device,
                     SJ_MAX_DEVICES);
        return TH_EINVAL;
    }
    std::lock_guard<std::mutex> lock(g_sj_mutex);
    if (g_sj_ptr[device] == nullptr) {
        // In host code `g_static_join_table` is only a handle for the symbol, not an address a kernel can use.
        void *p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_static_join_table);
        if (e != cudaSuccess || p == nullptr) {
            th_set_error("th_violation_join_static: cudaGetSymbolAddress on device %d -> %s", device,
                         cudaGetErrorString(e));
            return TH_ECUDA;
        }
        g_sj_ptr[device] = (i64 *)p;
    }
    *out = g_sj_ptr[device];
    return TH_OK;
}
}  // namespace

extern "C" int th_violation_join_static(const th_snapshot *h, int64_t now_us, int strict, int64_t p,
                                        const int64_t *proc_uuid, const int64_t *proc_owner, uint8_t *out_flag_u8,
                                        int64_t *out_res_id, void *stream) {
    TH_REQUIRE(h != nullptr && h->loaded, "th_violation_join_static: snapshot not loaded");
    TH_REQUIRE(p >= 0, "th_violation_join_static: negative p");
    if (h->n_uuids > TH_STATIC_JOIN_MAX_UUIDS) {
        th_set_error("th_violation_join_static: snapshot has %lld uuids, the built-in table holds %d; use "
                     "th_violation_join (which allocates) or shard the calendar",
                     (long long)h->n_uuids, TH_STATIC_JOIN_MAX_UUIDS);
        return TH_ECAPACITY;
    }
    if (p == 0) return TH_OK;
    TH_REQUIRE(proc_uuid && proc_owner && out_flag_u8 && out_res_id, "th_violation_join_static: NULL array");
    DeviceGuard guard(h->device);
    if (!guard.ok) { th_set_error("th_violation_join_static: cannot select device %d", h->device); return TH_ECUDA; }
    i64 *table = nullptr;
    TH_TRY(sj_table(h->device, &table));
    i64 *cur_res = table, *cur_owner = table + TH_STATIC_JOIN_MAX_UUIDS;
    cudaStream_t st = (cudaStream_t)stream;
    if (h->n_uuids > 0) {
        k3_current_kernel<<<th_grid_for(h->n_uuids, LK_WARPS, 16), LK_THREADS, 0, st>>>(th_grid_of(h), h->seg, h->p1, h->p3, h->p4,
                                                                                        h->n_uuids, now_us, cur_res,
                                                                                        cur_owner);
    }
    k5_violation_kernel<<<th_grid_for(p, 256 * 4, 16), 256, 0, st>>>(cur_res, cur_owner, h->n_uuids, strict ? 1 : 0, p,
                                                                     (const i64 *)proc_uuid, (const i64 *)proc_owner,
                                                                     (u8 *)out_flag_u8, (i64 *)out_res_id);
    TH_CUDA(cudaGetLastError());
    return TH_OK;
}
