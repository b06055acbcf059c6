// K2 -- batched interval-overlap search over a loaded snapshot (sm_100a).
//
// Replaces (reference, tensorhive/models/Reservation.py):
//   would_interfere :120-130                 -> th_conflict_bitmap(TH_MODE_STRICT), th_conflict_sets
//   filter_by_uuids_and_time_range :132-148  -> th_conflict_sets(TH_MODE_INCLUSIVE)
//   current_events :89-104                   -> th_conflict_sets(TH_MODE_INCLUSIVE, qs = qe = now)
//   upcoming_events_for_resource :106-118    -> th_conflict_sets(TH_MODE_UPCOMING)
//
// Bitmap kernel: the rows of a uuid with start < qe are a prefix of its segment, so the verdict is one
// lower-bound search over 16-byte {start, m1} entries; the last probe that moved the lower bound IS the
// prefix's last row, so its m1 (running max of non-cancelled `end`) is already in registers.  One lane
// per query, one ballot per warp = one 32-bit bitmap word per warp, stored coalesced.
//
// Set kernels: one warp per query.  The warp walks the candidate prefix backwards 32 rows at a time with
// 16-byte loads and stops as soon as the running maximum over the remaining prefix says no further row
// can match; a second pass re-walks the examined range forwards and compacts matches with ballots.
#include "th_common.cuh"

namespace {

struct DeviceGuard {
    int prev;
    bool ok;
    explicit DeviceGuard(int dev) : prev(0), ok(true) {
        if (cudaGetDevice(&prev) != cudaSuccess) ok = false;
        if (ok && prev != dev && cudaSetDevice(dev) != cudaSuccess) ok = false;
    }
    ~DeviceGuard() { cudaSetDevice(prev); }
};

// number of rows in [lo, hi) with start < key (strict) or start <= key (!strict)
template <bool STRICT>
__device__ __forceinline__ int th_prefix_end(const ThP1 *__restrict__ p1, int lo, int hi, i64 key) {
    while (lo < hi) {
        const int mid = (int)(((unsigned)lo + (unsigned)hi) >> 1);
        const i64 s = __ldg(&p1[mid].start);
        const bool left = STRICT ? (s < key) : (s <= key);
        if (left) lo = mid + 1; else hi = mid;
    }
    return lo;
}

constexpr int BM_THREADS = 256;

// One lane per query.  Memory instructions whose 32 lanes touch 32 different cache lines cost 32 L1
// wavefronts each, and that -- not bytes -- is what bounds this kernel, so the lookup is built to need as
// few divergent probes as possible: ONE 16-byte grid cell {pos, cnt, m1_before}; only if the candidate's
// bucket holds rows, a search over those (usually 1-2) 16-byte {start, m1} entries; only for self-updates a
// third probe for {a1, m2}.  The four query columns are read coalesced.
__global__ void __launch_bounds__(BM_THREADS) k2_bitmap_kernel(ThGrid g, const ThP1 *__restrict__ p1,
                                                               const ThP2 *__restrict__ p2, const ThP4 *__restrict__ p4,
                                                               i64 n_uuids, i64 m, const i64 *__restrict__ q_uuid,
                                                               const i64 *__restrict__ q_start,
                                                               const i64 *__restrict__ q_end,
                                                               const i64 *__restrict__ q_self, int mode,
                                                               u32 *__restrict__ out_bits) {
    const i64 i = (i64)blockIdx.x * BM_THREADS + threadIdx.x;
    bool hit = false;
    if (i < m) {
        const i64 u = q_uuid[i];
        const i64 qs = q_start[i], qe = q_end[i];
        const i64 self = (mode == TH_MODE_STRICT) ? q_self[i] : -1;
        if (u >= 0 && u < n_uuids) {
            i64 b = 0;
            if (qe > g.t0) {
                const u64 d = ((u64)qe - (u64)g.t0) >> g.shift;
                b = d < (u64)(g.nb - 1) ? (i64)d : (i64)(g.nb - 1);
            }
            const longlong2 c = th_ld16(&g.cells[u * g.nb + b]);   // {pos | cnt << 32, m1_before}
            int lo = (int)(c.x & 0xffffffffll);
            int hi = lo + (int)(c.x >> 32);
            if (mode == TH_MODE_STRICT) {
                i64 m1 = c.y;
                while (lo < hi) {
                    const int mid = (int)(((unsigned)lo + (unsigned)hi) >> 1);
                    const longlong2 e = th_ld16(&p1[mid]);   // {start, m1}
                    if (e.x < qe) { lo = mid + 1; m1 = e.y; } else hi = mid;
                }
                // lo - 1 is the last row with start < qe; it is at worst the uuid's sentinel
                if (self >= 0) {
                    const longlong2 t = th_ld16(&p2[lo - 1]);   // {a1, m2}
                    hit = (t.x == self) ? (t.y > qs) : (m1 > qs);
                } else {
                    hit = m1 > qs;
                }
            } else {   // TH_MODE_INCLUSIVE: any row, cancelled or not, with start <= qe and qs <= end
                while (lo < hi) {
                    const int mid = (int)(((unsigned)lo + (unsigned)hi) >> 1);
                    if (__ldg(&p1[mid].start) <= qe) lo = mid + 1; else hi = mid;
                }
                hit = __ldg(&p4[lo - 1].mall) >= qs;
            }
        }
    }
    const u32 word = __ballot_sync(0xffffffffu, hit);
    if ((threadIdx.x & 31) == 0 && i < m) out_bits[i >> 5] = word;
}

// ---- conflict sets ------------------------------------------------------------------------------------
constexpr int CS_THREADS = 256;
constexpr int CS_WARPS = CS_THREADS / 32;

struct RowView {
    const int *seg; const ThP1 *p1; const ThP3 *p3; const ThP4 *p4;
};

// does row j belong to the result list of (qs, qe, self) under `mode`?
__device__ __forceinline__ bool cs_match(int mode, i64 start, i64 end, i64 idc, i64 qs, i64 self) {
    if (mode == TH_MODE_STRICT) return end > qs && th_idc_id(idc) != self;
    if (mode == TH_MODE_INCLUSIVE) return end >= qs;
    return !th_idc_cancelled(idc) && (start >= qs || end > qs);   // TH_MODE_UPCOMING
}

// warp-uniform: can any row of [lo, c1) still match?  (looks at the running maximum stored at c1 - 1)
__device__ __forceinline__ bool cs_more(int mode, const RowView &v, int c1, i64 qs) {
    if (mode == TH_MODE_STRICT) return __ldg(&v.p4[c1 - 1].mall) > qs;
    if (mode == TH_MODE_INCLUSIVE) return __ldg(&v.p4[c1 - 1].mall) >= qs;
    const longlong2 e = th_ld16(&v.p1[c1 - 1]);   // {start, m1}: rows at or after qs always match
    return e.x >= qs || e.y > qs;
}

__global__ void __launch_bounds__(CS_THREADS) k2_sets_count_kernel(RowView v, i64 n_uuids, i64 m,
                                                                  const i64 *__restrict__ q_uuid,
                                                                  const i64 *__restrict__ q_start,
                                                                  const i64 *__restrict__ q_end,
                                                                  const i64 *__restrict__ q_self, int mode,
                                                                  int *__restrict__ out_cnt, int *__restrict__ out_beg,
                                                                  int *__restrict__ out_end) {
    const int lane = threadIdx.x & 31;
    const i64 wstride = (i64)gridDim.x * CS_WARPS;
    for (i64 q = (i64)blockIdx.x * CS_WARPS + (threadIdx.x >> 5); q < m; q += wstride) {
        const i64 u = q_uuid[q];
        int cnt = 0, beg = 0, pend = 0;
        if (u >= 0 && u < n_uuids) {
            const int lo = v.seg[u] + 1, hi = v.seg[u + 1];   // position seg[u] is the sentinel
            const i64 qs = q_start[q], qe = q_end[q];
            const i64 self = q_self ? q_self[q] : -1;
            pend = (mode == TH_MODE_STRICT) ? th_prefix_end<true>(v.p1, lo, hi, qe)
                                            : th_prefix_end<false>(v.p1, lo, hi, qe);
            int c1 = pend;
            beg = pend;
            while (c1 > lo && cs_more(mode, v, c1, qs)) {
                const int c0 = c1 - 32 > lo ? c1 - 32 : lo;
                const int j = c0 + lane;
                bool mt = false;
                if (j < c1) {
                    const longlong2 e = th_ld16(&v.p3[j]);   // {end, idc}
                    const i64 s = (mode == TH_MODE_UPCOMING) ? __ldg(&v.p1[j].start) : 0;
                    mt = cs_match(mode, s, e.x, e.y, qs, self);
                }
                cnt += __popc(__ballot_sync(0xffffffffu, mt));
                beg = c0;
                c1 = c0;
            }
        }
        if (lane == 0) { out_cnt[q] = cnt; out_beg[q] = beg; out_end[q] = pend; }
    }
}

__global__ void __launch_bounds__(CS_THREADS) k2_sets_fill_kernel(RowView v, i64 m, const i64 *__restrict__ q_start,
                                                                 const i64 *__restrict__ q_self, int mode,
                                                                 const int *__restrict__ beg_arr,
                                                                 const int *__restrict__ end_arr,
                                                                 const i64 *__restrict__ offsets,
                                                                 i64 *__restrict__ out_ids, u8 *__restrict__ out_canc) {
    const int lane = threadIdx.x & 31;
    const i64 wstride = (i64)gridDim.x * CS_WARPS;
    for (i64 q = (i64)blockIdx.x * CS_WARPS + (threadIdx.x >> 5); q < m; q += wstride) {
        const i64 o0 = offsets[q];
        if (offsets[q + 1] == o0) continue;
        const int beg = beg_arr[q], pend = end_arr[q];
        const i64 qs = q_start[q];
        const i64 self = q_self ? q_self[q] : -1;
        i64 o = o0;
        for (int c0 = beg; c0 < pend; c0 += 32) {
            const int j = c0 + lane;
            bool mt = false;
            longlong2 e = make_longlong2(0, 0);
            if (j < pend) {
                e = th_ld16(&v.p3[j]);
                const i64 s = (mode == TH_MODE_UPCOMING) ? __ldg(&v.p1[j].start) : 0;
                mt = cs_match(mode, s, e.x, e.y, qs, self);
            }
            const u32 bal = __ballot_sync(0xffffffffu, mt);
            if (mt) {
                const i64 pos = o + __popc(bal & ((1u << lane) - 1));
                out_ids[pos] = th_idc_id(e.y);
                out_canc[pos] = th_idc_cancelled(e.y) ? 1 : 0;
            }
            o += __popc(bal);
        }
    }
}

// rows come out of the fill pass in (start, id) order; SQLite returns them in rowid = ascending-id order.
// Ids are unique, so the rank of an id inside its list is the number of smaller ids.
__global__ void __launch_bounds__(CS_THREADS) k2_sets_order_kernel(i64 m, const i64 *__restrict__ offsets,
                                                                  const i64 *__restrict__ tmp_ids,
                                                                  const u8 *__restrict__ tmp_canc,
                                                                  i64 *__restrict__ out_ids, u8 *__restrict__ out_canc) {
    const int lane = threadIdx.x & 31;
    const i64 wstride = (i64)gridDim.x * CS_WARPS;
    for (i64 q = (i64)blockIdx.x * CS_WARPS + (threadIdx.x >> 5); q < m; q += wstride) {
        const i64 o0 = offsets[q], o1 = offsets[q + 1];
        const i64 c = o1 - o0;
        for (i64 k = lane; k < c; k += 32) {
            const i64 mine = tmp_ids[o0 + k];
            i64 rank = 0;
            for (i64 t = 0; t < c; ++t) rank += tmp_ids[o0 + t] < mine ? 1 : 0;
            out_ids[o0 + rank] = mine;
            out_canc[o0 + rank] = tmp_canc[o0 + k];
        }
    }
}

// ---- exclusive scan int32 counts -> int64 offsets[m + 1] ----------------------------------------------
constexpr int XS_THREADS = 256;
constexpr int XS_IPT = 8;
constexpr int XS_TILE = XS_THREADS * XS_IPT;

__device__ __forceinline__ i64 xs_block_exclusive(i64 v, i64 *wtot, i64 *total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    i64 inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const i64 t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) wtot[warp] = inc;
    __syncthreads();
    i64 pre = 0, tot = 0;
    for (int w = 0; w < XS_THREADS / 32; ++w) {
        if (w < warp) pre += wtot[w];
        tot += wtot[w];
    }
    *total = tot;
    __syncthreads();
    return pre + inc - v;
}

__global__ void __launch_bounds__(XS_THREADS) xs_tile_sum_kernel(const int *__restrict__ cnt, i64 m, i64 *tile_sum) {
    __shared__ i64 wtot[XS_THREADS / 32];
    const i64 j0 = (i64)blockIdx.x * XS_TILE + (i64)threadIdx.x * XS_IPT;
    i64 s = 0;
#pragma unroll
    for (int k = 0; k < XS_IPT; ++k)
        if (j0 + k < m) s += cnt[j0 + k];
    i64 total;
    xs_block_exclusive(s, wtot, &total);
    if (threadIdx.x == 0) tile_sum[blockIdx.x] = total;
}

__global__ void __launch_bounds__(XS_THREADS) xs_carry_kernel(i64 *tile_sum, i64 tiles, i64 *grand_total) {
    __shared__ i64 wtot[XS_THREADS / 32];
    const i64 chunk = (tiles + XS_THREADS - 1) / XS_THREADS;
    const i64 t0 = (i64)threadIdx.x * chunk;
    const i64 t1 = t0 + chunk < tiles ? t0 + chunk : tiles;
    i64 s = 0;
    for (i64 t = t0; t < t1; ++t) s += tile_sum[t];
    i64 total;
    i64 excl = xs_block_exclusive(s, wtot, &total);
    for (i64 t = t0; t < t1; ++t) {
        const i64 a = tile_sum[t];
        tile_sum[t] = excl;
        excl += a;
    }
    if (threadIdx.x == 0) *grand_total = total;
}

__global__ void __launch_bounds__(XS_THREADS) xs_write_kernel(const int *__restrict__ cnt, i64 m,
                                                              const i64 *__restrict__ tile_carry,
                                                              const i64 *__restrict__ grand_total,
                                                              i64 *__restrict__ offsets) {
    __shared__ i64 wtot[XS_THREADS / 32];
    const i64 j0 = (i64)blockIdx.x * XS_TILE + (i64)threadIdx.x * XS_IPT;
    i64 c[XS_IPT];
    i64 s = 0;
#pragma unroll
    for (int k = 0; k < XS_IPT; ++k) {
        c[k] = (j0 + k < m) ? cnt[j0 + k] : 0;
        s += c[k];
    }
    i64 total;
    i64 run = tile_carry[blockIdx.x] + xs_block_exclusive(s, wtot, &total);
#pragma unroll
    for (int k = 0; k < XS_IPT; ++k) {
        if (j0 + k < m) offsets[j0 + k] = run;
        run += c[k];
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) offsets[m] = *grand_total;
}

}  // namespace

extern "C" int th_conflict_bitmap(const th_snapshot *h, int64_t m, const int64_t *q_uuid, const int64_t *q_start,
                                  const int64_t *q_end, const int64_t *q_self_id, int mode, uint32_t *out_bits_u32,
                                  int mem, void *stream) {
    TH_REQUIRE(h != nullptr && h->loaded, "th_conflict_bitmap: snapshot not loaded");
    TH_REQUIRE(m >= 0, "th_conflict_bitmap: negative m");
    TH_REQUIRE(mode == TH_MODE_STRICT || mode == TH_MODE_INCLUSIVE, "th_conflict_bitmap: mode %d not supported", mode);
    TH_REQUIRE(mem == TH_MEM_HOST || mem == TH_MEM_DEVICE, "th_conflict_bitmap: bad mem flag %d", mem);
    if (m == 0) return TH_OK;
    TH_REQUIRE(q_uuid && q_start && q_end && out_bits_u32, "th_conflict_bitmap: NULL array");
    TH_REQUIRE(mode != TH_MODE_STRICT || q_self_id, "th_conflict_bitmap: q_self_id is NULL");
    DeviceGuard guard(h->device);
    if (!guard.ok) { th_set_error("th_conflict_bitmap: cannot select device %d", h->device); return TH_ECUDA; }
    cudaStream_t st = (cudaStream_t)stream;
    ThScratch sc(st);
    const size_t zm = (size_t)m, words = (zm + 31) / 32;
    const void *d_u, *d_s, *d_e, *d_f = nullptr;
    void *d_bits;
    TH_TRY(sc.in(q_uuid, 8 * zm, mem, &d_u));
    TH_TRY(sc.in(q_start, 8 * zm, mem, &d_s));
    TH_TRY(sc.in(q_end, 8 * zm, mem, &d_e));
    if (q_self_id) TH_TRY(sc.in(q_self_id, 8 * zm, mem, &d_f));
    TH_TRY(sc.out(out_bits_u32, 4 * words, mem, &d_bits));
    const unsigned blocks = (unsigned)((m + BM_THREADS - 1) / BM_THREADS);
    k2_bitmap_kernel<<<blocks, BM_THREADS, 0, st>>>(th_grid_of(h), h->p1, h->p2, h->p4, h->n_uuids, m, (const i64 *)d_u,
                                                    (const i64 *)d_s, (const i64 *)d_e, (const i64 *)d_f, mode,
                                                    (u32 *)d_bits);
    TH_CUDA(cudaGetLastError());
    TH_TRY(th_copy_back(out_bits_u32, d_bits, 4 * words, mem, st));
    if (mem == TH_MEM_HOST) TH_CUDA(cudaStreamSynchronize(st));
    return TH_OK;
}

extern "C" int th_conflict_sets(const th_snapshot *h, int64_t m, const int64_t *q_uuid, const int64_t *q_start,
                                const int64_t *q_end, const int64_t *q_self_id, int mode, int64_t *out_offsets,
                                int64_t *out_ids, uint8_t *out_cancelled_u8, int64_t capacity, int64_t *out_needed,
                                int mem, void *stream) {
    TH_REQUIRE(h != nullptr && h->loaded, "th_conflict_sets: snapshot not loaded");
    TH_REQUIRE(m >= 0 && capacity >= 0, "th_conflict_sets: negative size");
    TH_REQUIRE(mode >= TH_MODE_STRICT && mode <= TH_MODE_UPCOMING, "th_conflict_sets: bad mode %d", mode);
    TH_REQUIRE(mem == TH_MEM_HOST || mem == TH_MEM_DEVICE, "th_conflict_sets: bad mem flag %d", mem);
    TH_REQUIRE(out_offsets && out_needed, "th_conflict_sets: out_offsets / out_needed is NULL");
    TH_REQUIRE(m == 0 || (q_uuid && q_start && q_end), "th_conflict_sets: NULL query column");
    TH_REQUIRE(capacity == 0 || (out_ids && out_cancelled_u8), "th_conflict_sets: NULL output with capacity > 0");
    DeviceGuard guard(h->device);
    if (!guard.ok) { th_set_error("th_conflict_sets: cannot select device %d", h->device); return TH_ECUDA; }
    cudaStream_t st = (cudaStream_t)stream;
    ThScratch sc(st);
    const size_t zm = (size_t)m;
    *out_needed = 0;
    void *d_off;
    TH_TRY(sc.out(out_offsets, 8 * (zm + 1), mem, &d_off));
    if (m == 0) {
        TH_CUDA(cudaMemsetAsync(d_off, 0, 8, st));
        TH_TRY(th_copy_back(out_offsets, d_off, 8, mem, st));
        TH_CUDA(cudaStreamSynchronize(st));
        return TH_OK;
    }
    const void *d_u, *d_s, *d_e, *d_f = nullptr;
    TH_TRY(sc.in(q_uuid, 8 * zm, mem, &d_u));
    TH_TRY(sc.in(q_start, 8 * zm, mem, &d_s));
    TH_TRY(sc.in(q_end, 8 * zm, mem, &d_e));
    if (q_self_id && mode == TH_MODE_STRICT) TH_TRY(sc.in(q_self_id, 8 * zm, mem, &d_f));
    int *d_cnt, *d_beg, *d_end;
    TH_TRY(sc.alloc((void **)&d_cnt, 4 * zm));
    TH_TRY(sc.alloc((void **)&d_beg, 4 * zm));
    TH_TRY(sc.alloc((void **)&d_end, 4 * zm));
    RowView v;
    v.seg = h->seg; v.p1 = h->p1; v.p3 = h->p3; v.p4 = h->p4;
    const int wgrid = th_grid_for(m, CS_WARPS, 16);
    k2_sets_count_kernel<<<wgrid, CS_THREADS, 0, st>>>(v, h->n_uuids, m, (const i64 *)d_u, (const i64 *)d_s,
                                                       (const i64 *)d_e, (const i64 *)d_f, mode, d_cnt, d_beg, d_end);
    TH_CUDA(cudaGetLastError());
    const i64 tiles = (m + XS_TILE - 1) / XS_TILE;
    i64 *d_tile, *d_total;
    TH_TRY(sc.alloc((void **)&d_tile, 8 * (size_t)tiles));
    TH_TRY(sc.alloc((void **)&d_total, 8));
    xs_tile_sum_kernel<<<(unsigned)tiles, XS_THREADS, 0, st>>>(d_cnt, m, d_tile);
    xs_carry_kernel<<<1, XS_THREADS, 0, st>>>(d_tile, tiles, d_total);
    xs_write_kernel<<<(unsigned)tiles, XS_THREADS, 0, st>>>(d_cnt, m, d_tile, d_total, (i64 *)d_off);
    TH_CUDA(cudaGetLastError());
    i64 total = 0;
    TH_CUDA(cudaMemcpyAsync(&total, d_total, 8, cudaMemcpyDeviceToHost, st));
    TH_TRY(th_copy_back(out_offsets, d_off, 8 * (zm + 1), mem, st));
    TH_CUDA(cudaStreamSynchronize(st));
    *out_needed = total;
    if (total > capacity) {
        th_set_error("th_conflict_sets: %lld result rows exceed capacity %lld", (long long)total, (long long)capacity);
        return TH_ECAPACITY;
    }
    if (total == 0) return TH_OK;
    void *d_ids, *d_canc;
    TH_TRY(sc.out(out_ids, 8 * (size_t)total, mem, &d_ids));
    TH_TRY(sc.out(out_cancelled_u8, (size_t)total, mem, &d_canc));
    if (mode == TH_MODE_UPCOMING) {
        k2_sets_fill_kernel<<<wgrid, CS_THREADS, 0, st>>>(v, m, (const i64 *)d_s, (const i64 *)d_f, mode, d_beg, d_end,
                                                          (const i64 *)d_off, (i64 *)d_ids, (u8 *)d_canc);
    } else {
        i64 *t_ids; u8 *t_canc;
        TH_TRY(sc.alloc((void **)&t_ids, 8 * (size_t)total));
        TH_TRY(sc.alloc((void **)&t_canc, (size_t)total));
        k2_sets_fill_kernel<<<wgrid, CS_THREADS, 0, st>>>(v, m, (const i64 *)d_s, (const i64 *)d_f, mode, d_beg, d_end,
                                                          (const i64 *)d_off, t_ids, t_canc);
        k2_sets_order_kernel<<<wgrid, CS_THREADS, 0, st>>>(m, (const i64 *)d_off, t_ids, t_canc, (i64 *)d_ids,
                                                           (u8 *)d_canc);
    }
    TH_CUDA(cudaGetLastError());
    TH_TRY(th_copy_back(out_ids, d_ids, 8 * (size_t)total, mem, st));
    TH_TRY(th_copy_back(out_cancelled_u8, d_canc, (size_t)total, mem, st));
    TH_CUDA(cudaStreamSynchronize(st));
    return TH_OK;
}
