// Snapshot load: upload the `reservations` columns, sort them by (uuid_id, start, id) with K1 and build
// the per-row index every query kernel probes (segment offsets + running maxima of `end`).
//
// Reference: class Reservation columns (tensorhive/models/Reservation.py:14-37).  The reference has no
// index at all; every predicate is a full table scan inside SQLite.  The index built here lets
//   - would_interfere (:120-130) be answered by ONE lower-bound search plus one 16-byte probe:
//       rows with start < qe are a prefix of the uuid's segment; a conflict exists iff the largest `end`
//       among the non-cancelled rows of that prefix exceeds qs.  `id != self` is handled by keeping the
//       top-2 maxima with the id of the top one (m1/a1/m2).
//   - the list-returning predicates walk backwards from the prefix end only while the running maximum
//     over ALL rows (mall) says a match can still exist.
#include "th_common.cuh"
#include <stdlib.h>

void th_snapshot_release_internal(th_snapshot *h);

namespace {

struct LoadStats {
    i64 min_start, max_start, min_id, max_id, min_uuid, max_uuid;
    u64 id_inversions;
};

__global__ void __launch_bounds__(256) ld_stats_kernel(const i64 *__restrict__ id, const i64 *__restrict__ uuid,
                                                       const i64 *__restrict__ start, i64 n, LoadStats *out) {
    i64 mn_s = TH_I64_MAX, mx_s = TH_I64_MIN, mn_i = TH_I64_MAX, mx_i = TH_I64_MIN, mn_u = TH_I64_MAX, mx_u = TH_I64_MIN;
    u64 inv = 0;
    for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (i64)gridDim.x * blockDim.x) {
        const i64 s = start[i], d = id[i], u = uuid[i];
        mn_s = s < mn_s ? s : mn_s; mx_s = s > mx_s ? s : mx_s;
        mn_i = d < mn_i ? d : mn_i; mx_i = d > mx_i ? d : mx_i;
        mn_u = u < mn_u ? u : mn_u; mx_u = u > mx_u ? u : mx_u;
        if (i + 1 < n && id[i + 1] <= d) ++inv;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        i64 t;
        t = __shfl_xor_sync(0xffffffffu, mn_s, o); mn_s = t < mn_s ? t : mn_s;
        t = __shfl_xor_sync(0xffffffffu, mx_s, o); mx_s = t > mx_s ? t : mx_s;
        t = __shfl_xor_sync(0xffffffffu, mn_i, o); mn_i = t < mn_i ? t : mn_i;
        t = __shfl_xor_sync(0xffffffffu, mx_i, o); mx_i = t > mx_i ? t : mx_i;
        t = __shfl_xor_sync(0xffffffffu, mn_u, o); mn_u = t < mn_u ? t : mn_u;
        t = __shfl_xor_sync(0xffffffffu, mx_u, o); mx_u = t > mx_u ? t : mx_u;
        inv += __shfl_xor_sync(0xffffffffu, inv, o);
    }
    // one set of atomics per CTA, not per warp: thousands of same-address atomics serialise in L2
    __shared__ i64 sh[6][8];
    __shared__ u64 sh_inv[8];
    const int warp = threadIdx.x >> 5;
    if ((threadIdx.x & 31) == 0) {
        sh[0][warp] = mn_s; sh[1][warp] = mx_s; sh[2][warp] = mn_i; sh[3][warp] = mx_i; sh[4][warp] = mn_u; sh[5][warp] = mx_u;
        sh_inv[warp] = inv;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < 8; ++w) {
            mn_s = sh[0][w] < mn_s ? sh[0][w] : mn_s; mx_s = sh[1][w] > mx_s ? sh[1][w] : mx_s;
            mn_i = sh[2][w] < mn_i ? sh[2][w] : mn_i; mx_i = sh[3][w] > mx_i ? sh[3][w] : mx_i;
            mn_u = sh[4][w] < mn_u ? sh[4][w] : mn_u; mx_u = sh[5][w] > mx_u ? sh[5][w] : mx_u;
            inv += sh_inv[w];
        }
        atomicMin(&out->min_start, mn_s); atomicMax(&out->max_start, mx_s);
        atomicMin(&out->min_id, mn_i);    atomicMax(&out->max_id, mx_i);
        atomicMin(&out->min_uuid, mn_u);  atomicMax(&out->max_uuid, mx_u);
        if (inv) atomicAdd((u64 *)&out->id_inversions, inv);
    }
}

// records keyed by id (only when the caller's rows are not already in ascending-id order)
__global__ void __launch_bounds__(256) ld_idrec_kernel(const i64 *__restrict__ id, i64 n, i64 min_id, ThRec *rec,
                                                       i64 base = 0) {   // rows [base, base + n) of the columns
    for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (i64)gridDim.x * blockDim.x) {
        ThRec r;
        r.k0 = (u64)id[base + i] - (u64)min_id; r.k1 = 0; r.idx = (u32)(base + i);
        rec[i] = r;
    }
}

// records sorted by id: any two neighbours with the same id? (ids must be unique: the merge, the id ranking of the row
// lists and the `id != self` exclusion all rely on it)
__global__ void __launch_bounds__(256) ld_dup_kernel(const ThRec *__restrict__ rec, i64 n, int *flag) {
    for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i + 1 < n; i += (i64)gridDim.x * blockDim.x)
        if (rec[i].k0 == rec[i + 1].k0) *flag = 1;
}

// records keyed by (uuid, start - min_start); `order` (optional) gives the ascending-id visiting order
__global__ void __launch_bounds__(256) ld_keyrec_kernel(const i64 *__restrict__ uuid, const i64 *__restrict__ start,
                                                        const ThRec *__restrict__ order, i64 n, i64 min_start,
                                                        ThRec *rec) {
    for (i64 j = (i64)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += (i64)gridDim.x * blockDim.x) {
        const u32 i = order ? order[j].idx : (u32)j;
        ThRec r;
        r.k0 = (u64)start[i] - (u64)min_start; r.k1 = (u32)uuid[i]; r.idx = i;
        rec[j] = r;
    }
}

// Merge of two runs that are each in (uuid, start, id) order -- the survivors of a snapshot and a sorted delta
// (SURVEY.md section 8f item 1).  Ids are unique, so the order is total and every record's place in the output is its
// own index plus the number of records of the OTHER run that precede it: one binary search per record, no passes.
__device__ __forceinline__ bool mg_less(const ThRec &a, const ThRec &b, const i64 *__restrict__ id) {
    if (a.k1 != b.k1) return a.k1 < b.k1;
    if (a.k0 != b.k0) return a.k0 < b.k0;
    return id[a.idx] < id[b.idx];
}
__global__ void __launch_bounds__(256) mg_merge_kernel(const ThRec *__restrict__ a, i64 na, const ThRec *__restrict__ b,
                                                       i64 nb, const i64 *__restrict__ id, ThRec *__restrict__ out) {
    for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < na + nb; i += (i64)gridDim.x * blockDim.x) {
        const bool from_a = i < na;
        const ThRec x = from_a ? a[i] : b[i - na];
        const ThRec *other = from_a ? b : a;
        i64 lo = 0, hi = from_a ? nb : na;
        while (lo < hi) {                       // how many records of the other run come before x
            const i64 mid = lo + ((hi - lo) >> 1);
            if (mg_less(other[mid], x, id)) lo = mid + 1; else hi = mid;
        }
        out[(from_a ? i : i - na) + lo] = x;
    }
}

// gather the payload columns into snapshot order (row j of uuid u lands at position j + u + 1: one sentinel
// per uuid precedes its rows) and derive the sentinel positions seg[u] = (#rows with uuid < u) + u
__global__ void __launch_bounds__(256) ld_finalize_kernel(const ThRec *__restrict__ rec, i64 n, i64 n_uuids,
                                                          i64 min_start, const i64 *__restrict__ id,
                                                          const i64 *__restrict__ end, const i64 *__restrict__ owner,
                                                          const u8 *__restrict__ canc, ThP1 *p1, ThP3 *p3, ThP4 *p4,
                                                          u32 *uuid_out, int *seg) {
    for (i64 j = (i64)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += (i64)gridDim.x * blockDim.x) {
        const ThRec r = rec[j];
        const u32 i = r.idx;
        const i64 cur = r.k1;
        const i64 pos = j + cur + 1;
        p1[pos].start = (i64)r.k0 + min_start;
        ThP3 e;
        e.end = end[i];
        e.idc = id[i] | (canc[i] ? TH_I64_MIN : 0);
        p3[pos] = e;
        p4[pos].owner = owner[i];
        uuid_out[pos] = r.k1;
        const i64 prev = j > 0 ? (i64)rec[j - 1].k1 : -1;
        for (i64 u = prev + 1; u <= cur; ++u) seg[u] = (int)(j + u);
        if (j == n - 1)
            for (i64 u = cur + 1; u <= n_uuids; ++u) seg[u] = (int)(n + u);
    }
}

// empty table: only sentinels
__global__ void __launch_bounds__(256) ld_seg_empty_kernel(i64 n_uuids, int *seg) {
    for (i64 u = (i64)blockIdx.x * blockDim.x + threadIdx.x; u <= n_uuids; u += (i64)gridDim.x * blockDim.x)
        seg[u] = (int)u;
}

__global__ void __launch_bounds__(256) ld_sentinel_kernel(i64 n_uuids, const int *__restrict__ seg, ThP1 *p1, ThP2 *p2,
                                                          ThP3 *p3, ThP4 *p4, u32 *uuid_out) {
    for (i64 u = (i64)blockIdx.x * blockDim.x + threadIdx.x; u < n_uuids; u += (i64)gridDim.x * blockDim.x) {
        const int pos = seg[u];
        ThP1 a; a.start = TH_I64_MIN; a.m1 = TH_I64_MIN;
        ThP2 b; b.a1 = -1; b.m2 = TH_I64_MIN;
        ThP3 c; c.end = TH_I64_MIN; c.idc = -1;          // cancelled, id = 2^63 - 1: never a real row
        ThP4 d; d.owner = -1; d.mall = TH_I64_MIN;
        p1[pos] = a; p2[pos] = b; p3[pos] = c; p4[pos] = d;
        uuid_out[pos] = (u32)u;
    }
}

// Per-uuid grid plan: uuid u gets nb_u = about x4 / 4 cells per row of ITS rows, of power-of-two width, over ITS span of
// starts [first real row, last real row] (rows are start-sorted, so those are positions seg[u] + 1 and seg[u + 1] - 1).
// Also accumulates, in *crowd, sum over uuids of rows_u^2 / (buckets uuid u would occupy under the SHARED plan): divided by
// n that is the number of rows a random row shares its bucket with -- ~1 on a balanced calendar, hundreds when a few hot
// uuids hold most rows or one uuid's span dwarfs the others'.
__global__ void __launch_bounds__(256) ld_ugrid_kernel(i64 n_uuids, const int *__restrict__ seg, const ThP1 *__restrict__ p1,
                                                       i64 x4, int shared_shift, ThUGrid *ug, int *nb_out, double *crowd) {
    for (i64 u = (i64)blockIdx.x * blockDim.x + threadIdx.x; u < n_uuids; u += (i64)gridDim.x * blockDim.x) {
        const int a = seg[u] + 1, b = seg[u + 1];
        ThUGrid e;
        e.t0 = 0; e.base = 0; e.shift = 0; e.nb = 1; e.pad = 0;
        if (b > a) {
            const i64 rows = b - a;
            e.t0 = p1[a].start;
            const u64 span = (u64)p1[b - 1].start - (u64)e.t0;
            i64 target = (x4 * rows + 3) / 4;
            if (target < 1) target = 1;
            int shift = 0;
            // (unsigned: a span of 2^63 us or more -- starts near both int64 extremes -- must not read as negative)
            while (shift < 63 && (span >> shift) >= (u64)target) ++shift;
            e.shift = shift;
            e.nb = (int)((span >> shift) + 1);
            const double used = (double)(span >> shared_shift) + 1.0;
            atomicAdd(crowd, (double)rows * (double)rows / used);
        }
        ug[u] = e;
        nb_out[u] = e.nb;
    }
}
__global__ void __launch_bounds__(256) ld_ugrid_shared_kernel(i64 n_uuids, i64 t0, int shift, int nb, ThUGrid *ug) {
    for (i64 u = (i64)blockIdx.x * blockDim.x + threadIdx.x; u < n_uuids; u += (i64)gridDim.x * blockDim.x) {
        ThUGrid e;
        e.t0 = t0; e.base = u * nb; e.shift = shift; e.nb = nb; e.pad = 0;
        ug[u] = e;
    }
}
__global__ void __launch_bounds__(256) ld_ugrid_base_kernel(i64 n_uuids, const i64 *__restrict__ base, ThUGrid *ug) {
    for (i64 u = (i64)blockIdx.x * blockDim.x + threadIdx.x; u < n_uuids; u += (i64)gridDim.x * blockDim.x)
        ug[u].base = base[u];
}

// time grid: first[c] = first position whose cell is >= c.  The sentinel sits in bucket 0 of its uuid, so
// first[base_u] = seg[u] and the last bucket of uuid u ends at seg[u + 1].
__device__ __forceinline__ i64 ld_cell_of(const ThP1 *__restrict__ p1, const u32 *__restrict__ uuid,
                                          const ThUGrid *__restrict__ ug, i64 pos) {
    const i64 s = p1[pos].start;
    const ThUGrid e = ug[uuid[pos]];
    i64 b = 0;
    if (s != TH_I64_MIN && s > e.t0) {
        const u64 d = ((u64)s - (u64)e.t0) >> e.shift;
        b = d < (u64)(e.nb - 1) ? (i64)d : (i64)(e.nb - 1);
    }
    return e.base + b;
}
__global__ void __launch_bounds__(256) ld_grid_kernel(const ThP1 *__restrict__ p1, const u32 *__restrict__ uuid, i64 nt,
                                                      i64 n_cells, const ThUGrid *__restrict__ ug, int *first) {
    for (i64 pos = (i64)blockIdx.x * blockDim.x + threadIdx.x; pos < nt; pos += (i64)gridDim.x * blockDim.x) {
        const i64 cur = ld_cell_of(p1, uuid, ug, pos);
        const i64 prev = pos > 0 ? ld_cell_of(p1, uuid, ug, pos - 1) : -1;
        for (i64 c = prev + 1; c <= cur; ++c) first[c] = (int)pos;
        if (pos == nt - 1)
            for (i64 c = cur + 1; c <= n_cells; ++c) first[c] = (int)nt;
    }
}
// first[] -> 32-byte cells {pos, cnt, m1 of the row before the bucket (same uuid) or INT64_MIN, first row {start, m1}}
__global__ void __launch_bounds__(256) ld_cells_kernel(const int *__restrict__ first, const ThP1 *__restrict__ p1,
                                                       i64 n_cells, ThCell *cells) {
    for (i64 c = (i64)blockIdx.x * blockDim.x + threadIdx.x; c < n_cells; c += (i64)gridDim.x * blockDim.x) {
        ThCell x;
        x.pos = first[c];
        x.cnt = first[c + 1] - x.pos;
        x.start0 = TH_I64_MAX;                                          // empty bucket: no row before any key
        ThP1 r; r.start = TH_I64_MAX; r.m1 = 0;
        if (x.cnt > 0) r = p1[x.pos];
        // bucket 0 of a uuid starts at its sentinel (start = INT64_MIN, reserved): nothing of this uuid lies before it;
        // every other bucket -- empty ones included -- is preceded by a position of the same uuid
        x.m1_before = (x.cnt > 0 && r.start == TH_I64_MIN) ? TH_I64_MIN : p1[x.pos - 1].m1;
        x.m1_0 = x.m1_before;
        if (x.cnt > 0) { x.start0 = r.start; x.m1_0 = r.m1; }
        cells[c] = x;
    }
}

// ---- segmented running maxima ----------------------------------------------------------------------------
struct St {
    i64 m1, a1, m2, mall;
    int flag;   // a segment head lies inside the span this state summarises
};
__device__ __forceinline__ St st_identity() {
    St s; s.m1 = TH_I64_MIN; s.a1 = -1; s.m2 = TH_I64_MIN; s.mall = TH_I64_MIN; s.flag = 0; return s;
}
// L covers earlier rows than R.  Ties keep the earlier row as top-1; the later equal value becomes m2.
__device__ __forceinline__ St st_combine(const St &L, const St &R) {
    if (R.flag) return R;
    St o;
    o.flag = L.flag;
    o.mall = L.mall > R.mall ? L.mall : R.mall;
    if (R.m1 > L.m1) {
        o.m1 = R.m1; o.a1 = R.a1;
        o.m2 = L.m1 > R.m2 ? L.m1 : R.m2;
    } else {
        o.m1 = L.m1; o.a1 = L.a1;
        o.m2 = L.m2 > R.m1 ? L.m2 : R.m1;
    }
    return o;
}
__device__ __forceinline__ St st_shfl_up(const St &s, int o) {
    St t;
    t.m1 = __shfl_up_sync(0xffffffffu, s.m1, o);
    t.a1 = __shfl_up_sync(0xffffffffu, s.a1, o);
    t.m2 = __shfl_up_sync(0xffffffffu, s.m2, o);
    t.mall = __shfl_up_sync(0xffffffffu, s.mall, o);
    t.flag = __shfl_up_sync(0xffffffffu, s.flag, o);
    return t;
}

constexpr int SC_THREADS = 256;
constexpr int SC_IPT = 8;
constexpr int SC_TILE = SC_THREADS * SC_IPT;

// block-wide inclusive scan of one St per thread; returns the thread's EXCLUSIVE prefix and (in *total,
// valid for every thread) the block aggregate
__device__ __forceinline__ St st_block_exclusive(const St &agg, St *smem_wtot, St *total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    St inc = agg;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        St t = st_shfl_up(inc, o);
        if (lane >= o) inc = st_combine(t, inc);
    }
    St excl = st_shfl_up(inc, 1);
    if (lane == 0) excl = st_identity();
    if (lane == 31) smem_wtot[warp] = inc;
    __syncthreads();
    St pre = st_identity();
    for (int w = 0; w < warp; ++w) pre = st_combine(pre, smem_wtot[w]);
    St tot = pre;
    for (int w = warp; w < SC_THREADS / 32; ++w) tot = st_combine(tot, smem_wtot[w]);
    *total = tot;
    __syncthreads();
    return st_combine(pre, excl);
}

// A: tile-local segmented scan; writes tile-local m1/a1/m2/mall and the tile aggregate
__global__ void __launch_bounds__(SC_THREADS) sc_local_kernel(const u32 *__restrict__ uuid, const ThP3 *__restrict__ p3,
                                                              i64 n, ThP1 *p1, ThP2 *p2, ThP4 *p4, St *tile_agg) {
    __shared__ St wtot[SC_THREADS / 32];
    const i64 tile_start = (i64)blockIdx.x * SC_TILE;
    const i64 j0 = tile_start + (i64)threadIdx.x * SC_IPT;
    St loc[SC_IPT];
    St run = st_identity();
    u32 prev_u = (j0 > 0 && j0 < n) ? uuid[j0 - 1] : 0xffffffffu;
#pragma unroll
    for (int k = 0; k < SC_IPT; ++k) {
        const i64 j = j0 + k;
        if (j < n) {
            const u32 u = uuid[j];
            const ThP3 e = p3[j];
            St x;
            const bool c = th_idc_cancelled(e.idc);
            x.m1 = c ? TH_I64_MIN : e.end;
            x.a1 = c ? -1 : th_idc_id(e.idc);
            x.m2 = TH_I64_MIN;
            x.mall = e.end;
            x.flag = (j == 0 || u != prev_u) ? 1 : 0;
            run = st_combine(run, x);
            prev_u = u;
        }
        loc[k] = run;
    }
    St total;
    const St excl = st_block_exclusive(run, wtot, &total);
#pragma unroll
    for (int k = 0; k < SC_IPT; ++k) {
        const i64 j = j0 + k;
        if (j < n) {
            const St f = st_combine(excl, loc[k]);
            p1[j].m1 = f.m1;
            ThP2 q; q.a1 = f.a1; q.m2 = f.m2;
            p2[j] = q;
            p4[j].mall = f.mall;
        }
    }
    if (threadIdx.x == 0) tile_agg[blockIdx.x] = total;
}

// B: exclusive scan of the tile aggregates (one block; each thread walks a contiguous chunk of tiles)
__global__ void __launch_bounds__(SC_THREADS) sc_carry_kernel(St *tile_agg, i64 tiles) {
    __shared__ St wtot[SC_THREADS / 32];
    const i64 chunk = (tiles + SC_THREADS - 1) / SC_THREADS;
    const i64 t0 = (i64)threadIdx.x * chunk;
    const i64 t1 = t0 + chunk < tiles ? t0 + chunk : tiles;
    St run = st_identity();
    for (i64 t = t0; t < t1; ++t) run = st_combine(run, tile_agg[t]);
    St total;
    St excl = st_block_exclusive(run, wtot, &total);
    for (i64 t = t0; t < t1; ++t) {
        const St a = tile_agg[t];
        tile_agg[t] = excl;          // becomes the carry INTO tile t
        excl = st_combine(excl, a);
    }
}

// C: rows whose segment began in an earlier tile fold the carry in
__global__ void __launch_bounds__(SC_THREADS) sc_fix_kernel(const u32 *__restrict__ uuid, const int *__restrict__ seg,
                                                            i64 n, const St *__restrict__ carry, ThP1 *p1, ThP2 *p2,
                                                            ThP4 *p4) {
    for (i64 j = (i64)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += (i64)gridDim.x * blockDim.x) {
        const i64 tile = j / SC_TILE;
        if (tile == 0) continue;
        const i64 tile_start = tile * SC_TILE;
        if ((i64)seg[uuid[j]] >= tile_start) continue;
        St loc;
        loc.m1 = p1[j].m1;
        const ThP2 q = p2[j];
        loc.a1 = q.a1; loc.m2 = q.m2;
        loc.mall = p4[j].mall;
        loc.flag = 0;
        const St f = st_combine(carry[tile], loc);
        p1[j].m1 = f.m1;
        ThP2 w; w.a1 = f.a1; w.m2 = f.m2;
        p2[j] = w;
        p4[j].mall = f.mall;
    }
}

__global__ void __launch_bounds__(256) ex_export_kernel(i64 nt, i64 n_uuids, const ThP1 *__restrict__ p1,
                                                        const ThP3 *__restrict__ p3, const ThP4 *__restrict__ p4,
                                                        const u32 *__restrict__ uuid, const int *__restrict__ seg,
                                                        i64 *o_id, i64 *o_uuid, i64 *o_start, i64 *o_end, i64 *o_owner,
                                                        u8 *o_canc, i64 *o_seg) {
    const i64 stride = (i64)gridDim.x * blockDim.x;
    for (i64 pos = (i64)blockIdx.x * blockDim.x + threadIdx.x; pos < nt; pos += stride) {
        const i64 u = uuid[pos];
        if (pos == seg[u]) continue;                   // sentinel
        const i64 j = pos - (u + 1);
        const ThP3 e = p3[pos];
        if (o_id) o_id[j] = th_idc_id(e.idc);
        if (o_uuid) o_uuid[j] = u;
        if (o_start) o_start[j] = p1[pos].start;
        if (o_end) o_end[j] = e.end;
        if (o_owner) o_owner[j] = p4[pos].owner;
        if (o_canc) o_canc[j] = th_idc_cancelled(e.idc) ? 1 : 0;
    }
    if (o_seg)   // offsets in the caller's (sentinel-free) row numbering
        for (i64 u = (i64)blockIdx.x * blockDim.x + threadIdx.x; u <= n_uuids; u += stride) o_seg[u] = seg[u] - u;
}

// ---- delta maintenance: survivors of the old table + upserted rows -> columns for a device-side rebuild -----------
__global__ void __launch_bounds__(256) dl_killrec_kernel(const i64 *__restrict__ up_id, i64 n_up, const i64 *__restrict__ del_id,
                                                        i64 n_del, ThRec *rec) {
    for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < n_up + n_del; i += (i64)gridDim.x * blockDim.x) {
        ThRec r;
        r.k0 = (u64)(i < n_up ? up_id[i] : del_id[i - n_up]); r.k1 = 0; r.idx = (u32)i;
        rec[i] = r;
    }
}
// keep[pos] = 1 for every real row whose id is neither replaced nor deleted
__global__ void __launch_bounds__(256) dl_keep_kernel(i64 nt, const u32 *__restrict__ uuid, const int *__restrict__ seg,
                                                     const ThP3 *__restrict__ p3, const ThRec *__restrict__ kill, i64 n_kill,
                                                     int *keep) {
    for (i64 pos = (i64)blockIdx.x * blockDim.x + threadIdx.x; pos < nt; pos += (i64)gridDim.x * blockDim.x) {
        int k = 0;
        if ((i64)seg[uuid[pos]] != pos) {
            const u64 id = (u64)th_idc_id(p3[pos].idc);
            i64 lo = 0, hi = n_kill;
            while (lo < hi) {
                const i64 mid = (lo + hi) >> 1;
                if (kill[mid].k0 < id) lo = mid + 1; else hi = mid;
            }
            k = !(lo < n_kill && kill[lo].k0 == id);
        }
        keep[pos] = k;
    }
}
__global__ void __launch_bounds__(256) dl_gather_kernel(i64 nt, const int *__restrict__ keep, const i64 *__restrict__ off,
                                                       const ThP1 *__restrict__ p1, const ThP3 *__restrict__ p3,
                                                       const ThP4 *__restrict__ p4, const u32 *__restrict__ uuid, i64 *o_id,
                                                       i64 *o_uuid, i64 *o_start, i64 *o_end, i64 *o_owner, u8 *o_canc) {
    for (i64 pos = (i64)blockIdx.x * blockDim.x + threadIdx.x; pos < nt; pos += (i64)gridDim.x * blockDim.x) {
        if (!keep[pos]) continue;
        const i64 j = off[pos];
        const ThP3 e = p3[pos];
        o_id[j] = th_idc_id(e.idc);
        o_uuid[j] = uuid[pos];
        o_start[j] = p1[pos].start;
        o_end[j] = e.end;
        o_owner[j] = p4[pos].owner;
        o_canc[j] = th_idc_cancelled(e.idc) ? 1 : 0;
    }
}
__global__ void __launch_bounds__(256) dl_append_kernel(i64 n_up, i64 base, const i64 *__restrict__ id, const i64 *__restrict__ uuid,
                                                       const i64 *__restrict__ start, const i64 *__restrict__ end,
                                                       const i64 *__restrict__ owner, const u8 *__restrict__ canc, i64 *o_id,
                                                       i64 *o_uuid, i64 *o_start, i64 *o_end, i64 *o_owner, u8 *o_canc) {
    for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < n_up; i += (i64)gridDim.x * blockDim.x) {
        o_id[base + i] = id[i]; o_uuid[base + i] = uuid[i]; o_start[base + i] = start[i]; o_end[base + i] = end[i];
        o_owner[base + i] = owner[i]; o_canc[base + i] = canc[i];
    }
}

}  // namespace

// `presorted` >= 0: rows [0, presorted) are already in (uuid, start, id) order (th_snapshot_apply_delta hands over
// the survivors of the source snapshot that way); only the rows behind them are sorted, then the two runs are merged.
static int th_load_impl(th_snapshot *h, int64_t n, int64_t n_uuids, const int64_t *id, const int64_t *uuid_id,
                        const int64_t *start_us, const int64_t *end_us, const int64_t *owner,
                        const uint8_t *cancelled_u8, int mem, void *stream, int64_t presorted) {
    TH_REQUIRE(h != nullptr, "th_snapshot_load: handle is NULL");
    TH_REQUIRE(n >= 0 && n < (1ll << 30), "th_snapshot_load: n=%lld outside [0, 2^30)", (long long)n);
    TH_REQUIRE(n_uuids >= 0 && n_uuids < (1ll << 30), "th_snapshot_load: n_uuids=%lld outside [0, 2^30)",
               (long long)n_uuids);
    TH_REQUIRE(mem == TH_MEM_HOST || mem == TH_MEM_DEVICE, "th_snapshot_load: bad mem flag %d", mem);
    TH_REQUIRE(n == 0 || (id && uuid_id && start_us && end_us && owner && cancelled_u8),
               "th_snapshot_load: a column pointer is NULL");
    TH_REQUIRE(n == 0 || n_uuids > 0, "th_snapshot_load: rows given but n_uuids == 0");
    DeviceGuard guard(h->device);
    if (!guard.ok) { th_set_error("th_snapshot_load: cannot select device %d", h->device); return TH_ECUDA; }
    cudaStream_t st = th_call_stream(stream, mem, h->device);

    const i64 nt = n + n_uuids;
    const size_t zn = (size_t)n, znt = (size_t)(nt > 0 ? nt : 1);
    // Everything up to the payload gather works on scratch only: the handle is not touched before the input has been
    // validated and sorted, so a rejected (re)load -- bad uuid range, negative or duplicate ids, out of memory for the
    // scratch -- leaves a previously loaded snapshot exactly as it was.
    auto claim_buffers = [&]() -> int {
        // A reload whose table fits the buffers of the previous one reuses them (cudaMalloc/cudaFree of ~80 MB cost
        // more than the sort itself); otherwise everything is released and reallocated.
        h->loaded = 0;
        if (znt > (size_t)h->cap_nt || n_uuids + 1 > h->cap_uuids) {
            th_snapshot_release_internal(h);
            TH_CUDA(cudaMalloc((void **)&h->seg, sizeof(int) * (size_t)(n_uuids + 1)));
            TH_CUDA(cudaMalloc((void **)&h->p1, sizeof(ThP1) * znt));
            TH_CUDA(cudaMalloc((void **)&h->p2, sizeof(ThP2) * znt));
            TH_CUDA(cudaMalloc((void **)&h->p3, sizeof(ThP3) * znt));
            TH_CUDA(cudaMalloc((void **)&h->p4, sizeof(ThP4) * znt));
            TH_CUDA(cudaMalloc((void **)&h->uuid, sizeof(u32) * znt));
            h->cap_nt = (i64)znt;
            h->cap_uuids = n_uuids + 1;
        }
        h->n_uuids = n_uuids;
        return TH_OK;
    };
    i64 new_max_id = 0;

    ThScratch sc(st);
    int passes_total = 0;
    i64 min_start = 0, max_start = 0;
    const int ugrid = th_grid_for(n_uuids + 1, 256, 8);
    if (n == 0) {
        TH_TRY(claim_buffers());
        ld_seg_empty_kernel<<<ugrid, 256, 0, st>>>(n_uuids, h->seg);
        TH_CUDA(cudaGetLastError());
    } else {
        const void *d_id, *d_uuid, *d_start, *d_end, *d_owner, *d_canc;
        TH_TRY(sc.in(id, 8 * zn, mem, &d_id));
        TH_TRY(sc.in(uuid_id, 8 * zn, mem, &d_uuid));
        TH_TRY(sc.in(start_us, 8 * zn, mem, &d_start));
        TH_TRY(sc.in(end_us, 8 * zn, mem, &d_end));
        TH_TRY(sc.in(owner, 8 * zn, mem, &d_owner));
        TH_TRY(sc.in(cancelled_u8, zn, mem, &d_canc));

        // 1. column statistics decide the digit plan
        LoadStats *d_stats = nullptr;
        TH_TRY(sc.alloc((void **)&d_stats, sizeof(LoadStats)));
        LoadStats init;
        init.min_start = TH_I64_MAX; init.max_start = TH_I64_MIN; init.min_id = TH_I64_MAX; init.max_id = TH_I64_MIN;
        init.min_uuid = TH_I64_MAX; init.max_uuid = TH_I64_MIN; init.id_inversions = 0;
        TH_CUDA(cudaMemcpyAsync(d_stats, &init, sizeof(init), cudaMemcpyHostToDevice, st));
        const int grid = th_grid_for(n, 256 * 4, 8);
        ld_stats_kernel<<<grid, 256, 0, st>>>((const i64 *)d_id, (const i64 *)d_uuid, (const i64 *)d_start, n, d_stats);
        TH_CUDA(cudaGetLastError());
        LoadStats stats;
        TH_CUDA(cudaMemcpyAsync(&stats, d_stats, sizeof(stats), cudaMemcpyDeviceToHost, st));
        TH_CUDA(cudaStreamSynchronize(st));
        TH_REQUIRE(stats.min_uuid >= 0 && stats.max_uuid < n_uuids,
                   "th_snapshot_load: uuid_id range [%lld, %lld] outside [0, %lld)", (long long)stats.min_uuid,
                   (long long)stats.max_uuid, (long long)n_uuids);
        TH_REQUIRE(stats.min_id >= 0, "th_snapshot_load: negative reservation id %lld", (long long)stats.min_id);
        TH_REQUIRE(stats.min_start > TH_I64_MIN, "th_snapshot_load: start_us == INT64_MIN is reserved");
        min_start = stats.min_start;
        max_start = stats.max_start;
        new_max_id = stats.max_id;
        ThRec *ra = nullptr, *rb = nullptr;
        TH_TRY(sc.alloc((void **)&ra, sizeof(ThRec) * zn));
        TH_TRY(sc.alloc((void **)&rb, sizeof(ThRec) * zn));
        int passes = 0;
        const ThRec *order = nullptr;
        ThRec *sorted = nullptr;
        ThRec *key_a = ra, *key_b = rb;
        if (presorted >= 0) {
            const i64 nb = n - presorted;
            ld_keyrec_kernel<<<grid, 256, 0, st>>>((const i64 *)d_uuid, (const i64 *)d_start, nullptr, presorted,
                                                   stats.min_start, ra);                 // run A: as it stands
            TH_CUDA(cudaGetLastError());
            sorted = ra;
            if (nb > 0) {
                // run B: the delta rows, by id first so that the stable key sort leaves ties in id order
                ThRec *ba = nullptr, *bb = nullptr, *bc = nullptr, *b_id = nullptr, *b_key = nullptr;
                TH_TRY(sc.alloc((void **)&ba, sizeof(ThRec) * (size_t)nb));
                TH_TRY(sc.alloc((void **)&bb, sizeof(ThRec) * (size_t)nb));
                TH_TRY(sc.alloc((void **)&bc, sizeof(ThRec) * (size_t)nb));
                const int bgrid = th_grid_for(nb, 256 * 4, 8);
                ld_idrec_kernel<<<bgrid, 256, 0, st>>>((const i64 *)d_id, nb, stats.min_id, ba, presorted);
                TH_CUDA(cudaGetLastError());
                TH_TRY(th_radix_sort(ba, bb, nb, th_bits_for((u64)stats.max_id - (u64)stats.min_id), 0, &b_id, &passes, st));
                passes_total += passes;
                ThRec *other = (b_id == ba) ? bb : ba;
                ld_keyrec_kernel<<<bgrid, 256, 0, st>>>((const i64 *)d_uuid, (const i64 *)d_start, b_id, nb, stats.min_start, bc);
                TH_CUDA(cudaGetLastError());
                TH_TRY(th_radix_sort(bc, other, nb, th_bits_for((u64)stats.max_start - (u64)stats.min_start),
                                     th_bits_for((u64)stats.max_uuid), &b_key, &passes, st));
                passes_total += passes;
                mg_merge_kernel<<<grid, 256, 0, st>>>(ra, presorted, b_key, nb, (const i64 *)d_id, rb);
                TH_CUDA(cudaGetLastError());
                sorted = rb;
            }
        } else {
        if (stats.id_inversions > 0) {
            // rows are not in ascending-id order: establish it first so that the stable (uuid, start)
            // sort leaves equal keys in id order (SQLite's scan order for ties).
            ld_idrec_kernel<<<grid, 256, 0, st>>>((const i64 *)d_id, n, stats.min_id, ra);
            TH_CUDA(cudaGetLastError());
            TH_TRY(th_radix_sort(ra, rb, n, th_bits_for((u64)stats.max_id - (u64)stats.min_id), 0, &sorted, &passes, st));
            passes_total += passes;
            int *d_dup = nullptr, dup = 0;
            TH_TRY(sc.alloc((void **)&d_dup, sizeof(int)));
            TH_CUDA(cudaMemsetAsync(d_dup, 0, sizeof(int), st));
            ld_dup_kernel<<<grid, 256, 0, st>>>(sorted, n, d_dup);
            TH_CUDA(cudaMemcpyAsync(&dup, d_dup, sizeof(int), cudaMemcpyDeviceToHost, st));
            TH_CUDA(cudaStreamSynchronize(st));
            TH_REQUIRE(dup == 0, "th_snapshot_load: reservation ids are not unique");
            order = sorted;
            ThRec *rc_buf = nullptr;
            TH_TRY(sc.alloc((void **)&rc_buf, sizeof(ThRec) * zn));
            key_a = rc_buf;
            key_b = (sorted == ra) ? rb : ra;
        }
        // 2. (uuid, start) records and the main sort
        ld_keyrec_kernel<<<grid, 256, 0, st>>>((const i64 *)d_uuid, (const i64 *)d_start, order, n, stats.min_start,
                                               key_a);
        TH_CUDA(cudaGetLastError());
        TH_TRY(th_radix_sort(key_a, key_b, n, th_bits_for((u64)stats.max_start - (u64)stats.min_start),
                             th_bits_for((u64)stats.max_uuid), &sorted, &passes, st));
        passes_total += passes;
        }
        // 3. payload gather + sentinel positions: from here on the handle's own buffers are written
        TH_TRY(claim_buffers());
        ld_finalize_kernel<<<grid, 256, 0, st>>>(sorted, n, n_uuids, stats.min_start, (const i64 *)d_id,
                                                 (const i64 *)d_end, (const i64 *)d_owner, (const u8 *)d_canc, h->p1,
                                                 h->p3, h->p4, h->uuid, h->seg);
        TH_CUDA(cudaGetLastError());
    }
    h->max_id = new_max_id;
    size_t bytes = sizeof(int) * (size_t)h->cap_uuids + (sizeof(ThP1) * 4 + sizeof(u32)) * (size_t)h->cap_nt;
    if (nt > 0) {
        ld_sentinel_kernel<<<ugrid, 256, 0, st>>>(n_uuids, h->seg, h->p1, h->p2, h->p3, h->p4, h->uuid);
        TH_CUDA(cudaGetLastError());
        // 4. running maxima over every segment (sentinel = head of its segment)
        const int grid_nt = th_grid_for(nt, 256 * 4, 8);
        const i64 tiles = (nt + SC_TILE - 1) / SC_TILE;
        St *agg = nullptr;
        TH_TRY(sc.alloc((void **)&agg, sizeof(St) * (size_t)tiles));
        sc_local_kernel<<<(unsigned)tiles, SC_THREADS, 0, st>>>(h->uuid, h->p3, nt, h->p1, h->p2, h->p4, agg);
        TH_CUDA(cudaGetLastError());
        if (tiles > 1) {
            sc_carry_kernel<<<1, SC_THREADS, 0, st>>>(agg, tiles);
            sc_fix_kernel<<<grid_nt, SC_THREADS, 0, st>>>(h->uuid, h->seg, nt, agg, h->p1, h->p2, h->p4);
            TH_CUDA(cudaGetLastError());
        }
        // 5. time grid, one per uuid: plan -> scan of the cell counts -> cells
        i64 x4 = 6;
        if (const char *env = getenv("TH_GRID_CELLS_PER_ROW_X4")) { x4 = atoll(env); if (x4 < 1) x4 = 1; if (x4 > 5) x4 = x4 > 64 ? 64 : x4; }
        if (h->ugrid == nullptr || n_uuids > h->cap_ugrid) {
            if (h->ugrid) { TH_CUDA(cudaFree(h->ugrid)); h->ugrid = nullptr; }
            TH_CUDA(cudaMalloc((void **)&h->ugrid, sizeof(ThUGrid) * (size_t)(n_uuids > 0 ? n_uuids : 1)));
            h->cap_ugrid = n_uuids > 0 ? n_uuids : 1;
        }
        int *nb_u = nullptr;
        i64 *base = nullptr, *totals = nullptr;
        double *crowd = nullptr;
        TH_TRY(sc.alloc((void **)&nb_u, sizeof(int) * (size_t)(n_uuids > 0 ? n_uuids : 1)));
        TH_TRY(sc.alloc((void **)&base, 8 * (size_t)(n_uuids + 1)));
        TH_TRY(sc.alloc((void **)&totals, 16));
        TH_TRY(sc.alloc((void **)&crowd, sizeof(double)));
        TH_CUDA(cudaMemsetAsync(crowd, 0, sizeof(double), st));
        // the shared plan: one power-of-two width for all uuids over the global span, ~x4 / 4 cells per average row count
        int s_shift = 0;
        i64 s_nb = 1;
        {
            const u64 span = (u64)max_start - (u64)min_start;
            i64 target = n_uuids > 0 ? (x4 * ((n + n_uuids - 1) / n_uuids) + 3) / 4 : 1;
            if (target < 1) target = 1;
            while (s_shift < 63 && (span >> s_shift) >= (u64)target) ++s_shift;      // unsigned, see ld_ugrid_kernel
            s_nb = (i64)(span >> s_shift) + 1;
        }
        ld_ugrid_kernel<<<ugrid, 256, 0, st>>>(n_uuids, h->seg, h->p1, x4, s_shift, h->ugrid, nb_u, crowd);
        TH_CUDA(cudaGetLastError());
        TH_TRY(th_exclusive_scan_i32(nb_u, n_uuids, base, totals, sc, st));
        ld_ugrid_base_kernel<<<ugrid, 256, 0, st>>>(n_uuids, base, h->ugrid);
        i64 tot[2] = {0, 0};
        double crowd_h = 0.0;
        TH_CUDA(cudaMemcpyAsync(tot, totals, 16, cudaMemcpyDeviceToHost, st));
        TH_CUDA(cudaMemcpyAsync(&crowd_h, crowd, sizeof(double), cudaMemcpyDeviceToHost, st));
        TH_CUDA(cudaStreamSynchronize(st));
        i64 n_cells = tot[0];
        // balanced calendar (a random row shares its bucket with < 3 rows under the shared plan): keep the shared plan
        const char *force = getenv("TH_GRID_PLAN");           // diagnostics: "shared" / "per_uuid"
        bool shared = n > 0 && crowd_h / (double)n < 3.0 && n_uuids * s_nb < (1ll << 31) - 1;
        if (force && force[0] == 's' && n_uuids * s_nb < (1ll << 31) - 1) shared = true;
        if (force && force[0] == 'p') shared = false;
        h->grid_t0 = 0; h->grid_shift = 0; h->grid_nb = 0;
        if (shared) {
            ld_ugrid_shared_kernel<<<ugrid, 256, 0, st>>>(n_uuids, min_start, s_shift, (int)s_nb, h->ugrid);
            TH_CUDA(cudaGetLastError());
            n_cells = n_uuids * s_nb;
            h->grid_t0 = min_start; h->grid_shift = s_shift; h->grid_nb = (int)s_nb;
        }
        TH_REQUIRE(n_cells < (1ll << 31) - 1, "th_snapshot_load: time grid too large (%lld cells)", (long long)n_cells);
        int *first = nullptr;
        TH_TRY(sc.alloc((void **)&first, sizeof(int) * (size_t)(n_cells + 1)));
        if (n_cells > h->cap_cells || h->grid == nullptr) {
            if (h->grid) { TH_CUDA(cudaFree(h->grid)); h->grid = nullptr; }
            TH_CUDA(cudaMalloc((void **)&h->grid, sizeof(ThCell) * (size_t)(n_cells > 0 ? n_cells : 1)));
            h->cap_cells = n_cells > 0 ? n_cells : 1;
        }
        bytes += sizeof(ThCell) * (size_t)h->cap_cells + sizeof(ThUGrid) * (size_t)h->cap_ugrid;
        h->n_cells = n_cells;
        ld_grid_kernel<<<grid_nt, 256, 0, st>>>(h->p1, h->uuid, nt, n_cells, h->ugrid, first);
        ld_cells_kernel<<<th_grid_for(n_cells, 256 * 4, 8), 256, 0, st>>>(first, h->p1, n_cells, h->grid);
        TH_CUDA(cudaGetLastError());
    } else {
        h->n_cells = 0;
        h->grid_t0 = 0; h->grid_shift = 0; h->grid_nb = 0;
    }
    TH_CUDA(cudaStreamSynchronize(st));
    h->radix_passes = passes_total;
    h->n = n;
    h->nt = nt;
    h->device_bytes = (i64)bytes;
    h->loaded = 1;
    return TH_OK;
}

extern "C" int th_snapshot_load(th_snapshot *h, int64_t n, int64_t n_uuids, const int64_t *id, const int64_t *uuid_id,
                                const int64_t *start_us, const int64_t *end_us, const int64_t *owner,
                                const uint8_t *cancelled_u8, int mem, void *stream) {
    return th_load_impl(h, n, n_uuids, id, uuid_id, start_us, end_us, owner, cancelled_u8, mem, stream, -1);
}

extern "C" int th_snapshot_export(const th_snapshot *h, int64_t *id, int64_t *uuid_id, int64_t *start_us,
                                  int64_t *end_us, int64_t *owner, uint8_t *cancelled_u8, int64_t *seg_offsets,
                                  int mem, void *stream) {
    TH_REQUIRE(h != nullptr && h->loaded, "th_snapshot_export: snapshot not loaded");
    TH_REQUIRE(mem == TH_MEM_HOST || mem == TH_MEM_DEVICE, "th_snapshot_export: bad mem flag %d", mem);
    DeviceGuard guard(h->device);
    if (!guard.ok) { th_set_error("th_snapshot_export: cannot select device %d", h->device); return TH_ECUDA; }
    cudaStream_t st = th_call_stream(stream, mem, h->device);
    ThScratch sc(st);
    const size_t zn = (size_t)h->n, zu = (size_t)h->n_uuids + 1;
    void *d_id, *d_uuid, *d_start, *d_end, *d_owner, *d_canc, *d_seg;
    TH_TRY(sc.out(id, 8 * zn, mem, &d_id));
    TH_TRY(sc.out(uuid_id, 8 * zn, mem, &d_uuid));
    TH_TRY(sc.out(start_us, 8 * zn, mem, &d_start));
    TH_TRY(sc.out(end_us, 8 * zn, mem, &d_end));
    TH_TRY(sc.out(owner, 8 * zn, mem, &d_owner));
    TH_TRY(sc.out(cancelled_u8, zn, mem, &d_canc));
    TH_TRY(sc.out(seg_offsets, 8 * zu, mem, &d_seg));
    const i64 work = h->nt > h->n_uuids + 1 ? h->nt : h->n_uuids + 1;
    ex_export_kernel<<<th_grid_for(work, 256 * 4, 8), 256, 0, st>>>(h->nt, h->n_uuids, h->p1, h->p3, h->p4, h->uuid,
                                                                    h->seg, (i64 *)d_id, (i64 *)d_uuid, (i64 *)d_start,
                                                                    (i64 *)d_end, (i64 *)d_owner, (u8 *)d_canc,
                                                                    (i64 *)d_seg);
    TH_CUDA(cudaGetLastError());
    TH_TRY(th_copy_back(id, d_id, 8 * zn, mem, st));
    TH_TRY(th_copy_back(uuid_id, d_uuid, 8 * zn, mem, st));
    TH_TRY(th_copy_back(start_us, d_start, 8 * zn, mem, st));
    TH_TRY(th_copy_back(end_us, d_end, 8 * zn, mem, st));
    TH_TRY(th_copy_back(owner, d_owner, 8 * zn, mem, st));
    TH_TRY(th_copy_back(cancelled_u8, d_canc, zn, mem, st));
    TH_TRY(th_copy_back(seg_offsets, d_seg, 8 * zu, mem, st));
    if (mem == TH_MEM_HOST) TH_CUDA(cudaStreamSynchronize(st));
    return TH_OK;
}


// Snapshot maintenance without re-sending the table (the reference commits every write
// through CRUDModel.save / destroy, tensorhive/models/CRUDModel.py:21-50): `dst` becomes `src` minus the rows whose
// id is upserted or deleted, plus the upserted rows.  Only the delta crosses PCIe; survivors are gathered from the
// source snapshot on the device in their sorted order, the upserted rows are sorted (K1 on the delta only) and merged in
// with one binary search per record, and the index is rebuilt.  dst may be src.
extern "C" int th_snapshot_apply_delta(th_snapshot *dst, const th_snapshot *src, int64_t n_uuids, int64_t n_upsert,
                                       const int64_t *id, const int64_t *uuid_id, const int64_t *start_us,
                                       const int64_t *end_us, const int64_t *owner, const uint8_t *cancelled_u8,
                                       int64_t n_delete, const int64_t *delete_ids, int mem, void *stream) {
    TH_REQUIRE(dst != nullptr && src != nullptr && src->loaded, "th_snapshot_apply_delta: source snapshot not loaded");
    TH_REQUIRE(dst->device == src->device, "th_snapshot_apply_delta: snapshots live on different devices");
    TH_REQUIRE(n_upsert >= 0 && n_delete >= 0, "th_snapshot_apply_delta: negative size");
    TH_REQUIRE(n_uuids >= src->n_uuids, "th_snapshot_apply_delta: n_uuids=%lld is smaller than the source's %lld",
               (long long)n_uuids, (long long)src->n_uuids);
    TH_REQUIRE(mem == TH_MEM_HOST || mem == TH_MEM_DEVICE, "th_snapshot_apply_delta: bad mem flag %d", mem);
    TH_REQUIRE(n_upsert == 0 || (id && uuid_id && start_us && end_us && owner && cancelled_u8),
               "th_snapshot_apply_delta: a column pointer is NULL");
    TH_REQUIRE(n_delete == 0 || delete_ids, "th_snapshot_apply_delta: delete_ids is NULL");
    DeviceGuard guard(src->device);
    if (!guard.ok) { th_set_error("th_snapshot_apply_delta: cannot select device %d", src->device); return TH_ECUDA; }
    cudaStream_t st = th_call_stream(stream, mem, src->device);
    ThScratch sc(st);
    const size_t zu = (size_t)n_upsert, zd = (size_t)n_delete;
    const void *u_id, *u_uuid, *u_start, *u_end, *u_owner, *u_canc, *d_del;
    TH_TRY(sc.in(id, 8 * zu, mem, &u_id));
    TH_TRY(sc.in(uuid_id, 8 * zu, mem, &u_uuid));
    TH_TRY(sc.in(start_us, 8 * zu, mem, &u_start));
    TH_TRY(sc.in(end_us, 8 * zu, mem, &u_end));
    TH_TRY(sc.in(owner, 8 * zu, mem, &u_owner));
    TH_TRY(sc.in(cancelled_u8, zu, mem, &u_canc));
    TH_TRY(sc.in(delete_ids, 8 * zd, mem, &d_del));
    // ids that leave the old table, sorted for the per-row membership test
    const i64 n_kill = n_upsert + n_delete;
    ThRec *ka = nullptr, *kb = nullptr, *kill = nullptr;
    TH_TRY(sc.alloc((void **)&ka, sizeof(ThRec) * (size_t)(n_kill > 0 ? n_kill : 1)));
    TH_TRY(sc.alloc((void **)&kb, sizeof(ThRec) * (size_t)(n_kill > 0 ? n_kill : 1)));
    kill = ka;
    if (n_kill > 0) {
        dl_killrec_kernel<<<th_grid_for(n_kill, 256, 8), 256, 0, st>>>((const i64 *)u_id, n_upsert, (const i64 *)d_del,
                                                                       n_delete, ka);
        TH_CUDA(cudaGetLastError());
        int passes = 0;
        TH_TRY(th_radix_sort(ka, kb, n_kill, 64, 0, &kill, &passes, st));      // as u64: a negative id sorts last, matches nothing
    }
    // ids must be unique over the whole delta: two upserts of one id would both land in the table, and an id that is
    // upserted AND deleted has no defined order (callers collapse their batch first; Calendar.apply_delta does)
    int *d_dup = nullptr;
    TH_TRY(sc.alloc((void **)&d_dup, sizeof(int)));
    TH_CUDA(cudaMemsetAsync(d_dup, 0, sizeof(int), st));
    if (n_kill > 1) {
        ld_dup_kernel<<<th_grid_for(n_kill, 256, 8), 256, 0, st>>>(kill, n_kill, d_dup);
        TH_CUDA(cudaGetLastError());
    }
    // survivors
    const i64 nt = src->nt;
    int *keep = nullptr;
    i64 *off = nullptr, *totals = nullptr;
    TH_TRY(sc.alloc((void **)&keep, 4 * (size_t)(nt > 0 ? nt : 1)));
    TH_TRY(sc.alloc((void **)&off, 8 * (size_t)(nt + 1)));
    TH_TRY(sc.alloc((void **)&totals, 16));
    if (nt > 0) {
        dl_keep_kernel<<<th_grid_for(nt, 256 * 4, 8), 256, 0, st>>>(nt, src->uuid, src->seg, src->p3, kill, n_kill, keep);
        TH_CUDA(cudaGetLastError());
    }
    TH_TRY(th_exclusive_scan_i32(keep, nt, off, totals, sc, st));
    i64 tot[2] = {0, 0};
    int dup = 0;
    TH_CUDA(cudaMemcpyAsync(tot, totals, 16, cudaMemcpyDeviceToHost, st));
    TH_CUDA(cudaMemcpyAsync(&dup, d_dup, sizeof(int), cudaMemcpyDeviceToHost, st));
    TH_CUDA(cudaStreamSynchronize(st));
    TH_REQUIRE(dup == 0, "th_snapshot_apply_delta: an id appears more than once in the delta (upserts + deletes)");
    const i64 n_keep = tot[0];
    const i64 n_new = n_keep + n_upsert;
    TH_REQUIRE(n_new < (1ll << 30), "th_snapshot_apply_delta: %lld rows exceed the 2^30 limit", (long long)n_new);
    const size_t zn = (size_t)(n_new > 0 ? n_new : 1);
    i64 *c_id, *c_uuid, *c_start, *c_end, *c_owner;
    u8 *c_canc;
    TH_TRY(sc.alloc((void **)&c_id, 8 * zn));
    TH_TRY(sc.alloc((void **)&c_uuid, 8 * zn));
    TH_TRY(sc.alloc((void **)&c_start, 8 * zn));
    TH_TRY(sc.alloc((void **)&c_end, 8 * zn));
    TH_TRY(sc.alloc((void **)&c_owner, 8 * zn));
    TH_TRY(sc.alloc((void **)&c_canc, zn));
    if (nt > 0) {
        dl_gather_kernel<<<th_grid_for(nt, 256 * 4, 8), 256, 0, st>>>(nt, keep, off, src->p1, src->p3, src->p4, src->uuid, c_id,
                                                                      c_uuid, c_start, c_end, c_owner, c_canc);
        TH_CUDA(cudaGetLastError());
    }
    if (n_upsert > 0) {
        dl_append_kernel<<<th_grid_for(n_upsert, 256, 8), 256, 0, st>>>(n_upsert, n_keep, (const i64 *)u_id, (const i64 *)u_uuid,
                                                                        (const i64 *)u_start, (const i64 *)u_end,
                                                                        (const i64 *)u_owner, (const u8 *)u_canc, c_id, c_uuid,
                                                                        c_start, c_end, c_owner, c_canc);
        TH_CUDA(cudaGetLastError());
    }
    // the gathers above read `src`; if dst == src the rebuild below overwrites it only after them (stream order)
    // survivors come out of the source in (uuid, start, id) order: only the upserted rows are sorted, then merged in
    return th_load_impl(dst, n_new, n_uuids, (const int64_t *)c_id, (const int64_t *)c_uuid, (const int64_t *)c_start,
                        (const int64_t *)c_end, (const int64_t *)c_owner, c_canc, TH_MEM_DEVICE, (void *)st, n_keep);
}
