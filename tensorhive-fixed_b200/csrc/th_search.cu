// K2 -- batched interval-overlap search over a loaded snapshot (sm_100a).
//
// Replaces (reference, tensorhive/models/Reservation.py):
//   would_interfere :120-130                 -> th_conflict_bitmap(TH_MODE_STRICT), th_conflict_sets
//   filter_by_uuids_and_time_range :132-148  -> th_conflict_sets(TH_MODE_INCLUSIVE)
//   current_events :89-104                   -> th_conflict_sets(TH_MODE_INCLUSIVE, qs = qe = now)
//   upcoming_events_for_resource :106-118    -> th_conflict_sets(TH_MODE_UPCOMING)
//
// Bitmap kernel: the rows of a uuid with start < qe are a prefix of its segment, so the verdict is "does the
// running maximum m1 of non-cancelled `end` over that prefix exceed qs".  The prefix end is found through the
// snapshot's time grid (th_common.cuh): one 32-byte cell, then -- rarely -- a handful of 16-byte {start, m1} rows.
// One lane per query, one ballot per warp = one 32-bit bitmap word per warp, stored coalesced.
//
// Set kernels: one warp per query.  The warp walks the candidate prefix backwards 32 rows at a time with
// 16-byte loads and stops as soon as the running maximum over the remaining prefix says no further row
// can match; a second pass re-walks the examined range forwards and compacts matches with ballots.
#include "th_common.cuh"
#include <stdlib.h>
#include <string.h>
#include <atomic>

namespace {

constexpr int BM_THREADS = 256;
#ifndef BM_MIN_BLOCKS
#define BM_MIN_BLOCKS 4
#endif
#ifndef BM_PIPE
#define BM_PIPE 0
#endif
#ifndef BM_SEND_FENCE
#define BM_SEND_FENCE 0
#endif
#ifndef BM_CELL_LD      // how the 32-byte cell is read: 0 = one LDG.256, 1 = two LDG.128, 2 = LDG.128 + second half on demand
#define BM_CELL_LD 0
#endif
constexpr int BM_MAXW = 2048;   // bitmap words a CTA buffers for the fused exchange (8 KB)

// One lane per query, persistent warps.  Two things bound this kernel, and neither is bytes: (1) a memory
// instruction whose 32 lanes touch 32 different cache lines costs 32 L1 wavefronts and 32 L2 sector
// requests, and (2) every DEPENDENT round trip to L2/HBM adds ~1 us to the warp, whose latency is the maximum
// over its lanes.  So the lookup is three levels deep, each level issued for all lanes at once with
// predicated (not looped, not short-circuited) loads, and level 1 runs several tiles ahead:
//   level 1  a warp tile = 32 queries x 4 columns (1 KB) is staged in shared memory by coalesced 16-byte
//            cp.async copies (two per lane) into a per-warp 4-stage ring; warps never synchronise with each
//            other.  (TMA bulk copies, one mbarrier per stage, were measured too: same latency hiding but ~80
//            more warp instructions per tile for issue + barrier handling, and the kernel is issue bound);
//   level 2  ONE 32-byte grid cell {pos, cnt, m1_before, start0, m1_0}, a single 256-bit load (LDG.E.256).  m1 only
//            grows along a segment, so if the rows before the bucket already end after qs the verdict is "conflict";
//            buckets with at most one row (5 of 6 at ~1.5 cells per row) are settled by the cell as well;
//   level 3  rows 1..3 of the bucket as 16-byte {start, m1} probes, and, for self-updates whose prefix maximum
//            exceeds qs, the {a1, m2} entries of the rows the answer can land on.
// Buckets with more than four rows fall back to a binary search over the rest (about 1 % of the lanes).
constexpr int BM_WARPS = BM_THREADS / 32;
constexpr int BM_STAGES = 4;
constexpr int BM_PROBE = 4;

// level 2: the grid cell of (u, qe); a query for an unknown uuid reads cell 0 and is masked later
__device__ __forceinline__ const ThCell *k2_cell_of(const ThGrid &g, i64 n_uuids, i64 u, i64 qe) {
    return th_cell_of(g, (u < 0 || u >= n_uuids) ? 0 : u, qe);
}

// level 3: verdict from the cell and (if needed) the bucket's rows.  The cell carries the bucket's first row, so the
// row arrays are touched only when (a) the bucket holds two or more rows, its first row starts before qe and the running
// maximum up to that row does not settle the answer, or (b) the candidate is a self-update whose prefix maximum exceeds
// qs, so that the {a1, m2} entry of the landing row has to say whether that maximum is the candidate's own row.
template <int MODE>
__device__ __forceinline__ bool k2_resolve(const ThCellV c, const ThP1 *__restrict__ p1, const ThP2 *__restrict__ p2,
                                           const ThP4 *__restrict__ p4, i64 qs, i64 qe, i64 self) {
    const int pos = (int)(c.pc & 0xffffffffll);
    const int cnt = (int)(c.pc >> 32);
    static_assert(BM_PROBE == 4, "the probe offsets below are written out for four rows");
    if (MODE == TH_MODE_STRICT) {
        const bool ws = self >= 0;
        if (!ws && c.m1_before > qs) return true;     // decided by the rows before the bucket
        const bool in0 = cnt > 0 && c.start0 < qe;    // rows are start-sorted: the ones before qe are a prefix
        const bool more = in0 && cnt > 1;             // rows 1.. of the bucket may lie before qe as well
        const i64 m1c = in0 ? c.m1_0 : c.m1_before;   // running maximum as far as the cell knows it
        // m1 only grows along a segment: once it exceeds qs a candidate without self-exclusion is settled
        const bool rows = more && (ws || m1c <= qs);
        const bool two = ws && (more || m1c > qs);    // {a1, m2} wanted (m1 <= qs is "no conflict" for self-updates too)
        const longlong2 far = make_longlong2(TH_I64_MAX, 0);          // {start, m1}: never before qe
        const longlong2 none = make_longlong2(-1, TH_I64_MIN);       // {a1, m2}
        longlong2 e[BM_PROBE], t[BM_PROBE + 1];
        const ThP1 *r1 = p1 + pos;             // one address per array; the probes use immediate offsets from it
        const ThP2 *r2 = p2 + pos;
        e[1] = th_ld16_if_off<16>(r1, rows, far);
        e[2] = th_ld16_if_off<32>(r1, rows && cnt > 2, far);
        e[3] = th_ld16_if_off<48>(r1, rows && cnt > 3, far);
        t[0] = th_ld16_if_off<-16>(r2, two && !in0 && pos > 0, none);
        t[1] = th_ld16_if_off<0>(r2, two && in0, none);
        t[2] = th_ld16_if_off<16>(r2, two && more, none);
        t[3] = th_ld16_if_off<32>(r2, two && more && cnt > 2, none);
        t[4] = th_ld16_if_off<48>(r2, two && more && cnt > 3, none);
        int k = in0 ? 1 : 0;
        i64 m1 = m1c;
        longlong2 tt = in0 ? t[1] : t[0];
#pragma unroll
        for (int j = 1; j < BM_PROBE; ++j)
            if (e[j].x < qe) { k = j + 1; m1 = e[j].y; tt = t[j + 1]; }
        if (k == BM_PROBE && cnt > BM_PROBE) {        // long bucket: finish with a binary search
            int lo = pos + BM_PROBE, hi = pos + cnt;
            while (lo < hi) {
                const int mid = (int)(((unsigned)lo + (unsigned)hi) >> 1);
                const longlong2 x = th_ld16(&p1[mid]);
                if (x.x < qe) { lo = mid + 1; m1 = x.y; } else hi = mid;
            }
            if (ws) tt = th_ld16(&p2[lo - 1]);
        }
        // the last row with start < qe is at worst the uuid's sentinel (m1 = INT64_MIN)
        return (ws && tt.x == self) ? (tt.y > qs) : (m1 > qs);
    } else {   // TH_MODE_INCLUSIVE: any row, cancelled or not, with start <= qe and qs <= end
        // mall (running maximum of `end` over ALL rows, P4) is at least m1 (non-cancelled rows only), so the cell's
        // m1_before settles "yes" early; otherwise speculative probes on {start} (row 0 from the cell) and {mall}.
        if (qs == TH_I64_MIN) {
            // Every `end` qualifies, so the question is "does a row start at or before qe" -- and the sentinel's end of
            // -inf, which the comparisons below would accept, must not count as one.  Rare: a plain search of the bucket.
            int lo = pos, hi = pos + cnt;
            while (lo < hi) {
                const int mid = (int)(((unsigned)lo + (unsigned)hi) >> 1);
                if (th_ld8(&p1[mid].start) <= qe) lo = mid + 1; else hi = mid;
            }
            // row lo - 1 is the last one with start <= qe: a real row, or the uuid's sentinel (start = -inf)
            return lo > 0 && th_ld8(&p1[lo - 1].start) != TH_I64_MIN;
        }
        if (c.m1_before >= qs) return true;
        const bool in0 = cnt > 0 && c.start0 <= qe;
        const bool more = in0 && cnt > 1;
        const longlong2 far = make_longlong2(TH_I64_MAX, 0);
        const longlong2 low = make_longlong2(-1, TH_I64_MIN);          // {owner, mall}
        const ThP1 *r1 = p1 + pos;
        const ThP4 *r4 = p4 + pos;
        longlong2 e[BM_PROBE], w[BM_PROBE + 1];
        e[1] = th_ld16_if_off<16>(r1, more, far);
        e[2] = th_ld16_if_off<32>(r1, more && cnt > 2, far);
        e[3] = th_ld16_if_off<48>(r1, more && cnt > 3, far);
        w[0] = th_ld16_if_off<-16>(r4, !in0 && pos > 0, low);
        w[1] = th_ld16_if_off<0>(r4, in0, low);
        w[2] = th_ld16_if_off<16>(r4, more, low);
        w[3] = th_ld16_if_off<32>(r4, more && cnt > 2, low);
        w[4] = th_ld16_if_off<48>(r4, more && cnt > 3, low);
        int k = in0 ? 1 : 0;
        i64 mall = in0 ? w[1].y : w[0].y;
#pragma unroll
        for (int j = 1; j < BM_PROBE; ++j)      // (more && cnt > j: for qe = INT64_MAX the default of a row that was not probed is "before qe" too)
            if (more && cnt > j && e[j].x <= qe) { k = j + 1; mall = w[j + 1].y; }
        if (k == BM_PROBE && cnt > BM_PROBE) {        // long bucket: finish with a binary search
            int lo = pos + BM_PROBE, hi = pos + cnt;
            while (lo < hi) {
                const int mid = (int)(((unsigned)lo + (unsigned)hi) >> 1);
                if (th_ld8(&p1[mid].start) <= qe) lo = mid + 1; else hi = mid;
            }
            // lo >= 1: bucket 0 starts with the sentinel, later buckets are preceded by it
            mall = th_ld8(&p4[lo - 1].mall);
        }
        return mall >= qs;
    }
}

// PACKED = the wire format of th_conflict_bitmap_packed: uuid codes and self ids as uint32 columns (0xFFFFFFFF = no self
// id; q_self may be NULL = no candidate is a self-update), 24 or 20 bytes per query instead of 32.
template <int MODE, bool PACKED = false>
__global__ void __launch_bounds__(BM_THREADS, BM_MIN_BLOCKS) k2_bitmap_kernel(
    ThGrid g, const ThP1 *__restrict__ p1, const ThP2 *__restrict__ p2, const ThP4 *__restrict__ p4, i64 n_uuids, i64 m,
    const void *__restrict__ q_uuid_v, const i64 *__restrict__ q_start, const i64 *__restrict__ q_end,
    const void *__restrict__ q_self_v, int cols_aligned16, u32 *__restrict__ out_bits, ThPeerOut po) {
    __shared__ __align__(128) i64 sq[BM_WARPS][BM_STAGES][4][32];
    const i64 *q_uuid = static_cast<const i64 *>(q_uuid_v), *q_self = static_cast<const i64 *>(q_self_v);
    const u32 *q_uuid32 = static_cast<const u32 *>(q_uuid_v), *q_self32 = static_cast<const u32 *>(q_self_v);
    const bool with_self = MODE == TH_MODE_STRICT && (!PACKED || q_self_v != nullptr);
    constexpr int NCOL = MODE == TH_MODE_STRICT ? 4 : 3;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const i64 full_tiles = cols_aligned16 ? m / 32 : 0;      // warp tiles that are staged through shared memory
    const i64 tiles = (m + 31) / 32;
    // Each CTA owns a CONTIGUOUS range of tiles [t_begin, t_end); its 8 warps interleave inside it (warp w takes
    // t_begin + w, + 8, + 16, ...), so a CTA still reads 8 KB of contiguous query columns per round, and its bitmap
    // words form one contiguous run that the fused exchange can ship to the peers in 128-byte packets.
    const i64 per_cta = (tiles + gridDim.x - 1) / gridDim.x;
    const i64 t_begin = (i64)blockIdx.x * per_cta;
    const i64 t_end = t_begin + per_cta < tiles ? t_begin + per_cta : tiles;
    __shared__ u32 swords[BM_MAXW];
    const bool buffered = po.n > 0 && per_cta <= BM_MAXW;
    auto tile_of = [&](u32 it) -> i64 { return t_begin + warp + (i64)BM_WARPS * it; };
    u64 *const trace = po.n > 0 ? po.trace : nullptr;
    if (trace && threadIdx.x == 0) trace[4 * blockIdx.x + 0] = th_globaltimer_ns();

    // level 1: a warp tile (32 queries x 4 columns = 1 KB = 64 chunks of 16 bytes) is staged with two cp.async per
    // lane -- coalesced, 16 bytes each, no registers -- into a per-warp 4-stage ring; one commit group per tile, and
    // an (empty) group is committed even when there is nothing left to prefetch so that "all but the newest
    // BM_STAGES - 1 groups have landed" always means "this tile has landed".
    auto stage_in = [&](i64 tile, int st) {
        if (tile < t_end && tile < full_tiles) {
            const i64 base = tile * 32;
            if (!PACKED) {
#pragma unroll
                for (int c = lane; c < NCOL * 16; c += 32) {
                    const int col = c >> 4, off = (c & 15) * 2;      // 16 chunks per column, 2 elements per chunk
                    const i64 *src = (col == 0 ? q_uuid : col == 1 ? q_end : col == 2 ? q_start : q_self) + base + off;
                    th_cp_async16(&sq[warp][st][col][off], src);
                }
            } else {
                // 48 chunks: end (16), start (16), uuid u32 (8), self u32 (8, skipped when there is no self column)
                th_cp_async16(&sq[warp][st][lane < 16 ? 1 : 2][(lane & 15) * 2],
                              (lane < 16 ? q_end : q_start) + base + (lane & 15) * 2);
                if (lane < 8) th_cp_async16(reinterpret_cast<u32 *>(&sq[warp][st][0][0]) + lane * 4, q_uuid32 + base + lane * 4);
                else if (lane < 16 && with_self)
                    th_cp_async16(reinterpret_cast<u32 *>(&sq[warp][st][3][0]) + (lane - 8) * 4, q_self32 + base + (lane - 8) * 4);
            }
        }
        th_cp_async_commit();
    };
    auto fetch = [&](i64 tile, u32 it, i64 &u, i64 &qs, i64 &qe, i64 &self, ThCellV &cell) {
        const int st = it % BM_STAGES;
        if (tile < full_tiles) {
            th_cp_async_wait<BM_STAGES - 1>();
            __syncwarp();                                   // the other lanes' chunks are visible too
            qe = sq[warp][st][1][lane];
            qs = sq[warp][st][2][lane];
            if (!PACKED) {
                u = sq[warp][st][0][lane];
                self = MODE == TH_MODE_STRICT ? sq[warp][st][3][lane] : -1;
            } else {
                u = (i64)reinterpret_cast<const u32 *>(&sq[warp][st][0][0])[lane];
                const u32 f = with_self ? reinterpret_cast<const u32 *>(&sq[warp][st][3][0])[lane] : 0xffffffffu;
                self = f == 0xffffffffu ? -1 : (i64)f;
            }
            __syncwarp();                                   // every lane has read the stage: refill it
        } else {   // ragged last tile or unaligned caller buffers
            const i64 i = tile * 32 + lane;
            const i64 ic = i < m ? i : m - 1;
            qe = th_ld_stream(&q_end[ic]);
            qs = th_ld_stream(&q_start[ic]);
            if (!PACKED) {
                u = th_ld_stream(&q_uuid[ic]);
                self = MODE == TH_MODE_STRICT ? th_ld_stream(&q_self[ic]) : -1;
            } else {
                u = (i64)__ldg(&q_uuid32[ic]);
                const u32 f = with_self ? __ldg(&q_self32[ic]) : 0xffffffffu;
                self = f == 0xffffffffu ? -1 : (i64)f;
            }
        }
        stage_in(tile_of(it + BM_STAGES), st);
#if BM_CELL_LD == 0
        cell = n_uuids > 0 ? th_ld32(k2_cell_of(g, n_uuids, u, qe)) : ThCellV{0, 0, 0, 0};   // no uuids: no grid
#else
        {
            const ThCell *cp = n_uuids > 0 ? k2_cell_of(g, n_uuids, u, qe) : g.cells;
            const longlong2 a = th_ld16_if_off<0>(cp, n_uuids > 0, make_longlong2(0, 0));
#if BM_CELL_LD == 1
            const longlong2 b = th_ld16_if_off<16>(cp, n_uuids > 0, make_longlong2(0, 0));
#else
            // the first row only matters when the bucket has one and the rows before it do not already settle the verdict
            const bool want = n_uuids > 0 && (int)(a.x >> 32) > 0 && (MODE != TH_MODE_STRICT || self >= 0 || a.y <= qs);
            const longlong2 b = th_ld16_if_off<16>(cp, want, make_longlong2(TH_I64_MAX, a.y));
#endif
            cell.pc = a.x; cell.m1_before = a.y; cell.start0 = b.x; cell.m1_0 = b.y;
        }
#endif
    };

#pragma unroll
    for (int st = 0; st < BM_STAGES; ++st) stage_in(tile_of(st), st);
    // Measured on B200 (tools/k2_sweep.py): requesting the cell of tile k + 1 before the rows of tile k did not pay,
    // neither in plain C++ (ptxas sinks the prefetch again: 26.6 us vs 24.5 us cold at 1 M queries) nor as ordered
    // volatile PTX (24.8 us; 264 us vs 217 us at 16 M queries).  At ~320 warp instructions per 32 queries the kernel
    // is bound by issue slots, and the pipelined loop adds instructions; so each tile runs levels 2 and 3 in turn.
    // Also measured and dropped: pulling the index (cells, P1, P2 = 44 MB) into L2 with cp.async.bulk.prefetch.L2 at
    // kernel start -- 26.8 us vs 24.6 us cold: the prefetch competes with the query stream for HBM and the probes'
    // HBM latency was not what the warps were waiting for.
    // Also measured and dropped: compacting the lanes that need level 3 into a per-warp shared-memory queue and
    // running the row probes only on full batches of 32 (dense lanes, level 3 once per ~3 tiles): correct, but
    // 43.9 us cold / 483 us at 16 M queries -- the queue traffic, the extra ballots and atomicOr of late bits cost
    // far more issue slots than the idle lanes they save.
    // Also measured and dropped: level 2 of tile k + 1 under level 3 of tile k WITHOUT registers or scoreboards -- each
    // lane copies its cell into shared memory with a 16-byte cp.async gather, one commit group behind the query stage:
    // 24.5 us cold (same), 15.5 vs 14.5 us warm, 209 vs 213 us at 16 M queries, and 421 vs 375 us on 12.5 M rows x
    // 12.5 M queries, where every probe misses L2.  More probes in flight do not help: at that size the kernel moves
    // ~3 TB/s of scattered 32-byte sectors (about 2.7 per query), about what HBM delivers for this access pattern.
    // Also measured and dropped: a 64-byte record per grid cell (header + first three rows, one HBM atom per query)
    // read by a one-level fast path that defers self-updates with an unfiltered conflict and long buckets (~7 % of
    // the lanes) to a per-CTA queue resolved densely after the CTA's tiles: 32.5 us cold / 24 us warm, and 503 vs 387 us
    // on 12.5 M x 12.5 M -- the extra loads per lane and the thin, latency-bound tail cost more than the level saved.
    // Also measured and dropped: probes without default values (every use guarded by the probe's own predicate, saving
    // ~36 moves per tile) and per-lane precomputed staging addresses: ptxas then needs 64 registers plus 56 bytes of
    // spill (predicate pressure), 28.6 us cold / 19.8 us warm / 279 us at 16 M queries.
    auto emit = [&](i64 tile, bool hit) {
        const u32 word = __ballot_sync(0xffffffffu, hit);
        if (lane == 0) {
            if (out_bits) out_bits[tile] = word;
            if (buffered) swords[tile - t_begin] = word;
            else if (po.n > 0) {                                       // very large batches: one slot store per tile
                const u64 slot = (u64)word | ((u64)po.epoch << 32);
                for (int p = 0; p < po.n; ++p) po.dst[p][po.word_off + tile] = slot;
            }
        }
    };
#if BM_PIPE
    // Two tiles in flight per warp: the cell of tile k + 1 is requested before the rows of tile k are probed, so an
    // iteration exposes one round trip to L2 / HBM instead of two.  Only the next tile's CELL lives in registers across
    // the iteration: its query is read again from the shared-memory stage when its turn comes (a stage is refilled only
    // after its tile has been answered).  The loads are volatile asm, which keeps their order.
    {
        auto landed = [&](i64 tile) {                   // tile's stage is complete and visible to the whole warp
            if (tile < full_tiles) { th_cp_async_wait<BM_STAGES - 2>(); __syncwarp(); }
        };
        auto read_q = [&](i64 tile, int st, i64 &u, i64 &qs, i64 &qe, i64 &self) {
            if (tile < full_tiles) {
                qe = sq[warp][st][1][lane];
                qs = sq[warp][st][2][lane];
                if (!PACKED) {
                    u = sq[warp][st][0][lane];
                    self = MODE == TH_MODE_STRICT ? sq[warp][st][3][lane] : -1;
                } else {
                    u = (i64)reinterpret_cast<const u32 *>(&sq[warp][st][0][0])[lane];
                    const u32 f = with_self ? reinterpret_cast<const u32 *>(&sq[warp][st][3][0])[lane] : 0xffffffffu;
                    self = f == 0xffffffffu ? -1 : (i64)f;
                }
            } else {
                const i64 i = tile * 32 + lane;
                const i64 ic = i < m ? i : m - 1;
                qe = th_ld_stream(&q_end[ic]);
                qs = th_ld_stream(&q_start[ic]);
                if (!PACKED) {
                    u = th_ld_stream(&q_uuid[ic]);
                    self = MODE == TH_MODE_STRICT ? th_ld_stream(&q_self[ic]) : -1;
                } else {
                    u = (i64)__ldg(&q_uuid32[ic]);
                    const u32 f = with_self ? __ldg(&q_self32[ic]) : 0xffffffffu;
                    self = f == 0xffffffffu ? -1 : (i64)f;
                }
            }
        };
        auto cell_for = [&](i64 tile, int st) -> ThCellV {
            i64 u, qs, qe, self;
            read_q(tile, st, u, qs, qe, self);
            return n_uuids > 0 ? th_ld32(k2_cell_of(g, n_uuids, u, qe)) : ThCellV{0, 0, 0, 0};
        };
        u32 it = 0;
        i64 tile = tile_of(0);
        bool have = tile < t_end;
        ThCellV cell = {0, 0, 0, 0};
        if (have) { landed(tile); cell = cell_for(tile, 0); }
        while (have) {
            const int st = it % BM_STAGES;
            const i64 ntile = tile_of(it + 1);
            const bool nhave = ntile < t_end;
            ThCellV ncell = {0, 0, 0, 0};
            if (nhave) { landed(ntile); ncell = cell_for(ntile, (it + 1) % BM_STAGES); }
            i64 u, qs, qe, self;
            read_q(tile, st, u, qs, qe, self);
            const i64 i = tile * 32 + lane;
            const bool hit = (i < m) && (u >= 0 && u < n_uuids) && qe > TH_I64_MIN &&
                             k2_resolve<MODE>(cell, p1, p2, p4, qs, qe, self);
            emit(tile, hit);
            __syncwarp();                               // every lane has read this stage for the last time: refill it
            stage_in(tile_of(it + BM_STAGES), st);
            cell = ncell; tile = ntile; have = nhave; ++it;
        }
    }
#else
    u32 it = 0;
    for (i64 tile = tile_of(0); tile < t_end; tile = tile_of(++it)) {
        i64 u, qs, qe, self;
        ThCellV cell;
        fetch(tile, it, u, qs, qe, self, cell);
        const i64 i = tile * 32 + lane;
        // (qe = INT64_MIN cannot lie after any start, not even the sentinel's: no row, no conflict)
        const bool hit = (i < m) && (u >= 0 && u < n_uuids) && qe > TH_I64_MIN &&
                         k2_resolve<MODE>(cell, p1, p2, p4, qs, qe, self);
        emit(tile, hit);
    }
#endif
    if (buffered) {
        // Fused exchange: the CTA's run of words goes straight into every rank's slot array over NVLink as coalesced
        // 8-byte {word, epoch} stores (256-byte packets; a 4-byte store per word and tile was measured at +3 us per
        // extra peer on the whole kernel).  Nothing else is needed: the receiver sees each slot turn to this epoch.
        __syncthreads();
        if (trace && threadIdx.x == 0) trace[4 * blockIdx.x + 1] = th_globaltimer_ns();
        const int cnt = (int)(t_end > t_begin ? t_end - t_begin : 0);
        const u64 tag = (u64)po.epoch << 32;
        for (int p = 0; p < po.n; ++p) {
            u64 *dst = po.dst[p] + po.word_off + t_begin;
            for (int j = threadIdx.x; j < cnt; j += BM_THREADS) dst[j] = (u64)swords[j] | tag;
        }
#if BM_SEND_FENCE
        // not needed for correctness (a slot is one atomic 8-byte store); pushes the posted peer writes out at once
        if (threadIdx.x < cnt) __threadfence_system();
#endif
    }
    if (trace && threadIdx.x == 0) trace[4 * blockIdx.x + 2] = th_globaltimer_ns();
    if (po.n > 0 && po.unpack) {
        // Receive half, fused: a CTA that has sent its own words turns to the SAME word range of every rank.  The ranks
        // run the same grid over the same tile ranges, so the peer CTA that owns this range finished at about the same
        // time: the wait is one NVLink flight, not the slowest CTA of the slowest rank (polling the whole slot array from
        // every CTA cost 6-10 us per step: the early finishers' polls competed with the CTAs still answering queries).
        const i64 per_r = (po.recv_words + gridDim.x - 1) / gridDim.x;
        const i64 r_begin = (i64)blockIdx.x * per_r;
        const i64 r_cnt = r_begin < po.recv_words ? (r_begin + per_r < po.recv_words ? per_r : po.recv_words - r_begin) : 0;
        // (32-bit index arithmetic: a 64-bit division per slot cost more than the poll itself on small batches)
        const u32 rc32 = (u32)r_cnt, total32 = (u32)po.n * rc32;
        for (u32 k = threadIdx.x; k < total32; k += BM_THREADS) {
            const u32 r = k / rc32;
            const i64 j = r_begin + (k - r * rc32);
            const i64 idx = (i64)r * po.wpr + j;
            u32 w = 0;
            if (j < po.expect[r]) {
                u64 v = po.my_slots[idx];
                u64 t0 = 0;
                for (u32 spins = 0; (u32)(v >> 32) != po.epoch; ++spins) {
                    if ((spins & 63u) == 63u) {                    // bounded by wall time, checked every 64 polls
                        const u64 now = th_globaltimer_ns();
                        if (t0 == 0) t0 = now;
                        else if (now - t0 > po.timeout_ns) { *po.err = 1; v = 0; break; }   // late peer: word reads 0
                    }
                    __nanosleep(64);
                    v = po.my_slots[idx];
                }
                w = (u32)v;
            }
            po.my_out[idx] = w;
        }
        __syncthreads();
        if (trace && threadIdx.x == 0) trace[4 * blockIdx.x + 3] = th_globaltimer_ns();
        // slots beyond what any rank contributes this step read as 0
        const i64 tail = po.wpr - po.recv_words;
        for (int r = 0; r < po.n; ++r)
            for (i64 k = (i64)blockIdx.x * BM_THREADS + threadIdx.x; k < tail; k += (i64)gridDim.x * BM_THREADS)
                po.my_out[(i64)r * po.wpr + po.recv_words + k] = 0;
    }
}

// receivers: poll every expected slot until it carries this step's epoch, and write the compact words the caller
// reads (bounded spin; sets *err on timeout).  Slots a rank does not use this step read as 0.
__global__ void __launch_bounds__(256) k2_ll_unpack_kernel(const volatile u64 *__restrict__ slots, u32 *__restrict__ out,
                                                           int world, i64 wpr, ThPeerCounts counts, u32 epoch,
                                                           u64 timeout_ns, int *err) {
    const i64 total = (i64)world * wpr;
    for (i64 idx = (i64)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (i64)gridDim.x * blockDim.x) {
        const int r = (int)(idx / wpr);
        const i64 j = idx - (i64)r * wpr;
        u32 w = 0;
        if (j < counts.words[r]) {
            u64 v = slots[idx];
            u64 t0 = 0;
            for (u32 spins = 0; (u32)(v >> 32) != epoch; ++spins) {
                if ((spins & 63u) == 63u) {
                    const u64 now = th_globaltimer_ns();
                    if (t0 == 0) t0 = now;
                    else if (now - t0 > timeout_ns) { *err = 1; v = 0; break; }
                }
                __nanosleep(64);
                v = slots[idx];
            }
            w = (u32)v;
        }
        out[idx] = w;
    }
}

// ---- conflict sets ------------------------------------------------------------------------------------
constexpr int CS_THREADS = 256;
constexpr int CS_WARPS = CS_THREADS / 32;

struct RowView {
    const int *seg; const ThP1 *p1; const ThP3 *p3; const ThP4 *p4; ThGrid g;
};

// does row j belong to the result list of (qs, qe, self) under `mode`?
__device__ __forceinline__ bool cs_match(int mode, i64 start, i64 end, i64 idc, i64 qs, i64 self) {
    if (mode == TH_MODE_STRICT) return end > qs && th_idc_id(idc) != self;
    if (mode == TH_MODE_INCLUSIVE) return end >= qs;
    return !th_idc_cancelled(idc) && (start >= qs || end > qs);   // TH_MODE_UPCOMING
}

// warp-uniform: can any row of [lo, c1) still match?  (looks at the running maximum stored at c1 - 1)
__device__ __forceinline__ bool cs_more(int mode, const RowView &v, int c1, i64 qs) {
    if (mode == TH_MODE_STRICT) return th_ld8(&v.p4[c1 - 1].mall) > qs;
    if (mode == TH_MODE_INCLUSIVE) return th_ld8(&v.p4[c1 - 1].mall) >= qs;
    const longlong2 e = th_ld16(&v.p1[c1 - 1]);   // {start, m1}: rows at or after qs always match
    return e.x >= qs || e.y > qs;
}

// G lanes per query (32: one warp per query, for few long lists; 4: eight queries per warp, for large batches of
// short lists -- the walk is a chain of dependent probes, so what matters is how many queries are in flight).
template <int G>
__device__ __forceinline__ u32 cs_gmask(int lane) { return G == 32 ? 0xffffffffu : (((1u << (G & 31)) - 1u) << (lane & ~(G - 1))); }

template <int G>
__global__ void __launch_bounds__(CS_THREADS) k2_sets_count_kernel(RowView v, i64 n_uuids, i64 m,
                                                                  const i64 *__restrict__ q_uuid,
                                                                  const i64 *__restrict__ q_start,
                                                                  const i64 *__restrict__ q_end,
                                                                  const i64 *__restrict__ q_self, int mode,
                                                                  int *__restrict__ out_cnt, int *__restrict__ out_beg,
                                                                  int *__restrict__ out_end) {
    const int lane = threadIdx.x & 31, gl = lane & (G - 1);
    const u32 gmask = cs_gmask<G>(lane);
    const i64 gstride = (i64)gridDim.x * (CS_THREADS / G);
    for (i64 q = (i64)blockIdx.x * (CS_THREADS / G) + (threadIdx.x / G); q < m; q += gstride) {
        const i64 u = q_uuid[q];
        int cnt = 0, beg = 0, pend = 0;
        if (u >= 0 && u < n_uuids) {
            const int lo = v.seg[u] + 1;   // position seg[u] is the sentinel
            const i64 qs = q_start[q], qe = q_end[q];
            const i64 self = q_self ? q_self[q] : -1;
            pend = (mode == TH_MODE_STRICT) ? th_grid_prefix_end<true>(v.g, v.p1, u, qe)
                                            : th_grid_prefix_end<false>(v.g, v.p1, u, qe);
            int c1 = pend;
            beg = pend;
            while (c1 > lo && cs_more(mode, v, c1, qs)) {
                const int c0 = c1 - G > lo ? c1 - G : lo;
                const int j = c0 + gl;
                bool mt = false;
                if (j < c1) {
                    const longlong2 e = th_ld16(&v.p3[j]);   // {end, idc}
                    const i64 s = (mode == TH_MODE_UPCOMING) ? th_ld8(&v.p1[j].start) : 0;
                    mt = cs_match(mode, s, e.x, e.y, qs, self);
                }
                if (G == 1) cnt += mt ? 1 : 0;                  // one lane per query: nothing to vote on
                else cnt += __popc(__ballot_sync(gmask, mt) & gmask);
                beg = c0;
                c1 = c0;
            }
        }
        if (gl == 0) { out_cnt[q] = cnt; out_beg[q] = beg; out_end[q] = pend; }
    }
}

template <int G>
__global__ void __launch_bounds__(CS_THREADS) k2_sets_fill_kernel(RowView v, i64 m, const i64 *__restrict__ q_start,
                                                                 const i64 *__restrict__ q_self, int mode,
                                                                 const int *__restrict__ beg_arr,
                                                                 const int *__restrict__ end_arr,
                                                                 const i64 *__restrict__ offsets,
                                                                 i64 *__restrict__ out_ids, u8 *__restrict__ out_canc,
                                                                 const i64 *__restrict__ totals = nullptr, i64 capacity = 0) {
    // launched before the host has seen the totals (one synchronisation per call instead of two): nothing is written
    // when the result does not fit the caller's buffers
    if (totals != nullptr && totals[0] > capacity) return;
    const int lane = threadIdx.x & 31, gl = lane & (G - 1);
    const u32 gmask = cs_gmask<G>(lane);
    const i64 gstride = (i64)gridDim.x * (CS_THREADS / G);
    for (i64 q = (i64)blockIdx.x * (CS_THREADS / G) + (threadIdx.x / G); q < m; q += gstride) {
        const i64 o0 = offsets[q];
        if (offsets[q + 1] == o0) continue;
        const int beg = beg_arr[q], pend = end_arr[q];
        const i64 qs = q_start[q];
        const i64 self = q_self ? q_self[q] : -1;
        i64 o = o0;
        for (int c0 = beg; c0 < pend; c0 += G) {
            const int j = c0 + gl;
            bool mt = false;
            longlong2 e = make_longlong2(0, 0);
            if (j < pend) {
                e = th_ld16(&v.p3[j]);
                const i64 s = (mode == TH_MODE_UPCOMING) ? th_ld8(&v.p1[j].start) : 0;
                mt = cs_match(mode, s, e.x, e.y, qs, self);
            }
            if (G == 1) {                                       // one lane per query: nothing to vote on
                if (mt) {
                    out_ids[o] = th_idc_id(e.y);
                    out_canc[o] = th_idc_cancelled(e.y) ? 1 : 0;
                    ++o;
                }
                continue;
            }
            const u32 bal = __ballot_sync(gmask, mt) & gmask;
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
template <int G>
__global__ void __launch_bounds__(CS_THREADS) k2_sets_order_kernel(i64 m, const i64 *__restrict__ offsets,
                                                                  const i64 *__restrict__ tmp_ids,
                                                                  const u8 *__restrict__ tmp_canc,
                                                                  i64 *__restrict__ out_ids, u8 *__restrict__ out_canc,
                                                                  const i64 *__restrict__ totals = nullptr, i64 capacity = 0) {
    // (speculative launch: skipped when the result does not fit or a list is long enough to want the radix sort)
    if (totals != nullptr && (totals[0] > capacity || totals[1] > 64)) return;
    const int gl = threadIdx.x & (G - 1);
    const i64 gstride = (i64)gridDim.x * (CS_THREADS / G);
    for (i64 q = (i64)blockIdx.x * (CS_THREADS / G) + (threadIdx.x / G); q < m; q += gstride) {
        const i64 o0 = offsets[q], o1 = offsets[q + 1];
        const i64 c = o1 - o0;
        for (i64 k = gl; k < c; k += G) {
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

__global__ void __launch_bounds__(XS_THREADS) xs_tile_sum_kernel(const int *__restrict__ cnt, i64 m, i64 *tile_sum,
                                                                 i64 *max_cnt) {
    __shared__ i64 wtot[XS_THREADS / 32];
    const i64 j0 = (i64)blockIdx.x * XS_TILE + (i64)threadIdx.x * XS_IPT;
    i64 s = 0;
    int mx = 0;
#pragma unroll
    for (int k = 0; k < XS_IPT; ++k)
        if (j0 + k < m) { const int c = cnt[j0 + k]; s += c; mx = c > mx ? c : mx; }
    i64 total;
    xs_block_exclusive(s, wtot, &total);
    if (threadIdx.x == 0) tile_sum[blockIdx.x] = total;
    mx = __reduce_max_sync(0xffffffffu, mx);
    if ((threadIdx.x & 31) == 0 && mx > 0) atomicMax(max_cnt, (i64)mx);     // longest single list (picks the id-ordering path)
}

// long lists: records {k0 = id, k1 = query, idx = slot} for K1, and the scatter back in sorted order
__global__ void __launch_bounds__(256) k2_sets_rec_kernel(i64 m, const i64 *__restrict__ offsets,
                                                         const i64 *__restrict__ tmp_ids, ThRec *__restrict__ rec) {
    const int lane = threadIdx.x & 31;
    const i64 wstride = (i64)gridDim.x * (blockDim.x / 32);
    for (i64 q = (i64)blockIdx.x * (blockDim.x / 32) + (threadIdx.x >> 5); q < m; q += wstride) {
        const i64 o0 = offsets[q], o1 = offsets[q + 1];
        for (i64 k = o0 + lane; k < o1; k += 32) {
            ThRec r;
            r.k0 = (u64)tmp_ids[k]; r.k1 = (u32)q; r.idx = (u32)k;
            rec[k] = r;
        }
    }
}
__global__ void __launch_bounds__(256) k2_sets_unrec_kernel(i64 total, const ThRec *__restrict__ rec,
                                                           const u8 *__restrict__ tmp_canc, i64 *__restrict__ out_ids,
                                                           u8 *__restrict__ out_canc) {
    for (i64 k = (i64)blockIdx.x * blockDim.x + threadIdx.x; k < total; k += (i64)gridDim.x * blockDim.x) {
        const ThRec r = rec[k];          // sorted by (query, id): position k of the CSR
        out_ids[k] = (i64)r.k0;
        out_canc[k] = tmp_canc[r.idx];
    }
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

// ctas_per_sm > 0 caps the persistent grid below what fits (used when the queries stream in over PCIe: a fraction of the
// resident warps already keeps the link full, and the rest of each SM stays free for other streams)
static int k2_launch(const th_snapshot *h, i64 m, const void *d_u, const void *d_s, const void *d_e, const void *d_f,
                     int mode, u32 *d_bits, const ThPeerOut &po, cudaStream_t st, int ctas_per_sm = 0, bool packed = false) {
    const i64 n_tiles = ((m + 31) / 32 + BM_WARPS - 1) / BM_WARPS;   // blocks needed to give every warp at least one tile
    // persistent grid: exactly as many CTAs as can be resident at once (148 SMs x occupancy)
    // (per device and mode; the query is a pure function of both, so racing first uses write the same value)
    static std::atomic<int> occ_cache[64][2];
    static std::atomic<int> sms_cache[64];
    const int oi = mode == TH_MODE_STRICT ? 0 : 1;
    const int di = h->device >= 0 && h->device < 64 ? h->device : 0;
    int occ = occ_cache[di][oi].load(std::memory_order_relaxed);
    if (occ == 0) {
        int v = 0;
        if (mode == TH_MODE_STRICT) TH_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&v, k2_bitmap_kernel<TH_MODE_STRICT>, BM_THREADS, 0));
        else TH_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&v, k2_bitmap_kernel<TH_MODE_INCLUSIVE>, BM_THREADS, 0));
        occ = v > 0 ? v : 1;
        occ_cache[di][oi].store(occ, std::memory_order_relaxed);
    }
    int n_sms = sms_cache[di].load(std::memory_order_relaxed);
    if (n_sms == 0) {
        n_sms = TH_NUM_SMS;
        cudaDeviceGetAttribute(&n_sms, cudaDevAttrMultiProcessorCount, h->device);
        sms_cache[di].store(n_sms, std::memory_order_relaxed);
    }
    const i64 resident = (i64)n_sms * (ctas_per_sm > 0 && ctas_per_sm < occ ? ctas_per_sm : occ);
    const unsigned blocks = (unsigned)(n_tiles < resident ? (n_tiles > 0 ? n_tiles : 1) : resident);
    int al16 = ((((uintptr_t)d_u | (uintptr_t)d_s | (uintptr_t)d_e | (uintptr_t)d_f) & 15) == 0) ? 1 : 0;
    if (getenv("TH_K2_DIRECT")) al16 = 0;
#define K2_GO(MODE_, PACKED_)                                                                                              \
    k2_bitmap_kernel<MODE_, PACKED_><<<blocks, BM_THREADS, 0, st>>>(th_grid_of(h), h->p1, h->p2, h->p4, h->n_uuids, m, d_u, \
                                                                    (const i64 *)d_s, (const i64 *)d_e, d_f, al16, d_bits, po)
    if (mode == TH_MODE_STRICT) { if (packed) K2_GO(TH_MODE_STRICT, true); else K2_GO(TH_MODE_STRICT, false); }
    else { if (packed) K2_GO(TH_MODE_INCLUSIVE, true); else K2_GO(TH_MODE_INCLUSIVE, false); }
#undef K2_GO
    TH_CUDA(cudaGetLastError());
    return TH_OK;
}

int th_exclusive_scan_i32(const int *d_cnt, i64 m, i64 *d_offsets, i64 *d_totals, ThScratch &sc, cudaStream_t st) {
    const i64 tiles = (m + XS_TILE - 1) / XS_TILE;
    i64 *d_tile;
    TH_TRY(sc.alloc((void **)&d_tile, 8 * (size_t)(tiles > 0 ? tiles : 1)));
    TH_CUDA(cudaMemsetAsync(d_totals, 0, 16, st));
    if (m == 0) { TH_CUDA(cudaMemsetAsync(d_offsets, 0, 8, st)); return TH_OK; }
    xs_tile_sum_kernel<<<(unsigned)tiles, XS_THREADS, 0, st>>>(d_cnt, m, d_tile, d_totals + 1);
    xs_carry_kernel<<<1, XS_THREADS, 0, st>>>(d_tile, tiles, d_totals);
    xs_write_kernel<<<(unsigned)tiles, XS_THREADS, 0, st>>>(d_cnt, m, d_tile, d_totals, d_offsets);
    TH_CUDA(cudaGetLastError());
    return TH_OK;
}

// Page-locked (cudaHostAlloc / cudaHostRegister), device-mapped host columns can be read by the kernel in place.
// Returns 1 and the device-side aliases when EVERY non-NULL pointer is such memory, 0 when none is, -1 on a mix.
static int k2_map_pinned(const void *const *cols, int n, const void **mapped) {
    int host = 0, other = 0;
    for (int c = 0; c < n; ++c) {
        mapped[c] = cols[c];
        if (cols[c] == nullptr) continue;
        cudaPointerAttributes a;
        if (cudaPointerGetAttributes(&a, cols[c]) == cudaSuccess && a.type == cudaMemoryTypeHost && a.devicePointer != nullptr) {
            mapped[c] = a.devicePointer;
            ++host;
        } else {
            cudaGetLastError();
            ++other;
        }
    }
    return host == 0 ? 0 : (other == 0 ? 1 : -1);
}

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
    cudaStream_t st = th_call_stream(stream, mem, h->device);
    ThScratch sc(st);
    const size_t zm = (size_t)m, words = (zm + 31) / 32;
    const void *d_u, *d_s, *d_e, *d_f = nullptr;
    void *d_bits;
    // Pinned (page-locked, device-mapped) host columns are read by the kernel itself: its cp.async ring streams them
    // over PCIe while earlier tiles are being answered, so the host -> device transfer and the search overlap and no
    // staging buffers or copy launches are needed.  Pageable memory is staged as before.
    bool zero_copy = false;
    if (mem == TH_MEM_HOST && !getenv("TH_NO_ZERO_COPY")) {
        const void *cols[4] = {q_uuid, q_start, q_end, q_self_id};
        const void *mapped[4];
        zero_copy = k2_map_pinned(cols, 4, mapped) == 1;
        if (zero_copy) { d_u = mapped[0]; d_s = mapped[1]; d_e = mapped[2]; d_f = mapped[3]; }
    }
    // Controller-sized batches in pageable memory: through the calling thread's page-locked slab (th_api.cu), which the
    // kernel reads and writes in place -- no allocation, no cudaMemcpy.
    if (!zero_copy && mem == TH_MEM_HOST && m <= TH_SLAB_MAX_QUERIES && !getenv("TH_NO_ZERO_COPY")) {
        void *dev = nullptr;
        char *slab = (char *)th_thread_slab(h->device, &dev);
        if (slab != nullptr) {
            const size_t col = (8 * zm + 15) & ~(size_t)15;
            memcpy(slab, q_uuid, 8 * zm);
            memcpy(slab + col, q_start, 8 * zm);
            memcpy(slab + 2 * col, q_end, 8 * zm);
            if (q_self_id) memcpy(slab + 3 * col, q_self_id, 8 * zm);
            ThPeerOut none;
            none.n = 0;
            none.unpack = 0;
            const char *d = (const char *)dev;
            TH_TRY(k2_launch(h, m, d, d + col, d + 2 * col, q_self_id ? d + 3 * col : nullptr, mode, (u32 *)(d + 4 * col), none, st));
            TH_CUDA(cudaStreamSynchronize(st));
            memcpy(out_bits_u32, slab + 4 * col, 4 * words);
            return TH_OK;
        }
    }
    if (!zero_copy) {
        TH_TRY(sc.in(q_uuid, 8 * zm, mem, &d_u));
        TH_TRY(sc.in(q_start, 8 * zm, mem, &d_s));
        TH_TRY(sc.in(q_end, 8 * zm, mem, &d_e));
        if (q_self_id) TH_TRY(sc.in(q_self_id, 8 * zm, mem, &d_f));
    }
    // the bitmap (4 bytes per 32 queries) goes straight into a pinned output buffer the same way
    bool out_direct = false;
    if (zero_copy) {
        cudaPointerAttributes a;
        if (cudaPointerGetAttributes(&a, out_bits_u32) == cudaSuccess && a.type == cudaMemoryTypeHost && a.devicePointer) {
            d_bits = a.devicePointer;
            out_direct = true;
        } else {
            cudaGetLastError();
        }
    }
    if (!out_direct) TH_TRY(sc.out(out_bits_u32, 4 * words, mem, &d_bits));
    ThPeerOut none;
    none.n = 0;
    none.unpack = 0;
    // one CTA per SM keeps the link full (1 / 2 / 4 per SM: 0.659 / 0.660 / 0.660 ms per 1 M-query batch, staged 0.78)
    TH_TRY(k2_launch(h, m, d_u, d_s, d_e, d_f, mode, (u32 *)d_bits, none, st, zero_copy ? 1 : 0));
    TH_CUDA(cudaGetLastError());
    if (!out_direct) TH_TRY(th_copy_back(out_bits_u32, d_bits, 4 * words, mem, st));
    if (mem == TH_MEM_HOST) TH_CUDA(cudaStreamSynchronize(st));
    return TH_OK;
}

// The same verdicts from a narrower wire format (the host path is bound by PCIe bytes, nothing else): uuid codes and
// self ids travel as uint32 columns -- q_self_id_u32[i] = 0xFFFFFFFF means "not saved yet", and the whole column may
// be NULL when no candidate is a self-update -- so a query is 24 or 20 bytes instead of 32.  Ids must be below 2^32 - 1.
extern "C" int th_conflict_bitmap_packed(const th_snapshot *h, int64_t m, const uint32_t *q_uuid_u32, const int64_t *q_start,
                                         const int64_t *q_end, const uint32_t *q_self_id_u32, int mode,
                                         uint32_t *out_bits_u32, int mem, void *stream) {
    TH_REQUIRE(h != nullptr && h->loaded, "th_conflict_bitmap_packed: snapshot not loaded");
    TH_REQUIRE(m >= 0, "th_conflict_bitmap_packed: negative m");
    TH_REQUIRE(mode == TH_MODE_STRICT || mode == TH_MODE_INCLUSIVE, "th_conflict_bitmap_packed: mode %d not supported", mode);
    TH_REQUIRE(mem == TH_MEM_HOST || mem == TH_MEM_DEVICE, "th_conflict_bitmap_packed: bad mem flag %d", mem);
    if (m == 0) return TH_OK;
    TH_REQUIRE(q_uuid_u32 && q_start && q_end && out_bits_u32, "th_conflict_bitmap_packed: NULL array");
    DeviceGuard guard(h->device);
    if (!guard.ok) { th_set_error("th_conflict_bitmap_packed: cannot select device %d", h->device); return TH_ECUDA; }
    cudaStream_t st = th_call_stream(stream, mem, h->device);
    ThScratch sc(st);
    const size_t zm = (size_t)m, words = (zm + 31) / 32;
    const void *d_u, *d_s, *d_e, *d_f = nullptr;
    void *d_bits;
    bool zero_copy = false;
    if (mem == TH_MEM_HOST && !getenv("TH_NO_ZERO_COPY")) {
        const void *cols[4] = {q_uuid_u32, q_start, q_end, q_self_id_u32};
        const void *mapped[4];
        zero_copy = k2_map_pinned(cols, 4, mapped) == 1;
        if (zero_copy) { d_u = mapped[0]; d_s = mapped[1]; d_e = mapped[2]; d_f = mapped[3]; }
    }
    if (!zero_copy) {
        TH_TRY(sc.in(q_uuid_u32, 4 * zm, mem, &d_u));
        TH_TRY(sc.in(q_start, 8 * zm, mem, &d_s));
        TH_TRY(sc.in(q_end, 8 * zm, mem, &d_e));
        if (q_self_id_u32) TH_TRY(sc.in(q_self_id_u32, 4 * zm, mem, &d_f));
    }
    bool out_direct = false;
    if (zero_copy) {
        cudaPointerAttributes a;
        if (cudaPointerGetAttributes(&a, out_bits_u32) == cudaSuccess && a.type == cudaMemoryTypeHost && a.devicePointer) {
            d_bits = a.devicePointer;
            out_direct = true;
        } else {
            cudaGetLastError();
        }
    }
    if (!out_direct) TH_TRY(sc.out(out_bits_u32, 4 * words, mem, &d_bits));
    ThPeerOut none;
    none.n = 0;
    none.unpack = 0;
    TH_TRY(k2_launch(h, m, d_u, d_s, d_e, d_f, mode, (u32 *)d_bits, none, st, zero_copy ? 1 : 0, true));
    if (!out_direct) TH_TRY(th_copy_back(out_bits_u32, d_bits, 4 * words, mem, st));
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
    cudaStream_t st = th_call_stream(stream, mem, h->device);
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
    v.seg = h->seg; v.p1 = h->p1; v.p3 = h->p3; v.p4 = h->p4; v.g = th_grid_of(h);
    // large batches (would_interfere for a whole candidate list): 4 lanes per query, eight queries per warp (1 M
    // config-2 candidates, 1.25 M rows returned: 0.54 / 0.37 / 0.26 / 0.185 / 0.163 ms with 32 / 16 / 8 / 4 / 2 lanes; 4 keeps long ranges affordable -- the walk is
    // a chain of dependent probes, so queries in flight are what counts); a few calendar ranges: one warp per query
    // ... and ONE lane per query for batches of 256 K candidates and more, where a list is a row or two (1 M candidates:
    // 0.170 ms with four lanes, 0.146 with two, 0.129 with one: no votes at all, 32 queries in flight per warp)
    int lanes = m >= (1 << 18) ? 1 : (m >= 32768 ? 4 : 32);
    if (const char *e = getenv("TH_SETS_LANES")) { const int v = atoi(e); if (v == 1 || v == 2 || v == 4 || v == 32) lanes = v; }   // diagnostics
    const int wgrid = th_grid_for(m, CS_THREADS / lanes, 16);
#define CS_BY_LANES(KERNEL, ...)                                                           \
    do {                                                                                   \
        if (lanes == 1) KERNEL<1><<<wgrid, CS_THREADS, 0, st>>>(__VA_ARGS__);              \
        else if (lanes == 2) KERNEL<2><<<wgrid, CS_THREADS, 0, st>>>(__VA_ARGS__);         \
        else if (lanes == 4) KERNEL<4><<<wgrid, CS_THREADS, 0, st>>>(__VA_ARGS__);         \
        else KERNEL<32><<<wgrid, CS_THREADS, 0, st>>>(__VA_ARGS__);                        \
    } while (0)
    CS_BY_LANES(k2_sets_count_kernel, v, h->n_uuids, m, (const i64 *)d_u, (const i64 *)d_s, (const i64 *)d_e, (const i64 *)d_f, mode, d_cnt, d_beg, d_end);
    TH_CUDA(cudaGetLastError());
    i64 *d_total;
    TH_TRY(sc.alloc((void **)&d_total, 16));       // [0] = total rows, [1] = longest list
    TH_TRY(th_exclusive_scan_i32(d_cnt, m, (i64 *)d_off, d_total, sc, st));
    // Device arrays: fill (and rank short lists) speculatively, guarded on the device by the totals, so that the call has
    // ONE host synchronisation in the common case; the radix path for long lists and the host-array path (whose copies
    // need the total) keep the second one.
    const bool spec = mem == TH_MEM_DEVICE && capacity > 0 && m < (1ll << 32) && capacity < (1ll << 30);
    i64 *sp_ids = nullptr;
    u8 *sp_canc = nullptr;
    if (spec) {
        if (mode == TH_MODE_UPCOMING) {
            CS_BY_LANES(k2_sets_fill_kernel, v, m, (const i64 *)d_s, (const i64 *)d_f, mode, d_beg, d_end, (const i64 *)d_off, (i64 *)out_ids, (u8 *)out_cancelled_u8, (const i64 *)d_total, (i64)capacity);
        } else {
            TH_TRY(sc.alloc((void **)&sp_ids, 8 * (size_t)capacity));
            TH_TRY(sc.alloc((void **)&sp_canc, (size_t)capacity));
            CS_BY_LANES(k2_sets_fill_kernel, v, m, (const i64 *)d_s, (const i64 *)d_f, mode, d_beg, d_end, (const i64 *)d_off, sp_ids, sp_canc, (const i64 *)d_total, (i64)capacity);
            CS_BY_LANES(k2_sets_order_kernel, m, (const i64 *)d_off, sp_ids, sp_canc, (i64 *)out_ids, (u8 *)out_cancelled_u8, (const i64 *)d_total, (i64)capacity);
        }
        TH_CUDA(cudaGetLastError());
    }
    i64 tot2[2] = {0, 0};
    TH_CUDA(cudaMemcpyAsync(tot2, d_total, 16, cudaMemcpyDeviceToHost, st));
    TH_TRY(th_copy_back(out_offsets, d_off, 8 * (zm + 1), mem, st));
    TH_CUDA(cudaStreamSynchronize(st));
    const i64 total = tot2[0], longest = tot2[1];
    *out_needed = total;
    if (total > capacity) {
        th_set_error("th_conflict_sets: %lld result rows exceed capacity %lld", (long long)total, (long long)capacity);
        return TH_ECAPACITY;
    }
    if (total == 0) return TH_OK;
    if (spec) {
        if (mode == TH_MODE_UPCOMING || longest <= 64) return TH_OK;          // everything was written before the synchronisation
        // long lists: the speculative fill has left (start, id)-ordered rows in sp_ids / sp_canc; sort them by (query, id)
        ThRec *ra, *rb, *sorted;
        TH_TRY(sc.alloc((void **)&ra, sizeof(ThRec) * (size_t)total));
        TH_TRY(sc.alloc((void **)&rb, sizeof(ThRec) * (size_t)total));
        k2_sets_rec_kernel<<<wgrid, 256, 0, st>>>(m, (const i64 *)d_off, sp_ids, ra);
        TH_CUDA(cudaGetLastError());
        int passes = 0;
        TH_TRY(th_radix_sort(ra, rb, total, th_bits_for((u64)h->max_id), th_bits_for((u64)(m - 1)), &sorted, &passes, st));
        k2_sets_unrec_kernel<<<th_grid_for(total, 256 * 4, 8), 256, 0, st>>>(total, sorted, sp_canc, (i64 *)out_ids, (u8 *)out_cancelled_u8);
        TH_CUDA(cudaGetLastError());
        TH_CUDA(cudaStreamSynchronize(st));
        return TH_OK;
    }
    void *d_ids, *d_canc;
    TH_TRY(sc.out(out_ids, 8 * (size_t)total, mem, &d_ids));
    TH_TRY(sc.out(out_cancelled_u8, (size_t)total, mem, &d_canc));
    if (mode == TH_MODE_UPCOMING) {
        CS_BY_LANES(k2_sets_fill_kernel, v, m, (const i64 *)d_s, (const i64 *)d_f, mode, d_beg, d_end, (const i64 *)d_off, (i64 *)d_ids, (u8 *)d_canc);
    } else {
        i64 *t_ids; u8 *t_canc;
        TH_TRY(sc.alloc((void **)&t_ids, 8 * (size_t)total));
        TH_TRY(sc.alloc((void **)&t_canc, (size_t)total));
        CS_BY_LANES(k2_sets_fill_kernel, v, m, (const i64 *)d_s, (const i64 *)d_f, mode, d_beg, d_end, (const i64 *)d_off, t_ids, t_canc);
        if (longest <= 64 || total >= (1ll << 30) || m >= (1ll << 32)) {
            // short lists: rank by counting inside the warp (c^2 / 32 compares per list)
            CS_BY_LANES(k2_sets_order_kernel, m, (const i64 *)d_off, t_ids, t_canc, (i64 *)d_ids, (u8 *)d_canc);
        } else {
            // calendar views return thousands of rows per uuid: sort all of them by (query, id) with K1 instead
            ThRec *ra, *rb, *sorted;
            TH_TRY(sc.alloc((void **)&ra, sizeof(ThRec) * (size_t)total));
            TH_TRY(sc.alloc((void **)&rb, sizeof(ThRec) * (size_t)total));
            k2_sets_rec_kernel<<<wgrid, 256, 0, st>>>(m, (const i64 *)d_off, t_ids, ra);
            TH_CUDA(cudaGetLastError());
            int passes = 0;
            TH_TRY(th_radix_sort(ra, rb, total, th_bits_for((u64)h->max_id), th_bits_for((u64)(m - 1)), &sorted, &passes, st));
            k2_sets_unrec_kernel<<<th_grid_for(total, 256 * 4, 8), 256, 0, st>>>(total, sorted, t_canc, (i64 *)d_ids,
                                                                                 (u8 *)d_canc);
        }
    }
    TH_CUDA(cudaGetLastError());
    TH_TRY(th_copy_back(out_ids, d_ids, 8 * (size_t)total, mem, st));
    TH_TRY(th_copy_back(out_cancelled_u8, d_canc, (size_t)total, mem, st));
    TH_CUDA(cudaStreamSynchronize(st));
    return TH_OK;
}


// Fused K2 + all-gather: every rank's kernel stores its bitmap words, tagged with the step's
// epoch, directly into the slot array of every rank (peer-mapped memory over NVLink/NVSwitch); an unpack kernel on the
// same stream polls the slots it expects and writes the compact [world][words_per_rank] words the caller reads.
// No NCCL call, no fence, no host round trip.  Device pointers only.  m_per_rank (host, world entries, may be NULL =
// "every rank has m") tells the receiver how many queries each rank contributes this step.
#undef CS_BY_LANES

static int k2_launch_unpack(th_peer *peer, cudaStream_t st) {
    const int parity = (int)(peer->epoch & 1u);
    const i64 total = (i64)peer->world * peer->words_per_rank;
    i64 used = 0;
    for (int r = 0; r < peer->world; ++r) used += peer->counts.words[r];
    int blocks = (int)((total + 255) / 256);
    if (blocks > TH_NUM_SMS * 4) blocks = TH_NUM_SMS * 4;
    (void)used;
    k2_ll_unpack_kernel<<<blocks, 256, 0, st>>>(th_peer_slots(peer->base, peer->world, peer->words_per_rank, parity),
                                                th_peer_out(peer, parity), peer->world, peer->words_per_rank,
                                                peer->counts, peer->epoch, th_peer_timeout_ns(),
                                                (int *)th_peer_misc(peer->base, peer->world, peer->words_per_rank));
    TH_CUDA(cudaGetLastError());
    return TH_OK;
}

static int k2_exchange_impl(const th_snapshot *h, th_peer *peer, int64_t m, const void *q_uuid, const int64_t *q_start,
                            const int64_t *q_end, const void *q_self_id, bool packed, int mode,
                            const int64_t *m_per_rank, int wait, uint32_t **out_gathered, void *stream) {
    TH_REQUIRE(h != nullptr && h->loaded, "th_conflict_bitmap_exchange: snapshot not loaded");
    TH_REQUIRE(peer != nullptr && peer->connected, "th_conflict_bitmap_exchange: peer group not connected");
    TH_REQUIRE(peer->device == h->device, "th_conflict_bitmap_exchange: peer group and snapshot are on different devices");
    TH_REQUIRE(m >= 0 && (m + 31) / 32 <= peer->words_per_rank,
               "th_conflict_bitmap_exchange: %lld queries need %lld words, the peer group was created for %lld per rank",
               (long long)m, (long long)((m + 31) / 32), (long long)peer->words_per_rank);
    TH_REQUIRE(mode == TH_MODE_STRICT || mode == TH_MODE_INCLUSIVE, "th_conflict_bitmap_exchange: mode %d not supported", mode);
    TH_REQUIRE(m == 0 || (q_uuid && q_start && q_end), "th_conflict_bitmap_exchange: NULL array");
    TH_REQUIRE(packed || mode != TH_MODE_STRICT || m == 0 || q_self_id, "th_conflict_bitmap_exchange: q_self_id is NULL");
    const int world = peer->world;
    const i64 wpr = peer->words_per_rank;
    for (int r = 0; r < world; ++r) {
        const i64 mr = m_per_rank ? m_per_rank[r] : m;
        TH_REQUIRE(mr >= 0 && (mr + 31) / 32 <= wpr, "th_conflict_bitmap_exchange: m_per_rank[%d] = %lld does not fit", r,
                   (long long)mr);
        peer->counts.words[r] = (mr + 31) / 32;
    }
    TH_REQUIRE(peer->counts.words[peer->rank] == (m + 31) / 32, "th_conflict_bitmap_exchange: m_per_rank[rank] != m");
    DeviceGuard guard(h->device);
    if (!guard.ok) { th_set_error("th_conflict_bitmap_exchange: cannot select device %d", h->device); return TH_ECUDA; }
    cudaStream_t st = (cudaStream_t)stream;
    // 0 means "empty slot"; on wrap-around skip to 2 so that consecutive steps keep alternating slot generations
    if (++peer->epoch == 0) peer->epoch = 2;
    const u32 epoch = peer->epoch;
    const int parity = (int)(epoch & 1u);
    ThPeerOut po;
    po.n = world; po.rank = peer->rank; po.epoch = epoch;
    po.word_off = (i64)peer->rank * wpr;
    po.recv_words = 0;
    for (int p = 0; p < world; ++p) {
        po.dst[p] = th_peer_slots(peer->peer_base[p], world, wpr, parity);
        po.flags[p] = nullptr;
        po.expect[p] = peer->counts.words[p];
        if (po.expect[p] > po.recv_words) po.recv_words = po.expect[p];
    }
    po.timeout_ns = th_peer_timeout_ns();
    po.trace = nullptr;
    if (const char *tr = getenv("TH_PEER_TRACE")) po.trace = (u64 *)strtoull(tr, nullptr, 0);      // diagnostics: a device buffer
    // wait = 1: the receive half runs inside the K2 kernel (needs every rank's kernel to be running at the same time,
    // i.e. one rank per GPU); wait = 0: only the send half is enqueued and th_peer_wait adds the unpack kernel later
    // (ranks emulated on one GPU, where a kernel must never wait for one that is queued behind it).
    po.unpack = (wait && m > 0) ? 1 : 0;
    po.wpr = wpr;
    po.my_slots = th_peer_slots(peer->base, world, wpr, parity);
    po.my_out = th_peer_out(peer, parity);
    po.err = (int *)th_peer_misc(peer->base, world, wpr);
    // query columns: device memory, or page-locked host memory that the kernel streams over PCIe itself
    const void *cols[4] = {q_uuid, q_start, q_end, q_self_id};
    const void *mapped[4];
    const int pinned = m > 0 ? k2_map_pinned(cols, 4, mapped) : 0;
    TH_REQUIRE(pinned >= 0, "th_conflict_bitmap_exchange: query columns mix page-locked host memory with other memory");
    if (m > 0) TH_TRY(k2_launch(h, m, mapped[0], mapped[1], mapped[2], mapped[3], mode, nullptr, po, st, pinned == 1 ? 1 : 0, packed));
    if (wait && m == 0) TH_TRY(k2_launch_unpack(peer, st));
    if (out_gathered) *out_gathered = th_peer_out(peer, parity);
    return TH_OK;
}

extern "C" int th_conflict_bitmap_exchange(const th_snapshot *h, th_peer *peer, int64_t m, const int64_t *q_uuid,
                                           const int64_t *q_start, const int64_t *q_end, const int64_t *q_self_id,
                                           int mode, const int64_t *m_per_rank, int wait, uint32_t **out_gathered,
                                           void *stream) {
    return k2_exchange_impl(h, peer, m, q_uuid, q_start, q_end, q_self_id, false, mode, m_per_rank, wait, out_gathered, stream);
}

// the same step from the packed wire format of th_conflict_bitmap_packed (uint32 uuid / self columns, self may be NULL)
extern "C" int th_conflict_bitmap_exchange_packed(const th_snapshot *h, th_peer *peer, int64_t m, const uint32_t *q_uuid_u32,
                                                  const int64_t *q_start, const int64_t *q_end,
                                                  const uint32_t *q_self_id_u32, int mode, const int64_t *m_per_rank,
                                                  int wait, uint32_t **out_gathered, void *stream) {
    return k2_exchange_impl(h, peer, m, q_uuid_u32, q_start, q_end, q_self_id_u32, true, mode, m_per_rank, wait, out_gathered,
                            stream);
}

// device-side rendezvous of all ranks: raise my barrier flag on every peer, then wait for everybody's
// Two rounds.  After one round the LAST rank to arrive leaves at once (it has seen everybody's flag) while the others
// leave one NVLink flight later, when its flag reaches them: the ranks start the step a flight apart, and the early one
// then waits that flight for its peers' words (measured: the fused step cost ~6 us more than the plain kernel whatever
// the batch).  In the second round everybody needs a flag that was sent a flight after the last arrival, so with three
// or more ranks all of them leave together.
__global__ void k2_peer_barrier_kernel(ThPeerOut po, volatile u32 *my_bflags, int *err) {
    const int r = threadIdx.x;
    for (int round = 0; round < 2; ++round) {
        if (r < po.n) {
            volatile u32 *f = po.flags[r] + round * TH_MAX_PEERS + po.rank;
            *f = po.epoch;
            __threadfence_system();
            volatile u32 *mine = my_bflags + round * TH_MAX_PEERS;
            for (u64 spins = 0; (int)(mine[r] - po.epoch) < 0; ++spins) {
                if (spins > (1ull << 24)) { *err = 1; break; }
                __nanosleep(100);
            }
        }
        __syncwarp();
    }
}

// All ranks meet on the device (bench.py uses it to line the ranks up before a timed step, so that a step's time
// is the exchange's latency and not the ranks' arrival skew).  Enqueued on `stream`; no host synchronisation.
extern "C" int th_peer_barrier(th_peer *peer, void *stream) {
    TH_REQUIRE(peer != nullptr && peer->connected, "th_peer_barrier: peer group not connected");
    int prev = 0;
    TH_CUDA(cudaGetDevice(&prev));
    if (prev != peer->device) TH_CUDA(cudaSetDevice(peer->device));
    ThPeerOut po;
    po.n = peer->world; po.rank = peer->rank; po.epoch = ++peer->bepoch; po.word_off = 0; po.unpack = 0;
    for (int p = 0; p < peer->world; ++p) {
        po.dst[p] = nullptr;
        po.flags[p] = th_peer_bflags(peer->peer_base[p], peer->world, peer->words_per_rank);
    }
    k2_peer_barrier_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(
        po, th_peer_bflags(peer->base, peer->world, peer->words_per_rank),
        (int *)th_peer_misc(peer->base, peer->world, peer->words_per_rank));
    cudaError_t e = cudaGetLastError();
    if (prev != peer->device) cudaSetDevice(prev);
    if (e != cudaSuccess) { th_set_error("th_peer_barrier: launch -> %s", cudaGetErrorString(e)); return TH_ECUDA; }
    return TH_OK;
}

// the receive half on its own (for th_conflict_bitmap_exchange(wait = 0)): enqueues the unpack kernel of the latest step
extern "C" int th_peer_wait(th_peer *peer, void *stream) {
    TH_REQUIRE(peer != nullptr && peer->connected, "th_peer_wait: peer group not connected");
    int prev = 0;
    TH_CUDA(cudaGetDevice(&prev));
    if (prev != peer->device) TH_CUDA(cudaSetDevice(peer->device));
    const int rc = k2_launch_unpack(peer, (cudaStream_t)stream);
    if (prev != peer->device) cudaSetDevice(prev);
    return rc;
}

extern "C" int th_peer_check(th_peer *peer, void *stream) {
    TH_REQUIRE(peer != nullptr && peer->base, "th_peer_check: peer group not created");
    int e = 0;
    int *flag = (int *)th_peer_misc(peer->base, peer->world, peer->words_per_rank);
    TH_CUDA(cudaMemcpyAsync(&e, flag, sizeof(int), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    TH_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    if (e) {
        TH_CUDA(cudaMemsetAsync(flag, 0, sizeof(int), (cudaStream_t)stream));      // reported once
        th_set_error("th_peer_check: a rank's words did not arrive within TH_PEER_TIMEOUT_MS; the words of that step "
                     "read as 0 and must not be used");
        return TH_EINTERNAL;
    }
    return TH_OK;
}
