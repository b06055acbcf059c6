// K4 -- segmented reduction of usage samples into reservation windows (sm_100a).
//
// Replaces the reference's usage averager: per reservation a JSON file that is re-read and re-written on
// every tick (tensorhive/core/services/UsageLoggingService.py:85-116,159-171) and, on expiry,
// `avg(values) = sum(values) // len(values)` or -1 (:25-30, :218-219).  A sample belongs to a window iff
// it was logged while the reservation was in current_events(): start <= ts <= end, both inclusive
// (tensorhive/models/Reservation.py:96-97), so touching windows share their boundary sample.
//
// Eight lanes per window, four windows per warp (four lanes, eight windows for the uint8 format): the range search is the same instruction stream for every
// lane of a group, so sharing a warp between four windows cuts the issue slots per window by 4 (the first
// version, one warp per window, was instruction-issue bound at 70 % SM throughput, 36 % DRAM).  Samples of a
// uuid are ts-sorted, so the window is a contiguous index range [a, b): both ends are located by an
// interpolation guess followed by ONE coalesced 8-wide probe per end (exact for a regular logging cadence,
// which is what the service produces; the four probe loads of both ends are issued together), with a
// group-uniform galloping search from the window outwards only when a guess misses.  The range is then summed with 16-byte loads of
// gpu_util / mem_util.  The ts column is only probed, never streamed.
//
// The kernel is bound by the dependent chain per window (window columns -> uuid entry -> probes -> samples) times
// the resident warps, and by issue slots (about 230 warp instructions per window in the first version): the per-uuid
// entry table removes two levels of the chain and the int64 division, int32 positions relative to the window halve
// the bounds arithmetic.  Measured and dropped: prefetching the next windows' columns / entries one or two
// iterations ahead through registers (4.5-5.1 ms against 3.6 ms: the prefetches share scoreboards with the loads the
// warp is waiting for and cost 15-40 registers) or through a cp.async shared-memory ring (4.0 ms), four or eight pairs
// per lane and iteration (3.65 / 4.5 ms).  Round 2: the kernel is held at 40 registers (__launch_bounds__(256, 6), 48
// resident warps, no spills) -- with the converged-warp form below 48 warps win over 40 by 3-6 % on every variant.
#include "th_common.cuh"


namespace {

constexpr int US_THREADS = 256;
// lanes per window are a template parameter G (8 for int64 columns, 4 for uint8): US_THREADS / G windows per block
// and iteration

// Per-uuid entry of the sample log: where the segment starts, how long it is, its first timestamp and the
// interpolation slope (n - 1) / (ts[last] - ts[first]).  Built by a tiny kernel in front of K4 so that a window
// reaches its probes after ONE dependent 32-byte read (L2 resident: 2 MB for 65 536 uuids) instead of
// off[u] -> ts[lo], ts[hi - 1]; with int64 division gone from the hot loop as well.
struct __align__(16) UsSeg { i64 lo; i64 n; i64 t_lo; double inv; };

__device__ __forceinline__ UsSeg us_make_seg(const i64 *__restrict__ off, const i64 *__restrict__ ts, i64 u) {
    UsSeg e;
    e.lo = off[u];
    e.n = off[u + 1] - e.lo;
    e.t_lo = 0;
    e.inv = 0.0;
    if (e.n > 0) {
        e.t_lo = ts[e.lo];
        const double span = (double)ts[e.lo + e.n - 1] - (double)e.t_lo;
        if (e.n > 1 && span > 0.0) e.inv = (double)(e.n - 1) / span;
    }
    return e;
}

__global__ void __launch_bounds__(256) k4_segments_kernel(i64 n_uuids, const i64 *__restrict__ off,
                                                          const i64 *__restrict__ ts, UsSeg *__restrict__ tab) {
    const i64 u = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (u < n_uuids) tab[u] = us_make_seg(off, ts, u);
}

// 8-byte read-only load that executes only when `pred` holds (see th_ld16_if)
__device__ __forceinline__ i64 us_ld8_if(const i64 *p, bool pred, i64 dflt) {
    i64 v = dflt;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.s32 p, %1, 0;\n"
        "@p ld.global.nc" TH_L2H ".s64 %0, [%2];\n"
        "}\n"
        : "+l"(v)
        : "r"((int)pred), "l"(p));
    return v;
}

// start of the W-wide probe window for `key` inside a segment of n >= 1 samples (relative index)
template <int W>
__device__ __forceinline__ i64 us_window(i64 n, i64 key, i64 t_lo, double inv) {
    double x = ((double)key - (double)t_lo) * inv;        // only a guess: rounding cannot change the result
    x = fmin(fmax(x, 0.0), (double)(n - 1));
    i64 w0 = (i64)x - (W / 2 - 1);
    if (w0 > n - W) w0 = n - W;
    if (w0 < 0) w0 = 0;
    return w0;
}

// From the number of probe-window samples that fail the predicate (a prefix, the segment being ts-sorted) to the first
// index that satisfies it; a group-uniform search only when the whole window lies on one side of the answer.
template <bool GE, int W>
__device__ __forceinline__ i64 us_finish(const i64 *__restrict__ seg, i64 n, i64 key, i64 w0, int nfalse) {
    const i64 cntw = n - w0 < W ? n - w0 : W;
    // The guess missed.  Irregular logs (gaps, restarts) put the answer tens of samples away, rarely thousands: gallop
    // from the window in doubling steps until the predicate flips, then bisect that bracket -- about 2 log2(distance)
    // dependent loads instead of log2(n).
    auto holds = [&](i64 p) { const i64 x = th_ld8(&seg[p]); return GE ? (x >= key) : (x > key); };
    i64 l, r;   // the answer lies in [l, r]; the predicate holds at r (r == n counts as "holds")
    if (nfalse == 0) {
        if (w0 == 0) return 0;
        r = w0;
        for (i64 step = W;; step <<= 1) {
            const i64 p = r - step;
            if (p <= 0) { l = 0; break; }
            if (!holds(p)) { l = p + 1; break; }
            r = p;
        }
    } else if (nfalse == (int)cntw) {
        if (w0 + cntw == n) return n;
        l = w0 + cntw;
        for (i64 step = W;; step <<= 1) {
            const i64 p = l + step - 1;
            if (p >= n) { r = n; break; }
            if (holds(p)) { r = p; break; }
            l = p + 1;
        }
    } else {
        return w0 + nfalse;
    }
    while (l < r) {
        const i64 mid = l + ((r - l) >> 1);
        const i64 x = th_ld8(&seg[mid]);
        if (GE ? (x >= key) : (x > key)) r = mid; else l = mid + 1;
    }
    return l;
}

// [a, b) = samples of the segment with ws <= ts <= we, as indices relative to the segment.  Both ends are probed
// together: four independent loads per lane, then four ballots.  All arguments are uniform within the 8-lane group.
// The ballots are FULL-WARP (every lane of the warp calls this together; a group with nothing to search passes n = 0):
// a ballot over a partial mask that differs from group to group compiles into a loop over the groups -- 8 rounds of ~12
// instructions per ballot and shuffle, a fifth of the uint8 kernel's instructions.
template <int G, int W>
__device__ __forceinline__ void us_range(const i64 *__restrict__ seg, i64 n, i64 t_lo, double inv, i64 ws, i64 we, int gl,
                                         int gbase, i64 &a, i64 &b) {
    constexpr int R = W / G;                      // probe loads per lane and end
    const i64 wa = us_window<W>(n, ws, t_lo, inv), wb = us_window<W>(n, we, t_lo, inv);
    constexpr u32 GM = G == 32 ? 0xffffffffu : ((1u << (G & 31)) - 1u);
    i64 va[R], vb[R];
#pragma unroll
    for (int r = 0; r < R; ++r) va[r] = us_ld8_if(seg + wa + r * G + gl, wa + r * G + gl < n, TH_I64_MAX);
#pragma unroll
    for (int r = 0; r < R; ++r) vb[r] = us_ld8_if(seg + wb + r * G + gl, wb + r * G + gl < n, TH_I64_MAX);
    int fa = 0, fb = 0;                           // beyond the segment counts as "holds"
#pragma unroll
    for (int r = 0; r < R; ++r) fa += __popc((__ballot_sync(0xffffffffu, wa + r * G + gl < n && va[r] < ws) >> gbase) & GM);
#pragma unroll
    for (int r = 0; r < R; ++r) fb += __popc((__ballot_sync(0xffffffffu, wb + r * G + gl < n && vb[r] <= we) >> gbase) & GM);
    a = us_finish<true, W>(seg, n, ws, wa, fa);   // first ts >= ws
    b = us_finish<false, W>(seg, n, we, wb, fb);  // first ts >  we
    if (b < a) b = a;                             // end before start: empty window
}

__device__ __forceinline__ i64 us_floor_div(i64 a, i64 b) {   // b > 0; Python's //
    // utilisation sums of a window fit 32 bits unless it holds tens of millions of samples: a 32-bit division is ~20
    // instructions, the 64-bit one a subroutine of ~150, and every window pays two of them
    if ((((u64)a | (u64)b) >> 32) == 0) return (i64)((u32)a / (u32)b);
    i64 q = a / b;
    if ((a % b != 0) && (a < 0)) --q;
    return q;
}

__device__ __forceinline__ i64 us_sat_sub(i64 a, i64 b) {      // a - b for a >= b, clamped to INT64_MAX
    const u64 d = (u64)a - (u64)b;
    return d > (u64)TH_I64_MAX ? TH_I64_MAX : (i64)d;
}
__device__ __forceinline__ i64 us_sat_add(i64 a, i64 b) {      // a + b for a, b >= 0, clamped to INT64_MAX
    return a > TH_I64_MAX - b ? TH_I64_MAX : a + b;
}

// x / c for x >= 0, c > 0 through the double-precision reciprocal, corrected to the exact quotient: ~15 instructions
// against ~80 for the 64-bit integer division (two per window on an exact-cadence log: a fifth of that kernel's
// instructions).  (double)x is exact below 2^53 and the product is then off by at most a few units, which the remainder
// test repairs; larger x take the integer division.
__device__ __forceinline__ i64 us_div_nonneg(i64 x, i64 c, double rc) {
    if (x >= (1ll << 52)) return x / c;
    i64 q = (i64)((double)x * rc);
    i64 r = x - q * c;
    while (r < 0) { --q; r += c; }
    while (r >= c) { ++q; r -= c; }
    return q;
}

// ---- range sums --------------------------------------------------------------------------------------------------
// int64 columns, 16-byte aligned: element pairs, two pairs per lane and iteration, every load issued before any is
// consumed.  Positions are int32 offsets from the (even) first pair, so the bounds checks are single compares.
template <int G>
__device__ __forceinline__ void us_sum_i64(const i64 *__restrict__ gpu, const i64 *__restrict__ mem, i64 s_total, i64 a,
                                           i64 b, int gl, i64 &sg, i64 &sm) {
    const i64 base = a & ~1ll;
    const i64 span = b - base;
    if (span <= 0) return;
    if (span >= (1ll << 30)) {                       // a window of a billion samples: plain strided loop
        for (i64 i = a + gl; i < b; i += G) { sg += __ldg(&gpu[i]); sm += __ldg(&mem[i]); }
        return;
    }
    const int ra = (int)(a - base), rb = (int)span;
    const i64 room = s_total - 1 - base;             // pair j may be loaded as 16 bytes iff 2 j < room
    const int rl = room < (i64)rb ? (int)room : rb;
    const longlong2 *gp = reinterpret_cast<const longlong2 *>(gpu + base);
    const longlong2 *mp = reinterpret_cast<const longlong2 *>(mem + base);
    const longlong2 zero2 = make_longlong2(0, 0);
    for (int j = gl; 2 * j < rb; j += 2 * G) {
        const int k = j + G;
        longlong2 g0 = th_ld16_if(gp + j, 2 * j < rl, zero2);
        longlong2 m0 = th_ld16_if(mp + j, 2 * j < rl, zero2);
        longlong2 g1 = th_ld16_if(gp + k, 2 * k < rl, zero2);
        longlong2 m1 = th_ld16_if(mp + k, 2 * k < rl, zero2);
        if (2 * j < ra) { g0.x = 0; m0.x = 0; }              // first pair of a window that starts on an odd index
        if (2 * j + 1 >= rb) { g0.y = 0; m0.y = 0; }         // last pair of a window that ends on an odd index
        if (2 * k + 1 >= rb) { g1.y = 0; m1.y = 0; }
        sg += (g0.x + g0.y) + (g1.x + g1.y);
        sm += (m0.x + m0.y) + (m1.x + m1.y);
    }
    // the last element of an odd-length column has no 16-byte partner
    if (rl < rb && (room & 1) == 0 && gl == 0 && s_total - 1 >= a) { sg += __ldg(&gpu[s_total - 1]); sm += __ldg(&mem[s_total - 1]); }
}

// uint8 columns (utilisation is a percentage, one byte is lossless), 16-byte aligned:
// 16 samples per lane and load, bytes outside [a, b) masked, four bytes added per __dp4a.
// 0xff for bytes k of an 8-byte half that lie in [lo, hi) (0 <= lo, hi <= 8 after clamping): two 64-bit shifts instead of a
// compare-and-select per byte (the per-byte form was 36 % of the uint8 kernel's instructions: ISETP + SEL)
__device__ __forceinline__ u64 us_mask8(int lo, int hi) {
    lo = lo < 0 ? 0 : (lo > 8 ? 8 : lo);
    hi = hi < 0 ? 0 : (hi > 8 ? 8 : hi);
    if (hi <= lo) return 0ull;
    return (~0ull >> (8 * (8 - hi))) & (~0ull << (8 * lo));       // hi > lo: both shift counts are below 64
}
__device__ __forceinline__ int us_sum16(const uint4 v) {
    u32 acc = 0;   // unsigned bytes x 1: at most 16 * 255
    acc = __dp4a(v.x, 0x01010101u, acc);
    acc = __dp4a(v.y, 0x01010101u, acc);
    acc = __dp4a(v.z, 0x01010101u, acc);
    acc = __dp4a(v.w, 0x01010101u, acc);
    return (int)acc;
}
__device__ __forceinline__ uint4 us_mask16(uint4 v, int lo, int hi) {
    const u64 m0 = us_mask8(lo, hi), m1 = us_mask8(lo - 8, hi - 8);
    v.x &= (u32)m0; v.y &= (u32)(m0 >> 32); v.z &= (u32)m1; v.w &= (u32)(m1 >> 32);
    return v;
}
// Only the first and the last chunk of a range can be partial.  They go to lanes 0 and 1 of the group, once, with masks;
// the chunks between them are whole and are summed by all lanes without a mask.  (Round 2, from the ncu source page: with
// the masks inside the chunk loop 99 % of the WARP's iterations took the mask branch, because some lane of the warp's 32
// always held an edge chunk -- 75 of the loop's 123 instructions, three times per warp iteration.)
template <int G>
__device__ __forceinline__ void us_sum_u8(const u8 *__restrict__ gpu, const u8 *__restrict__ mem, i64 s_total, i64 a,
                                          i64 b, int gl, i64 &sg, i64 &sm) {
    static_assert(G >= 2, "the two edge chunks go to two lanes");
    if (b <= a) return;
    const i64 c0 = a >> 4, cl = (b - 1) >> 4;           // first and last 16-sample chunk
    // the edge chunk of this lane, if any: loaded before the interior loop, consumed after it
    const bool edge = gl == 0 || (gl == 1 && cl != c0);
    const i64 e0 = (gl == 0 ? c0 : cl) << 4;
    const int lo = gl == 0 ? (int)(a - e0) : 0;
    const int hi = b - e0 < 16 ? (int)(b - e0) : 16;
    const bool wide = edge && e0 + 16 <= s_total;       // the whole 16 bytes exist
    const longlong2 zero2 = make_longlong2(0, 0);
    const longlong2 eg = th_ld16_if(reinterpret_cast<const longlong2 *>(gpu + e0), wide, zero2);
    const longlong2 em = th_ld16_if(reinterpret_cast<const longlong2 *>(mem + e0), wide, zero2);
#pragma unroll 1
    for (i64 c = c0 + 1 + gl; c < cl; c += G) {         // interior chunks end before b <= s_total
        const uint4 g4 = th_ld16u(gpu + (c << 4));
        const uint4 m4 = th_ld16u(mem + (c << 4));
        sg += us_sum16(g4);
        sm += us_sum16(m4);
    }
    if (wide) {
        uint4 g4 = make_uint4((u32)eg.x, (u32)((u64)eg.x >> 32), (u32)eg.y, (u32)((u64)eg.y >> 32));
        uint4 m4 = make_uint4((u32)em.x, (u32)((u64)em.x >> 32), (u32)em.y, (u32)((u64)em.y >> 32));
        g4 = us_mask16(g4, lo, hi);
        m4 = us_mask16(m4, lo, hi);
        sg += us_sum16(g4);
        sm += us_sum16(m4);
    } else if (edge) {                                  // the column ends inside this chunk
        for (int t = lo; t < hi; ++t) { sg += gpu[e0 + t]; sm += mem[e0 + t]; }
    }
}

// ELEM = bytes per utilisation sample (8: int64 columns, 1: uint8 columns); VEC = the columns are 16-byte aligned;
// TAB = `tab` holds the per-uuid entries (otherwise each window derives its entry from off / ts itself -- used when
// there are far more uuids than windows).
// REG = entries with a NEGATIVE slope field describe an exact arithmetic run, ts[k] = t_lo + k * cadence with cadence = -inv
// (the sample ring establishes that while it appends, th_ring.cu): both range ends follow from two divisions and the ts
// column is not touched at all.
template <int ELEM, bool VEC, bool TAB, int G, int W, bool REG = false>
__global__ void __launch_bounds__(US_THREADS, 6) k4_usage_kernel(i64 s_total, i64 n_uuids, const i64 *__restrict__ off,
                                                              const UsSeg *__restrict__ tab, const i64 *__restrict__ ts,
                                                              const void *__restrict__ gpu_v, const void *__restrict__ mem_v,
                                                              i64 w, const i64 *__restrict__ win_uuid,
                                                              const i64 *__restrict__ win_start,
                                                              const i64 *__restrict__ win_end, i64 *__restrict__ o_sg,
                                                              i64 *__restrict__ o_sm, i64 *__restrict__ o_cnt,
                                                              i64 *__restrict__ o_ag, i64 *__restrict__ o_am) {
    const int lane = threadIdx.x & 31;
    constexpr int GROUPS_PER_BLOCK = US_THREADS / G;
    const int gl = lane & (G - 1);
    const int gbase = lane & ~(G - 1);
    constexpr int GPW = 32 / G;                               // windows per warp and iteration
    const i64 gstride = (i64)gridDim.x * GROUPS_PER_BLOCK;
    // The warp stays converged: a warp-uniform loop (a warp's last iteration may have idle groups) and full-warp ballots
    // and shuffles.  A/B on one B200 against per-group masks and a per-group loop, both at 40 registers (6 blocks per
    // SM; at 48 registers every variant loses 3-6 %): uint8 columns 1.27 -> 1.08 ms, int64 columns 2.77 -> 2.74 ms,
    // shuffled windows 3.70 -> 3.65 ms, ring with jittered stamps 0.168 -> 0.146 ms, ring on an exact cadence 0.141 -> 0.143 ms.
    for (i64 qb = (i64)blockIdx.x * GROUPS_PER_BLOCK + (i64)(threadIdx.x >> 5) * GPW; qb < w; qb += gstride) {
        const i64 q = qb + (lane / G);
        const bool live = q < w;
        const i64 u = live ? win_uuid[q] : -1;
        const i64 ws = live ? win_start[q] : 0, we = live ? win_end[q] : 0;
        i64 a = 0, b = 0;
        UsSeg e;
        e.lo = 0; e.n = 0; e.t_lo = 0; e.inv = 0.0;
        if (u >= 0 && u < n_uuids) {
            if (TAB) {
                const longlong2 x = th_ld16(&tab[u]);
                const longlong2 y = th_ld16(reinterpret_cast<const char *>(&tab[u]) + 16);
                e.lo = x.x; e.n = x.y; e.t_lo = y.x; e.inv = __longlong_as_double(y.y);
            } else {
                e = us_make_seg(off, ts, u);
            }
        }
        const bool reg = REG && e.inv < 0.0;                  // an exact arithmetic run: no search at all
        // a warp-uniform branch: a warp whose windows are all empty or arithmetic runs skips the search
        const bool need = !reg && e.n > 0;
        if (__any_sync(0xffffffffu, need))
            us_range<G, W>(ts + e.lo, need ? e.n : 0, e.t_lo, need ? e.inv : 0.0, ws, we, gl, gbase, a, b);
        if (reg) {
            const i64 c = (i64)(-e.inv);
            const double rc = 1.0 / (double)c;
            // (saturating: a window end at the int64 extremes must not wrap; anything beyond the run is clamped below)
            a = ws <= e.t_lo ? 0 : us_div_nonneg(us_sat_add(us_sat_sub(ws, e.t_lo), c - 1), c, rc);   // first k with t_lo + k c >= ws
            b = we < e.t_lo ? 0 : us_sat_add(us_div_nonneg(us_sat_sub(we, e.t_lo), c, rc), 1);         // one past the last k with t_lo + k c <= we
            a = a < e.n ? a : e.n;
            b = b < e.n ? b : e.n;
            if (b < a) b = a;
        }
        a += e.lo;
        b += e.lo;
        i64 sg = 0, sm = 0;
        if (ELEM == 8) {
            const i64 *gpu = static_cast<const i64 *>(gpu_v), *mem = static_cast<const i64 *>(mem_v);
            if (VEC) us_sum_i64<G>(gpu, mem, s_total, a, b, gl, sg, sm);
            else for (i64 i = a + gl; i < b; i += G) { sg += __ldg(&gpu[i]); sm += __ldg(&mem[i]); }
        } else {
            const u8 *gpu = static_cast<const u8 *>(gpu_v), *mem = static_cast<const u8 *>(mem_v);
            if (VEC) us_sum_u8<G>(gpu, mem, s_total, a, b, gl, sg, sm);
            else for (i64 i = a + gl; i < b; i += G) { sg += gpu[i]; sm += mem[i]; }
        }
#pragma unroll
        for (int o = G / 2; o > 0; o >>= 1) {
            sg += __shfl_xor_sync(0xffffffffu, sg, o);        // o < G: partners stay inside the group
            sm += __shfl_xor_sync(0xffffffffu, sm, o);
        }
        if (gl == 0 && live) {
            const i64 c = b - a;
            if (o_sg) o_sg[q] = sg;
            if (o_sm) o_sm[q] = sm;
            if (o_cnt) o_cnt[q] = c;
            if (o_ag) o_ag[q] = c > 0 ? us_floor_div(sg, c) : -1;
            if (o_am) o_am[q] = c > 0 ? us_floor_div(sm, c) : -1;
        }
    }
}

template <int ELEM, bool VEC, int G, int W>
void us_launch(bool tab_mode, int grid, cudaStream_t st, i64 s, i64 n_uuids, const i64 *off, const UsSeg *tab, const i64 *ts,
               const void *g, const void *m, i64 w, const i64 *wu, const i64 *ws, const i64 *we, i64 *sg, i64 *sm, i64 *c,
               i64 *ag, i64 *am) {
    if (tab_mode)
        k4_usage_kernel<ELEM, VEC, true, G, W><<<grid, US_THREADS, 0, st>>>(s, n_uuids, off, tab, ts, g, m, w, wu, ws, we, sg, sm, c, ag, am);
    else
        k4_usage_kernel<ELEM, VEC, false, G, W><<<grid, US_THREADS, 0, st>>>(s, n_uuids, off, tab, ts, g, m, w, wu, ws, we, sg, sm, c, ag, am);
}

}  // namespace

// Reduction over a log whose per-uuid entries the caller has already built on the device (th_ring.cu): `tab` holds
// n_uuids UsSeg-layout entries {lo, n, t_lo, inv or -cadence}; the sample columns are uint8, 16-byte aligned, with
// `cap` addressable elements.  Window arrays and outputs follow `mem` as everywhere else.
int th_usage_reduce_table_u8(int device, const void *tab, int64_t n_uuids, int64_t cap, const int64_t *ts_us,
                             const uint8_t *gpu_u8, const uint8_t *mem_u8, int64_t w, const int64_t *win_uuid,
                             const int64_t *win_start, const int64_t *win_end, int64_t *out_sum_gpu,
                             int64_t *out_sum_mem, int64_t *out_count, int64_t *out_avg_gpu, int64_t *out_avg_mem,
                             int mem, cudaStream_t st) {
    if (w == 0) return TH_OK;
    TH_REQUIRE(win_uuid && win_start && win_end, "th_ring_reduce: NULL window array");
    ThScratch sc(st);
    const size_t zw = (size_t)w;
    const void *d_wu, *d_ws, *d_we;
    void *d_sg, *d_sm, *d_c, *d_ag, *d_am;
    TH_TRY(sc.in(win_uuid, 8 * zw, mem, &d_wu));
    TH_TRY(sc.in(win_start, 8 * zw, mem, &d_ws));
    TH_TRY(sc.in(win_end, 8 * zw, mem, &d_we));
    TH_TRY(sc.out(out_sum_gpu, 8 * zw, mem, &d_sg));
    TH_TRY(sc.out(out_sum_mem, 8 * zw, mem, &d_sm));
    TH_TRY(sc.out(out_count, 8 * zw, mem, &d_c));
    TH_TRY(sc.out(out_avg_gpu, 8 * zw, mem, &d_ag));
    TH_TRY(sc.out(out_avg_mem, 8 * zw, mem, &d_am));
    const int grid = th_grid_for(w, US_THREADS / 4, 64);
    k4_usage_kernel<1, true, true, 4, 8, true><<<grid, US_THREADS, 0, st>>>(
        (i64)cap, (i64)n_uuids, (const i64 *)nullptr, (const UsSeg *)tab, (const i64 *)ts_us, (const void *)gpu_u8,
        (const void *)mem_u8, (i64)w, (const i64 *)d_wu, (const i64 *)d_ws,
        (const i64 *)d_we, (i64 *)d_sg, (i64 *)d_sm, (i64 *)d_c, (i64 *)d_ag, (i64 *)d_am);
    TH_CUDA(cudaGetLastError());
    TH_TRY(th_copy_back(out_sum_gpu, d_sg, 8 * zw, mem, st));
    TH_TRY(th_copy_back(out_sum_mem, d_sm, 8 * zw, mem, st));
    TH_TRY(th_copy_back(out_count, d_c, 8 * zw, mem, st));
    TH_TRY(th_copy_back(out_avg_gpu, d_ag, 8 * zw, mem, st));
    TH_TRY(th_copy_back(out_avg_mem, d_am, 8 * zw, mem, st));
    if (mem == TH_MEM_HOST) TH_CUDA(cudaStreamSynchronize(st));
    (void)device;
    return TH_OK;
}

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
    cudaStream_t st = th_call_stream(stream, mem, device);
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
            // lanes per window, measured on config 3: 8 / 4 / 2 lanes = 2.98 / 3.34 / 4.9 ms for int64 columns (a
            // 100-sample window is 50 16-byte pairs) and 2.00 / 1.82 / 2.9 ms for uint8 columns (7 16-byte chunks)
            const int lanes = elem == 1 ? 4 : 8;
            const int grid = th_grid_for(w, US_THREADS / lanes, 64);
            // per-uuid entries: worth a pass over the uuids unless there are far more of them than windows
            const bool tab_mode = n_uuids > 0 && n_uuids <= 4 * w + 1024;
            void *d_tab = nullptr;
            if (tab_mode) {
                US_TRY(sc.alloc(&d_tab, sizeof(UsSeg) * zu));
                k4_segments_kernel<<<(unsigned)((n_uuids + 255) / 256), 256, 0, st>>>(n_uuids, (const i64 *)d_off,
                                                                                     (const i64 *)d_ts, (UsSeg *)d_tab);
            }
#define US_ARGS tab_mode, grid, st, s, n_uuids, (const i64 *)d_off, (const UsSeg *)d_tab, (const i64 *)d_ts, d_g, d_m, w, \
                (const i64 *)d_wu, (const i64 *)d_ws, (const i64 *)d_we, (i64 *)d_sg, (i64 *)d_sm, (i64 *)d_c, (i64 *)d_ag, (i64 *)d_am
            // probe width 8 (16 / 8 / 4 measured on config 3: 2.96 / 2.80 / -- ms for int64 columns, 1.69 / 1.35 / 1.23 ms
            // for uint8; on an irregular log -- a fifth of the samples missing, 0.9 s jitter -- 3.39 / 3.41 / -- and
            // 2.15 / 1.91 / 1.90 ms: 8 keeps a margin of +-3 samples around the guess at no cost on either kind of log)
            if (elem == 1 && vec) us_launch<1, true, 4, 8>(US_ARGS);
            else if (elem == 1) us_launch<1, false, 4, 8>(US_ARGS);
            else if (vec) us_launch<8, true, 8, 8>(US_ARGS);
            else us_launch<8, false, 8, 8>(US_ARGS);
#undef US_ARGS
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
