// Shared declarations for libth_b200.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include "../../include/th_b200.h"

#ifndef TH_NUM_SMS
#define TH_NUM_SMS 148  // B200: 2 dies x 74 SMs; grids are sized in multiples of this
#endif

typedef long long i64;
typedef unsigned long long u64;
typedef unsigned int u32;
typedef unsigned char u8;

#define TH_I64_MIN (-0x7fffffffffffffffLL - 1)
#define TH_I64_MAX (0x7fffffffffffffffLL)

// ---- error plumbing -------------------------------------------------------------------------------
void th_set_error(const char *fmt, ...);

#define TH_CUDA(call)                                                                          \
    do {                                                                                       \
        cudaError_t e__ = (call);                                                              \
        if (e__ != cudaSuccess) {                                                              \
            th_set_error("%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__)); \
            return e__ == cudaErrorMemoryAllocation ? TH_ENOMEM : TH_ECUDA;                    \
        }                                                                                      \
    } while (0)

#define TH_TRY(call)              \
    do {                          \
        int r__ = (call);         \
        if (r__ != TH_OK) return r__; \
    } while (0)

#define TH_REQUIRE(cond, ...)         \
    do {                              \
        if (!(cond)) {                \
            th_set_error(__VA_ARGS__); \
            return TH_EINVAL;         \
        }                             \
    } while (0)

// ---- records moved by the radix sort (K1) ------------------------------------------------------------
// 16 bytes so that one record is one 128-bit load/store.
struct __align__(16) ThRec {
    u64 k0;   // low key word  (start_us - min_start, or id - min_id)
    u32 k1;   // high key word (uuid_id, or query index)
    u32 idx;  // original row
};

// ---- per-row index entries of a loaded snapshot -------------------------------------------------------
// All pairs are 16 B so that every probe is one 128-bit load.
struct __align__(16) ThP1 { i64 start; i64 m1;  };  // m1  = max end over non-cancelled rows of the segment prefix
struct __align__(16) ThP2 { i64 a1;    i64 m2;  };  // a1  = id of the row attaining m1, m2 = max over the others
struct __align__(16) ThP3 { i64 end;   i64 idc; };  // idc = id | cancelled << 63
struct __align__(16) ThP4 { i64 owner; i64 mall; }; // mall = max end over ALL rows of the segment prefix

// Row layout of a loaded snapshot: rows sorted by (uuid, start, id); every uuid's segment is PRECEDED by one
// sentinel row (start = end = m1 = m2 = mall = INT64_MIN, cancelled), so a search prefix is never empty and
// the hot kernel needs no per-uuid bounds: position seg[u] is the sentinel, real rows are
// (seg[u], seg[u + 1]).  nt = n + n_uuids positions in total.
//
// Time grid, ONE PER UUID: cell (u, b) covers starts in [t0_u + (b << shift_u), t0_u + ((b + 1) << shift_u)) of uuid u
// (the sentinel counts as bucket 0); uuid u owns cells [base_u, base_u + nb_u) with nb_u proportional to ITS row count
// and spread over ITS OWN span of starts, so a hot GPU with a hundred times the rows gets a hundred times the cells
// (round 1 cut every uuid's axis by one global span and the average row count: with skewed popularity most rows sat
// in buckets of hundreds, finished by bisection).  The per-uuid entry {t0, base, shift, nb} is 32 bytes, L2-resident
// (2 MB for 65 536 uuids).  A cell is 32 bytes = ONE probe (a single 256-bit load, LDG.E.256 -- new with sm_100): the
// first position of the bucket, how many rows it holds, the running maximum m1 of everything before it, and the
// bucket's FIRST row {start, m1}.  At ~1.5 cells per row half of the buckets are empty and a third hold one row, so
// most lower bounds are settled by the cell's own 32-byte sector without touching the row arrays at all (at
// config-5 size every scattered probe costs a 64-byte HBM atom: one atom per query instead of two or three).
struct __align__(32) ThCell { int pos; int cnt; i64 m1_before; i64 start0; i64 m1_0; };
struct ThCellV { i64 pc, m1_before, start0, m1_0; };   // as loaded: pc = pos | cnt << 32
struct __align__(32) ThUGrid { i64 t0; i64 base; int shift; int nb; i64 pad; };   // read as ThCellV: {t0, base, shift | nb << 32, -}
// A calendar whose uuids are about equally busy over about the same span keeps ONE plan for all uuids (nb > 0: cell
// (u, b) is cells[u * nb + b], no table read in the query path: 6-12 % faster than going through the table, measured);
// the load decides from the rows-per-bucket a random row would see under the shared plan (th_snapshot.cu).
struct ThGrid {
    const ThCell *cells;   // [sum of nb_u]
    const ThUGrid *ug;     // [n_uuids], always filled
    i64 t0;                // shared plan (nb > 0) ...
    int shift;
    int nb;                // ... or 0: per-uuid plans
};

struct th_snapshot {
    int device;
    int loaded;
    i64 n, n_uuids, nt;
    i64 device_bytes;
    i64 radix_passes;
    int *seg;      // [n_uuids + 1] position of each uuid's sentinel (int32: nt < 2^31)
    ThP1 *p1;      // [nt]
    ThP2 *p2;      // [nt]
    ThP3 *p3;      // [nt]
    ThP4 *p4;      // [nt]
    u32 *uuid;     // [nt] uuid of each position (export / rebuild)
    ThCell *grid;  // [n_cells]
    ThUGrid *ugrid;  // [n_uuids]
    i64 n_cells;
    i64 grid_t0;     // shared plan, grid_nb == 0 when the plans are per uuid
    int grid_shift, grid_nb;
    i64 max_id;                         // largest reservation id (bounds the digit plan when lists are id-sorted)
    i64 cap_nt, cap_uuids, cap_cells, cap_ugrid;   // allocated capacities: a reload that fits reuses the buffers
};
static inline ThGrid th_grid_of(const th_snapshot *h) {
    ThGrid g; g.cells = h->grid; g.ug = h->ugrid; g.t0 = h->grid_t0; g.shift = h->grid_shift; g.nb = h->grid_nb; return g;
}

// ---- peer exchange (multi-GPU): where a fused kernel deposits its bitmap words on every rank --------------
// "LL" protocol: every 32-bit bitmap word travels as ONE 8-byte store {word, epoch}.  A naturally aligned 8-byte
// store is single-copy atomic (also across NVLink), so the receiver knows a word has arrived as soon as its slot
// shows the current epoch -- no fences, no completion counters, no flag round trip after the data.
#define TH_MAX_PEERS 16
struct ThPeerOut {
    int n;                       // 0 = no exchange (single GPU)
    int rank;
    u32 epoch;                   // tag written into the high half of every slot of this step
    i64 word_off;                // rank * words_per_rank
    u64 *dst[TH_MAX_PEERS];      // slot array (this step's generation) of every rank, peer-mapped
    u32 *flags[TH_MAX_PEERS];    // th_peer_barrier only: barrier flag array of every rank
    // receive half inside the same kernel (unpack != 0): this rank's own slots / compact output / expectations
    int unpack;
    i64 wpr;
    const volatile u64 *my_slots;
    u32 *my_out;
    int *err;
    i64 expect[TH_MAX_PEERS];    // bitmap words each rank contributes this step
    u64 *trace;                  // diagnostics (TH_PEER_TRACE): 4 globaltimer stamps per CTA, or NULL
    i64 recv_words;              // max over expect[]: the word range the CTAs split between them for the receive half
    u64 timeout_ns;              // how long a receiver waits for a peer's word before it reports TH_EINTERNAL
};
struct ThPeerCounts { i64 words[TH_MAX_PEERS]; };   // bitmap words each rank contributes this step
struct th_peer {
    int device, world, rank;
    i64 words_per_rank;          // capacity per rank
    u32 epoch;
    u32 bepoch;                  // generation of th_peer_barrier
    void *base;                  // IPC-shared: slots [2 generations][world][words_per_rank] u64, then err + barrier flags
    u32 *out;                    // local: compact words [2 generations][world][words_per_rank]
    void *peer_base[TH_MAX_PEERS];
    ThPeerCounts counts;         // of the latest step
    int connected;
};
static inline size_t th_peer_bytes(int world, i64 wpr) { return (size_t)2 * world * wpr * 8 + (size_t)(2 * TH_MAX_PEERS + 8) * 4; }
static inline u64 *th_peer_slots(void *base, int world, i64 wpr, int parity) { return (u64 *)base + (size_t)parity * world * wpr; }
// after the slots: err | 7 spare | barrier flags round 1 [TH_MAX_PEERS] | round 2 [TH_MAX_PEERS]
static inline u32 *th_peer_misc(void *base, int world, i64 wpr) { return (u32 *)((u64 *)base + (size_t)2 * world * wpr); }
static inline u32 *th_peer_bflags(void *base, int world, i64 wpr) { return th_peer_misc(base, world, wpr) + 8; }
static inline u32 *th_peer_out(const th_peer *p, int parity) { return p->out + (size_t)parity * p->world * p->words_per_rank; }

// L2 sector-promotion hint for scattered probes.  Measured on B200 (tools/micro/probe_gran.cu, ncu dram__bytes_read): ONE
// scattered 16- or 32-byte read costs ~110 bytes of DRAM traffic with every plain flavour of ld.global (.nc, .cg, .cv,
// L1::no_allocate, cp.async; cudaLimitMaxL2FetchGranularity 32 / 64 / 128 makes no difference) -- the L2 fills the whole
// 128-byte line -- and 59 bytes with the .L2::64B qualifier (SASS LDG.E.LTC64B).  Every index probe of this library
// carries it: at config-5 size the interval search is bound by exactly those bytes.
#ifndef TH_L2H
#define TH_L2H ".L2::64B"
#endif

// ---- device helpers -------------------------------------------------------------------------------------
__device__ __forceinline__ longlong2 th_ld16(const void *p) {
    // read-only 128-bit load; the index entries are immutable while a snapshot is live
    longlong2 v;
    asm volatile("ld.global.nc" TH_L2H ".v2.s64 {%0, %1}, [%2];" : "=l"(v.x), "=l"(v.y) : "l"(p));
    return v;
}
// read-only 64-bit probe with the same hint
__device__ __forceinline__ i64 th_ld8(const i64 *p) {
    i64 v;
    asm volatile("ld.global.nc" TH_L2H ".s64 %0, [%1];" : "=l"(v) : "l"(p));
    return v;
}
__device__ __forceinline__ uint4 th_ld16u(const void *p) {
    uint4 v;
    asm volatile("ld.global.nc" TH_L2H ".v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return v;
}
// streaming 64-bit load (read once, do not keep in L1)
__device__ __forceinline__ i64 th_ld_stream(const i64 *p) {
    i64 v;
    asm volatile("ld.global.nc.L1::no_allocate.b64 %0, [%1];" : "=l"(v) : "l"(p));
    return v;
}

// 16-byte read-only load that executes only when `pred` holds (otherwise `dflt` is returned).  Written as
// predicated PTX so that several probes can be in flight together: given plain C++ the compiler turns
// "load A; load B; use A to decide whether B matters" into two DEPENDENT round trips.
__device__ __forceinline__ longlong2 th_ld16_if(const void *p, bool pred, longlong2 dflt) {
    longlong2 v = dflt;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.s32 p, %2, 0;\n"
        "@p ld.global.nc" TH_L2H ".v2.s64 {%0, %1}, [%3];\n"
        "}\n"
        : "+l"(v.x), "+l"(v.y)
        : "r"((int)pred), "l"(p));
    return v;
}

__device__ __forceinline__ u64 th_globaltimer_ns() {
    u64 t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
// bound of the exchange's receive half (TH_PEER_TIMEOUT_MS in the environment, default 2000)
u64 th_peer_timeout_ns();

__device__ __forceinline__ u32 th_smem_addr(const void *p) { return (u32)__cvta_generic_to_shared(p); }

// same, at a compile-time byte offset from `base`: neighbouring probes share ONE 64-bit address computation
template <int OFF>
__device__ __forceinline__ longlong2 th_ld16_if_off(const void *base, bool pred, longlong2 dflt) {
    longlong2 v = dflt;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.s32 p, %2, 0;\n"
        "@p ld.global.nc" TH_L2H ".v2.s64 {%0, %1}, [%3+%4];\n"
        "}\n"
        : "+l"(v.x), "+l"(v.y)
        : "r"((int)pred), "l"(base), "n"(OFF));
    return v;
}

// 32-byte read-only load (sm_100: LDG.E.256); `p` must be 32-byte aligned.  Executes only when `pred` holds.
__device__ __forceinline__ ThCellV th_ld32_if(const void *p, bool pred, ThCellV dflt) {
    ThCellV v = dflt;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.s32 p, %4, 0;\n"
        "@p ld.global.nc" TH_L2H ".v4.b64 {%0, %1, %2, %3}, [%5];\n"
        "}\n"
        : "+l"(v.pc), "+l"(v.m1_before), "+l"(v.start0), "+l"(v.m1_0)
        : "r"((int)pred), "l"(p));
    return v;
}
__device__ __forceinline__ ThCellV th_ld32(const void *p) {
    ThCellV v;
    asm volatile("ld.global.nc" TH_L2H ".v4.b64 {%0, %1, %2, %3}, [%4];"
                 : "=l"(v.pc), "=l"(v.m1_before), "=l"(v.start0), "=l"(v.m1_0)
                 : "l"(p));
    return v;
}

// ---- cp.async (LDGSTS): 16-byte global -> shared copies that bypass registers, grouped for multi-stage rings -----
__device__ __forceinline__ void th_cp_async16(void *smem_dst, const void *gmem_src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(th_smem_addr(smem_dst)), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void th_cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void th_cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__device__ __forceinline__ i64 th_idc_id(i64 idc) { return idc & TH_I64_MAX; }
__device__ __forceinline__ bool th_idc_cancelled(i64 idc) { return idc < 0; }

// the grid cell of (u, key); u must be a valid uuid.  One L2-resident 32-byte read of the uuid's entry, then arithmetic.
__device__ __forceinline__ const ThCell *th_cell_of(const ThGrid &g, i64 u, i64 key) {
    if (g.nb > 0) {                               // shared plan (warp-uniform branch)
        i64 b = 0;
        if (key > g.t0) {
            const u64 d = ((u64)key - (u64)g.t0) >> g.shift;
            b = d < (u64)(g.nb - 1) ? (i64)d : (i64)(g.nb - 1);
        }
        return g.cells + (u * g.nb + b);
    }
    const ThCellV e = th_ld32(&g.ug[u]);          // {t0, base, shift | nb << 32, -}
    const int shift = (int)(e.start0 & 0xffffffffll), nb = (int)(e.start0 >> 32);
    i64 b = 0;
    if (key > e.pc) {
        const u64 d = ((u64)key - (u64)e.pc) >> shift;
        b = d < (u64)(nb - 1) ? (i64)d : (i64)(nb - 1);
    }
    return g.cells + (e.m1_before + b);
}

// End of the prefix "rows of uuid u that start before `key`" (STRICT: start < key, else start <= key) as an absolute
// position, found through the time grid: one 16-byte cell, then a search over that bucket's rows only (rows of
// earlier buckets start before the bucket and hence before key, rows of later buckets after it).  The result is
// at least seg[u] + 1 because the sentinel (start = -inf) sits in bucket 0.  u must be a valid uuid.
template <bool STRICT>
__device__ __forceinline__ int th_grid_prefix_end(const ThGrid &g, const ThP1 *__restrict__ p1, i64 u, i64 key) {
    const ThCellV c = th_ld32(th_cell_of(g, u, key));
    int lo = (int)(c.pc & 0xffffffffll);
    const int cnt = (int)(c.pc >> 32);
    int hi = lo + cnt;
    {
        // the bucket's first row comes with the cell; rows 1..3 in one more round trip only when the bucket has them
        // and row 0 lies on the left of `key` (rows are start-sorted, so the ones on the left are a prefix); only
        // longer buckets go on to bisect the rest
        const bool in0 = cnt > 0 && (STRICT ? c.start0 < key : c.start0 <= key);
        if (!in0 || cnt == 1) return lo + (in0 ? 1 : 0);
        const longlong2 far = make_longlong2(TH_I64_MAX, 0);
        const ThP1 *r = p1 + lo;
        const longlong2 e1 = th_ld16_if_off<16>(r, cnt > 1, far);
        const longlong2 e2 = th_ld16_if_off<32>(r, cnt > 2, far), e3 = th_ld16_if_off<48>(r, cnt > 3, far);
        auto left = [&](const longlong2 &e, int j) { return cnt > j && (STRICT ? e.x < key : e.x <= key); };
        const int k = 1 + (int)left(e1, 1) + (int)left(e2, 2) + (int)left(e3, 3);
        if (k < 4 || cnt <= 4) return lo + k;
        lo += 4;
    }
    while (lo < hi) {
        const int mid = (int)(((unsigned)lo + (unsigned)hi) >> 1);
        const i64 s = th_ld8(&p1[mid].start);
        const bool left = STRICT ? (s < key) : (s <= key);
        if (left) lo = mid + 1; else hi = mid;
    }
    return lo;
}

__device__ __forceinline__ i64 th_warp_min(i64 v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        i64 t = __shfl_xor_sync(0xffffffffu, v, o);
        v = t < v ? t : v;
    }
    return v;
}
__device__ __forceinline__ i64 th_warp_sum(i64 v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// makes `dev` current for the lifetime of the object and restores the caller's device afterwards
struct DeviceGuard {
    int prev;
    bool ok;
    explicit DeviceGuard(int dev) : prev(0), ok(true) {
        if (cudaGetDevice(&prev) != cudaSuccess) ok = false;
        if (ok && prev != dev && cudaSetDevice(dev) != cudaSuccess) ok = false;
    }
    ~DeviceGuard() { cudaSetDevice(prev); }
};

// The stream a call runs on: the caller's, except that a TH_MEM_HOST call given NULL gets a private non-blocking stream of
// the CALLING THREAD on `device` (created on first use; `device` must be current).  Host-array calls return only after
// synchronising their stream, so nothing orders them against other streams -- and on the legacy default stream the
// service threads (3-4 of them plus greenlets) would serialise against each other and against
// every other stream of the process.
cudaStream_t th_call_stream(void *stream, int mem, int device);

// page-locked, device-mapped slab of the calling thread on `device` (current), TH_SLAB_BYTES long; NULL if unavailable
#define TH_SLAB_BYTES (64 << 10)
#define TH_SLAB_MAX_QUERIES 1024
void *th_thread_slab(int device, void **dev_alias);

// ---- host-side scratch: stream-ordered device buffers used to stage TH_MEM_HOST calls -----------------
struct ThScratch {
    cudaStream_t st;
    void *ptrs[24];
    int count;
    explicit ThScratch(cudaStream_t s) : st(s), count(0) {}
    ~ThScratch() {
        for (int i = 0; i < count; ++i) cudaFreeAsync(ptrs[i], st);
    }
    int alloc(void **out, size_t bytes);
    // device copy of a host array (or the pointer itself when mem == TH_MEM_DEVICE)
    int in(const void *src, size_t bytes, int mem, const void **out);
    // device buffer standing in for a host output (or the pointer itself)
    int out(void *dst, size_t bytes, int mem, void **out_dev);
};
int th_copy_back(void *host_dst, const void *dev_src, size_t bytes, int mem, cudaStream_t st);
// Stream-ordered scratch from the library's own memory pool on the CURRENT device.  The pool keeps what it has
// allocated (release threshold = unlimited): with the device's default pool every stream synchronisation hands the
// staging buffers back to the driver, and the next TH_MEM_HOST call pays for mapping them again.
cudaError_t th_malloc_async(void **p, size_t bytes, cudaStream_t st);

// exclusive scan of int32 counts into int64 offsets[m + 1] (th_search.cu); totals[0] = sum, totals[1] = max count.
// All pointers are device memory; `totals` must hold two i64.  Enqueued on `st`, no synchronisation.
int th_exclusive_scan_i32(const int *d_cnt, i64 m, i64 *d_offsets, i64 *d_totals, ThScratch &sc, cudaStream_t st);

// ---- K1 internal entry (th_sort.cu) ----------------------------------------------------------------------
// Sorts n records by (k1, k0) ascending, stable, using only the low k0_bits / k1_bits of each word.
// `a` holds the input; the sorted result is returned in *sorted (either a or b).  `passes` gets the
// number of digit passes that ran.
int th_radix_sort(ThRec *a, ThRec *b, i64 n, int k0_bits, int k1_bits, ThRec **sorted, int *passes,
                  cudaStream_t st);

static inline int th_bits_for(u64 maxval) {
    int b = 0;
    while (maxval) { ++b; maxval >>= 1; }
    return b;
}
static inline int th_grid_for(i64 work_items, int per_block, int max_waves = 32) {
    i64 blocks = (work_items + per_block - 1) / per_block;
    if (blocks < 1) blocks = 1;
    // round up to whole waves of the 148 SMs once the grid is larger than one wave
    i64 cap = (i64)TH_NUM_SMS * max_waves;
    if (blocks > cap) blocks = cap;
    if (blocks > TH_NUM_SMS) blocks = ((blocks + TH_NUM_SMS - 1) / TH_NUM_SMS) * TH_NUM_SMS;
    return (int)blocks;
}
