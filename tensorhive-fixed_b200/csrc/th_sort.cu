// K1 -- stable LSD radix sort of 16-byte records, onesweep style (sm_100a).
//
// Replaces: SQLite's full scan + `ORDER BY _start` (tensorhive/models/Reservation.py:117) and, more
// generally, the absence of any index on `reservations` in the reference.
//
// Shape of the algorithm
//   1. one histogram kernel counts every digit of every pass up front (one read of the keys);
//   2. a 256-thread scan turns each pass's histogram into global digit bases;
//   3. per pass ONE kernel reads each record once and writes it once ("onesweep"): a tile of 2048
//      records is ranked with warp-level multi-split (__match_any_sync peers, per-warp digit counters in
//      shared memory), the tile's digit counts are published and the exclusive prefix over earlier
//      tiles is obtained by decoupled look-back (flag+count packed in one 32-bit word), records are
//      regrouped by digit in shared memory and leave in coalesced runs of 16-byte stores.
//   Tiles take their index from an atomic ticket so a tile only ever waits on tiles that are already
//   resident; a watchdog turns a would-be hang into TH_EINTERNAL.
//
// HBM traffic: 16 B read + 16 B written per record per pass, plus 16 B per record for the histogram.
// The digit plan covers only the key bits that can differ (callers pass bit widths), so a
// (4096-uuid, 47-bit start) key runs 6 + 2 passes, not 12.
#include "th_common.cuh"
#include <atomic>

namespace {

constexpr int RS_THREADS = 256;
constexpr int RS_WARPS = RS_THREADS / 32;
// 8 records per thread: 64 registers, no spills, 4 CTAs per SM (16 per thread needed 128 registers or, capped at 85 for
// 3 CTAs per SM, spilled the records to local memory: 4.07 against 3.91 ms for the 12.5 M-row reload)
constexpr int RS_IPT = 8;
constexpr int RS_TILE = RS_THREADS * RS_IPT;     // 2048 records = 32 KB of shared memory
constexpr int RS_RADIX = 256;
constexpr int RS_MAXPASS = 12;                   // 64 + 32 key bits / 8
constexpr u32 RS_VAL_MASK = (1u << 30) - 1;
constexpr u32 RS_FLAG_AGG = 1u << 30;            // tile count published, prefix not known yet
constexpr u32 RS_FLAG_INC = 2u << 30;            // inclusive prefix published
constexpr u32 RS_SPIN_LIMIT = 1u << 22;

struct RsPlan {
    int npass;
    u8 word[RS_MAXPASS];   // 0 = k0, 1 = k1
    u8 shift[RS_MAXPASS];
    u8 bits[RS_MAXPASS];
};

__device__ __forceinline__ u32 rs_digit(const ThRec &r, int word, int shift, u32 mask) {
    return word ? ((r.k1 >> shift) & mask) : ((u32)(r.k0 >> shift) & mask);
}

// ---- 1. all histograms in one read --------------------------------------------------------------------
__global__ void __launch_bounds__(RS_THREADS) rs_hist_kernel(const ThRec *__restrict__ in, u32 n, RsPlan plan,
                                                             u32 *__restrict__ ghist) {
    __shared__ u32 h[RS_MAXPASS * RS_RADIX];
    for (int i = threadIdx.x; i < plan.npass * RS_RADIX; i += RS_THREADS) h[i] = 0;
    __syncthreads();
    const u32 lane = threadIdx.x & 31;
    const u32 stride = gridDim.x * RS_THREADS;
    const u32 rounds = (n + stride - 1) / stride;   // uniform trip count: every lane joins every match
    for (u32 it = 0; it < rounds; ++it) {
        const u32 i = it * stride + blockIdx.x * RS_THREADS + threadIdx.x;
        const bool valid = i < n;
        ThRec r;
        r.k0 = 0; r.k1 = 0; r.idx = 0;
        if (valid) r = in[i];
        for (int p = 0; p < plan.npass; ++p) {
            const u32 mask = (1u << plan.bits[p]) - 1;
            const u32 d = valid ? rs_digit(r, plan.word[p], plan.shift[p], mask) : RS_RADIX;
            // High digits of a timestamp are the same for the whole warp: one atomic for all 32 lanes.  Otherwise
            // the digits are spread out and plain shared-memory atomics rarely collide.  (__match_any_sync on every
            // digit of every record made this kernel 3x the cost of a full sort pass.)
            int same = 0;
            __match_all_sync(0xffffffffu, d, &same);
            if (same) {
                if (lane == 0 && valid) atomicAdd(&h[p * RS_RADIX + d], 32u);
            } else if (valid) {
                atomicAdd(&h[p * RS_RADIX + d], 1u);
            }
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < plan.npass * RS_RADIX; i += RS_THREADS)
        if (h[i]) atomicAdd(&ghist[i], h[i]);
}

// ---- 2. exclusive scan of each pass's 256 bins ---------------------------------------------------------
__global__ void __launch_bounds__(RS_RADIX) rs_scan_kernel(u32 *__restrict__ ghist) {
    __shared__ u32 wtot[RS_RADIX / 32];
    u32 *h = ghist + blockIdx.x * RS_RADIX;
    const u32 lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const u32 c = h[threadIdx.x];
    u32 inc = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        u32 t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= (u32)o) inc += t;
    }
    if (lane == 31) wtot[warp] = inc;
    __syncthreads();
    u32 pre = 0;
    for (u32 w = 0; w < warp; ++w) pre += wtot[w];
    h[threadIdx.x] = pre + inc - c;
}

// ---- 3. one digit pass ------------------------------------------------------------------------------------
struct RsSmem {
    ThRec rec[RS_TILE];
    u32 whist[RS_WARPS][RS_RADIX];
    u32 dstart[RS_RADIX];
    u32 gbase[RS_RADIX];
    u32 wtot[RS_WARPS];
    u32 tile;
};

__global__ void __launch_bounds__(RS_THREADS, 4) rs_onesweep_kernel(const ThRec *__restrict__ in, ThRec *__restrict__ out,
                                                                 u32 n, int word, int shift, u32 mask,
                                                                 const u32 *__restrict__ ghist_excl, u32 *tile_counter,
                                                                 volatile u32 *lookback, int *err) {
    extern __shared__ __align__(16) unsigned char rs_smem_raw[];
    RsSmem &s = *reinterpret_cast<RsSmem *>(rs_smem_raw);
    const u32 tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    if (tid == 0) s.tile = atomicAdd(tile_counter, 1u);
    for (u32 i = tid; i < RS_WARPS * RS_RADIX; i += RS_THREADS) (&s.whist[0][0])[i] = 0;
    __syncthreads();
    const u32 tile = s.tile;
    const u32 tile_base = tile * RS_TILE;
    const u32 wbase = tile_base + warp * (RS_IPT * 32);

    // warp-striped load: item j of lane l is record wbase + 32 j + l, so (j, lane) order = memory order
    ThRec r[RS_IPT];
#pragma unroll
    for (int j = 0; j < RS_IPT; ++j) {
        const u32 gi = wbase + j * 32 + lane;
        if (gi < n) r[j] = in[gi];
        else { r[j].k0 = 0; r[j].k1 = 0; r[j].idx = 0; }
    }

    // warp-level multi-split: rank of each record among the warp's records with the same digit
    u32 pk[RS_IPT];   // digit (9 bits, 256 = padding) | rank-in-warp << 9
#pragma unroll
    for (int j = 0; j < RS_IPT; ++j) {
        const bool valid = (wbase + j * 32 + lane) < n;
        const u32 d = valid ? rs_digit(r[j], word, shift, mask) : (u32)RS_RADIX;
        const u32 peers = __match_any_sync(0xffffffffu, d);
        const u32 lt = peers & ((1u << lane) - 1);
        u32 base = 0;
        if (valid) base = s.whist[warp][d];
        __syncwarp();
        if (valid && lt == 0) s.whist[warp][d] = base + __popc(peers);
        __syncwarp();
        pk[j] = d | ((base + __popc(lt)) << 9);
    }
    __syncthreads();

    // thread b owns digit b: per-warp exclusive offsets and the tile's count
    const u32 b = tid;
    u32 cnt = 0;
#pragma unroll
    for (int w = 0; w < RS_WARPS; ++w) {
        const u32 t = s.whist[w][b];
        s.whist[w][b] = cnt;
        cnt += t;
    }
    lookback[(size_t)tile * RS_RADIX + b] = (tile == 0 ? RS_FLAG_INC : RS_FLAG_AGG) | cnt;

    // exclusive scan of the counts over the 256 digits -> where each digit starts inside the tile
    u32 inc = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        u32 t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= (u32)o) inc += t;
    }
    if (lane == 31) s.wtot[warp] = inc;
    __syncthreads();
    u32 wpre = 0;
    for (u32 w = 0; w < warp; ++w) wpre += s.wtot[w];
    const u32 ds = wpre + inc - cnt;
    s.dstart[b] = ds;

    // decoupled look-back over earlier tiles for digit b.  The walk is inherently sequential (stop at the first
    // tile whose inclusive prefix is known), but the loads are not: four predecessors are requested at once and
    // then consumed in order, so a tile that starts in a wave of ~450 resident tiles pays ~1/4 of the round trips.
    u32 excl = 0;
    if (tile > 0) {
        i64 t = (i64)tile - 1;
        u32 spins = 0;
        bool done = false;
        while (!done) {
            u32 v[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const i64 idx = t - k;
                v[k] = RS_FLAG_INC + 0u;
                if (idx >= 0) v[k] = lookback[(size_t)idx * RS_RADIX + b];
            }
            int used = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                if (done || used < k) continue;        // stop consuming after a not-ready or an inclusive entry
                const u32 f = v[k] >> 30;
                if (f == 0) continue;                  // not published yet: retry from here
                excl += v[k] & RS_VAL_MASK;
                used = k + 1;
                if (f == 2) done = true;
            }
            t -= used;
            if (!done && used < 4) {                   // hit an unpublished tile
                if (++spins > RS_SPIN_LIMIT) { *err = 1; break; }
                __nanosleep(40);
            } else {
                spins = 0;
            }
        }
        lookback[(size_t)tile * RS_RADIX + b] = RS_FLAG_INC | ((excl + cnt) & RS_VAL_MASK);
    }
    s.gbase[b] = ghist_excl[b] + excl - ds;   // modular: + position-in-tile (>= ds) gives the global slot
    __syncthreads();

    // regroup by digit in shared memory, then leave in coalesced runs
#pragma unroll
    for (int j = 0; j < RS_IPT; ++j) {
        const u32 d = pk[j] & 511u;
        if (d < RS_RADIX) {
            const u32 pos = s.dstart[d] + s.whist[warp][d] + (pk[j] >> 9);
            s.rec[pos] = r[j];
        }
    }
    __syncthreads();
    const u32 left = n - tile_base;
    const u32 nvalid = left < (u32)RS_TILE ? left : (u32)RS_TILE;
    for (u32 k = tid; k < nvalid; k += RS_THREADS) {
        const ThRec x = s.rec[k];
        const u32 pos = s.gbase[rs_digit(x, word, shift, mask)] + k;
        if (pos < n) out[pos] = x;   // the bound only matters after a watchdog trip
    }
}

void rs_add_passes(RsPlan &plan, int word, int nbits) {
    if (nbits <= 0) return;
    const int np = (nbits + 7) / 8;
    const int w = (nbits + np - 1) / np;   // balance digit widths, e.g. 47 bits -> 6 passes of 8,8,8,8,8,7
    for (int i = 0; i < np; ++i) {
        const int sh = i * w;
        int bits = nbits - sh < w ? nbits - sh : w;
        plan.word[plan.npass] = (u8)word;
        plan.shift[plan.npass] = (u8)sh;
        plan.bits[plan.npass] = (u8)bits;
        ++plan.npass;
    }
}

}  // namespace

// Small inputs (a delta of a few hundred rows, a kill list): one CTA, bitonic network in shared memory on
// (k1, k0, input position) -- the position makes it stable like the LSD passes -- instead of histogram + scan + one
// launch per digit and a host synchronisation for the look-back watchdog.
constexpr int RS_SMALL = 2048;
__global__ void __launch_bounds__(1024) rs_small_sort_kernel(const ThRec *__restrict__ in, ThRec *__restrict__ out, u32 n,
                                                             u64 mask0, u32 mask1) {
    __shared__ ThRec s[RS_SMALL];
    __shared__ u32 pos[RS_SMALL];
    u32 P = 2;
    while (P < n) P <<= 1;
    for (u32 i = threadIdx.x; i < P; i += blockDim.x) {
        if (i < n) { s[i] = in[i]; pos[i] = i; }
        else { ThRec pad; pad.k0 = ~0ull; pad.k1 = ~0u; pad.idx = 0; s[i] = pad; pos[i] = ~0u; }
    }
    __syncthreads();
    for (u32 k = 2; k <= P; k <<= 1) {
        for (u32 j = k >> 1; j > 0; j >>= 1) {
            for (u32 i = threadIdx.x; i < P; i += blockDim.x) {
                const u32 l = i ^ j;
                if (l > i) {
                    const ThRec x = s[i], y = s[l];
                    const u32 px = pos[i], py = pos[l];
                    const u32 x1 = x.k1 & mask1, y1 = y.k1 & mask1;
                    const u64 x0 = x.k0 & mask0, y0 = y.k0 & mask0;
                    bool x_after_y;                          // pads (pos = ~0) sort behind every record
                    if (px == ~0u || py == ~0u) x_after_y = px == ~0u && py != ~0u;
                    else x_after_y = x1 != y1 ? x1 > y1 : (x0 != y0 ? x0 > y0 : px > py);
                    const bool ascending = (i & k) == 0;
                    if (x_after_y == ascending) { s[i] = y; s[l] = x; pos[i] = py; pos[l] = px; }
                }
            }
            __syncthreads();
        }
    }
    for (u32 i = threadIdx.x; i < n; i += blockDim.x) out[i] = s[i];
}

int th_radix_sort(ThRec *a, ThRec *b, i64 n, int k0_bits, int k1_bits, ThRec **sorted, int *passes,
                  cudaStream_t st) {
    *sorted = a;
    if (passes) *passes = 0;
    if (n <= 1) return TH_OK;
    if (n >= (1ll << 30)) {
        th_set_error("th_radix_sort: n=%lld exceeds the 2^30-1 record limit of the look-back words", (long long)n);
        return TH_EINVAL;
    }
    if (k0_bits > 64) k0_bits = 64;
    if (k1_bits > 32) k1_bits = 32;
    RsPlan plan;
    plan.npass = 0;
    rs_add_passes(plan, 0, k0_bits);
    rs_add_passes(plan, 1, k1_bits);
    if (plan.npass == 0) return TH_OK;
    if (n <= RS_SMALL) {
        const u64 mask0 = k0_bits >= 64 ? ~0ull : ((1ull << k0_bits) - 1ull);
        const u32 mask1 = k1_bits >= 32 ? ~0u : ((1u << k1_bits) - 1u);
        rs_small_sort_kernel<<<1, 1024, 0, st>>>(a, b, (u32)n, mask0, mask1);
        TH_CUDA(cudaGetLastError());
        *sorted = b;
        if (passes) *passes = 1;
        return TH_OK;
    }

    // once per device; the attribute call is idempotent, so two threads racing on the first use is harmless
    static std::atomic<bool> attr_set[64];
    int dev = 0;
    TH_CUDA(cudaGetDevice(&dev));
    if (dev >= 64 || !attr_set[dev].load(std::memory_order_acquire)) {
        TH_CUDA(cudaFuncSetAttribute(rs_onesweep_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)sizeof(RsSmem)));
        if (dev < 64) attr_set[dev].store(true, std::memory_order_release);
    }

    const u32 tiles = (u32)((n + RS_TILE - 1) / RS_TILE);
    const size_t ghist_bytes = (size_t)plan.npass * RS_RADIX * sizeof(u32);
    const size_t look_bytes = ((size_t)tiles * RS_RADIX + 4) * sizeof(u32);   // + ticket + err + pad
    u32 *ghist = nullptr, *look = nullptr;
    TH_CUDA(th_malloc_async((void **)&ghist, ghist_bytes, st));
    cudaError_t e = th_malloc_async((void **)&look, look_bytes, st);
    if (e != cudaSuccess) {
        cudaFreeAsync(ghist, st);
        th_set_error("th_radix_sort: th_malloc_async(%zu) -> %s", look_bytes, cudaGetErrorString(e));
        return TH_ENOMEM;
    }
    int *err_dev = nullptr;
    e = th_malloc_async((void **)&err_dev, sizeof(int), st);
    int rc = TH_OK;
    int err_host = 0;
    do {
        if (e != cudaSuccess) { rc = TH_ENOMEM; th_set_error("th_radix_sort: err flag alloc failed"); break; }
#define RS_CK(x) if ((e = (x)) != cudaSuccess) { rc = TH_ECUDA; th_set_error("th_radix_sort: %s -> %s", #x, cudaGetErrorString(e)); break; }
        RS_CK(cudaMemsetAsync(ghist, 0, ghist_bytes, st));
        RS_CK(cudaMemsetAsync(err_dev, 0, sizeof(int), st));
        // two CTAs per SM at most: every CTA ends with up to npass x 256 global atomics
        rs_hist_kernel<<<th_grid_for(n, RS_THREADS * 8, 2), RS_THREADS, 0, st>>>(a, (u32)n, plan, ghist);
        rs_scan_kernel<<<plan.npass, RS_RADIX, 0, st>>>(ghist);
        RS_CK(cudaGetLastError());
        ThRec *src = a, *dst = b;
        bool failed = false;
        for (int p = 0; p < plan.npass; ++p) {
            if ((e = cudaMemsetAsync(look, 0, look_bytes, st)) != cudaSuccess) { failed = true; break; }
            u32 *ticket = look + (size_t)tiles * RS_RADIX;
            rs_onesweep_kernel<<<tiles, RS_THREADS, sizeof(RsSmem), st>>>(
                src, dst, (u32)n, plan.word[p], plan.shift[p], (1u << plan.bits[p]) - 1, ghist + p * RS_RADIX, ticket,
                look, err_dev);
            if ((e = cudaGetLastError()) != cudaSuccess) { failed = true; break; }
            ThRec *t = src; src = dst; dst = t;
        }
        if (failed) { rc = TH_ECUDA; th_set_error("th_radix_sort: pass launch -> %s", cudaGetErrorString(e)); break; }
        *sorted = src;
        if (passes) *passes = plan.npass;
        RS_CK(cudaMemcpyAsync(&err_host, err_dev, sizeof(int), cudaMemcpyDeviceToHost, st));
        RS_CK(cudaStreamSynchronize(st));
#undef RS_CK
        if (err_host) { rc = TH_EINTERNAL; th_set_error("th_radix_sort: look-back watchdog tripped"); }
    } while (0);
    cudaFreeAsync(ghist, st);
    cudaFreeAsync(look, st);
    if (err_dev) cudaFreeAsync(err_dev, st);
    return rc;
}
