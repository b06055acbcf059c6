// Peer group for the fused K2 + all-gather (the reference has no multi-GPU path at all).
//
// One process per GPU.  Every rank allocates one buffer with cudaMalloc -- two generations of 8-byte slots
// {bitmap word, epoch} [world][words_per_rank] plus barrier flags -- exports it as a CUDA IPC handle, and maps the
// buffers of all other ranks (NVLink peer access is enabled lazily by the mapping).  The handles travel through
// whatever the host side has (torch.distributed.all_gather_object in sharding.py).  The compact gathered words
// the caller reads live in a second, private allocation.
//
// Two generations: rank A can run at most one step ahead of rank B (A's unpack kernel of step k+1 needs B's slots
// of step k+1, which B writes only after finishing, in stream order, everything of step k), so step k+2 never
// overwrites slots of step k that somebody still polls.
#include "th_common.cuh"
#include <string.h>
#include <stdlib.h>

// How long a receiver polls for a peer's words before the step is reported as failed.  Rank skew is normal (a GC pause,
// a first-call module load, host staging on one rank), so the default is generous; a dead peer still ends the kernel.
u64 th_peer_timeout_ns() {
    static const u64 v = [] {
        const char *e = getenv("TH_PEER_TIMEOUT_MS");
        long long ms = e ? atoll(e) : 2000;
        if (ms < 1) ms = 1;
        if (ms > 600000) ms = 600000;
        return (u64)ms * 1000000ull;
    }();
    return v;
}

extern "C" int th_peer_create(int device, int world, int rank, int64_t words_per_rank, th_peer **out,
                              void *handle_out_64) {
    TH_REQUIRE(out && handle_out_64, "th_peer_create: NULL output");
    TH_REQUIRE(world >= 1 && world <= TH_MAX_PEERS, "th_peer_create: world %d outside [1, %d]", world, TH_MAX_PEERS);
    TH_REQUIRE(rank >= 0 && rank < world, "th_peer_create: rank %d outside [0, %d)", rank, world);
    TH_REQUIRE(words_per_rank > 0 && words_per_rank < (1ll << 28), "th_peer_create: bad words_per_rank");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle is 64 bytes");
    int prev = 0;
    TH_CUDA(cudaGetDevice(&prev));
    if (prev != device) TH_CUDA(cudaSetDevice(device));
    th_peer *p = new th_peer();
    memset(p, 0, sizeof(*p));
    p->device = device; p->world = world; p->rank = rank; p->words_per_rank = words_per_rank;
    const size_t bytes = th_peer_bytes(world, words_per_rank);
    cudaError_t e = cudaMalloc(&p->base, bytes);
    if (e == cudaSuccess) e = cudaMemset(p->base, 0, bytes);     // epoch 0 = "nothing has arrived"
    if (e == cudaSuccess) e = cudaMalloc((void **)&p->out, (size_t)2 * world * words_per_rank * 4);
    if (e == cudaSuccess) e = cudaMemset(p->out, 0, (size_t)2 * world * words_per_rank * 4);
    cudaIpcMemHandle_t hnd;
    if (e == cudaSuccess) e = cudaIpcGetMemHandle(&hnd, p->base);
    if (prev != device) cudaSetDevice(prev);
    if (e != cudaSuccess) {
        th_set_error("th_peer_create: %s", cudaGetErrorString(e));
        if (p->base) cudaFree(p->base);
        if (p->out) cudaFree(p->out);
        delete p;
        return e == cudaErrorMemoryAllocation ? TH_ENOMEM : TH_ECUDA;
    }
    memcpy(handle_out_64, &hnd, 64);
    *out = p;
    return TH_OK;
}

// all_handles: world x 64 bytes in rank order (this rank's own entry is ignored).  With local_ptrs != NULL the
// entries are device pointers of buffers in THIS process instead (single-process emulation of several ranks for
// tests; IPC handles cannot be opened by the process that exported them).
extern "C" int th_peer_connect(th_peer *p, const void *all_handles, void *const *local_ptrs) {
    TH_REQUIRE(p != nullptr && p->base, "th_peer_connect: peer group not created");
    TH_REQUIRE(all_handles || local_ptrs, "th_peer_connect: no handles");
    int prev = 0;
    TH_CUDA(cudaGetDevice(&prev));
    if (prev != p->device) TH_CUDA(cudaSetDevice(p->device));
    int rc = TH_OK;
    for (int r = 0; r < p->world && rc == TH_OK; ++r) {
        if (r == p->rank) { p->peer_base[r] = p->base; continue; }
        if (local_ptrs) { p->peer_base[r] = local_ptrs[r]; continue; }
        cudaIpcMemHandle_t hnd;
        memcpy(&hnd, (const char *)all_handles + (size_t)r * 64, 64);
        void *ptr = nullptr;
        cudaError_t e = cudaIpcOpenMemHandle(&ptr, hnd, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) {
            th_set_error("th_peer_connect: cudaIpcOpenMemHandle(rank %d) -> %s", r, cudaGetErrorString(e));
            rc = TH_ECUDA;
        }
        p->peer_base[r] = ptr;
    }
    if (prev != p->device) cudaSetDevice(prev);
    if (rc == TH_OK) p->connected = local_ptrs ? 2 : 1;
    return rc;
}

extern "C" int th_peer_local_base(th_peer *p, void **out) {
    TH_REQUIRE(p != nullptr && out, "th_peer_local_base: NULL");
    *out = p->base;
    return TH_OK;
}

extern "C" int th_peer_destroy(th_peer *p) {
    if (!p) return TH_OK;
    int prev = 0;
    cudaGetDevice(&prev);
    cudaSetDevice(p->device);
    cudaDeviceSynchronize();
    if (p->connected == 1)
        for (int r = 0; r < p->world; ++r)
            if (r != p->rank && p->peer_base[r]) cudaIpcCloseMemHandle(p->peer_base[r]);
    if (p->base) cudaFree(p->base);
    if (p->out) cudaFree(p->out);
    cudaSetDevice(prev);
    delete p;
    return TH_OK;
}
