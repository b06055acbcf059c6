// Error plumbing, staging scratch and handle lifecycle of libth_b200.so.
#include "th_common.cuh"
#include <string.h>

static thread_local char g_th_err[512] = "";

void th_set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_th_err, sizeof(g_th_err), fmt, ap);
    va_end(ap);
}

extern "C" const char *th_last_error(void) { return g_th_err; }
extern "C" int th_version(void) { return TH_ABI_VERSION; }

// ---- the library's stream-ordered pool ------------------------------------------------------------------
#include <mutex>
static std::mutex g_pool_mutex;
static cudaMemPool_t g_pools[64] = {nullptr};

cudaError_t th_malloc_async(void **p, size_t bytes, cudaStream_t st) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= 64) return cudaMallocAsync(p, bytes, st);
    cudaMemPool_t pool;
    {
        std::lock_guard<std::mutex> lock(g_pool_mutex);
        if (g_pools[dev] == nullptr) {
            cudaMemPoolProps props = {};
            props.allocType = cudaMemAllocationTypePinned;
            props.handleTypes = cudaMemHandleTypeNone;
            props.location.type = cudaMemLocationTypeDevice;
            props.location.id = dev;
            e = cudaMemPoolCreate(&g_pools[dev], &props);
            if (e != cudaSuccess) { g_pools[dev] = nullptr; return e; }
            unsigned long long keep = ~0ull;
            cudaMemPoolSetAttribute(g_pools[dev], cudaMemPoolAttrReleaseThreshold, &keep);
        }
        pool = g_pools[dev];
    }
    return cudaMallocFromPoolAsync(p, bytes, pool, st);
}

// Diagnostics: what the library's staging pool on `device` holds (0s when no call has needed it yet).
extern "C" int th_pool_stats(int device, int64_t *reserved_bytes, int64_t *used_bytes, int64_t *used_high_bytes,
                             int reset_high) {
    TH_REQUIRE(device >= 0 && device < 64, "th_pool_stats: device %d outside [0, 64)", device);
    cudaMemPool_t pool;
    {
        std::lock_guard<std::mutex> lock(g_pool_mutex);
        pool = g_pools[device];
    }
    unsigned long long r = 0, u = 0, hi = 0;
    if (pool != nullptr) {
        TH_CUDA(cudaMemPoolGetAttribute(pool, cudaMemPoolAttrReservedMemCurrent, &r));
        TH_CUDA(cudaMemPoolGetAttribute(pool, cudaMemPoolAttrUsedMemCurrent, &u));
        TH_CUDA(cudaMemPoolGetAttribute(pool, cudaMemPoolAttrUsedMemHigh, &hi));
        if (reset_high) {
            unsigned long long zero = 0;
            TH_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrUsedMemHigh, &zero));
        }
    }
    if (reserved_bytes) *reserved_bytes = (int64_t)r;
    if (used_bytes) *used_bytes = (int64_t)u;
    if (used_high_bytes) *used_high_bytes = (int64_t)hi;
    return TH_OK;
}

// ---- per-thread streams for host-array calls ------------------------------------------------------------------
namespace {
struct ThreadStreams {
    cudaStream_t s[64];
    ThreadStreams() { for (auto &x : s) x = nullptr; }
    ~ThreadStreams() {
        // the thread ends: its streams have no work left (every host-array call synchronises before returning)
        for (auto x : s) if (x) cudaStreamDestroy(x);
    }
};
thread_local ThreadStreams g_thread_streams;
}  // namespace

cudaStream_t th_call_stream(void *stream, int mem, int device) {
    if (stream != nullptr || mem != TH_MEM_HOST || device < 0 || device >= 64) return (cudaStream_t)stream;
    cudaStream_t &s = g_thread_streams.s[device];
    if (s == nullptr && cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking) != cudaSuccess) {
        cudaGetLastError();
        s = nullptr;                 // fall back to the default stream rather than fail the call
    }
    return s;
}

// ---- per-thread page-locked slab for controller-sized host calls ---------------------------------------------------
// One would_interfere from a REST controller is 32 bytes in and 4 bytes out: staging that through five stream-ordered
// allocations and five copies costs more than the search.  Small host-array batches are instead copied (by the CPU) into
// a page-locked, device-mapped slab of the calling thread, which the kernel reads and writes in place: no cudaMemcpy,
// no allocation, one launch and one synchronisation.
namespace {
struct ThreadSlab {
    void *host[64];
    ThreadSlab() { for (auto &x : host) x = nullptr; }
    ~ThreadSlab() { for (auto x : host) if (x) cudaFreeHost(x); }
};
thread_local ThreadSlab g_thread_slab;
}  // namespace

void *th_thread_slab(int device, void **dev_alias) {
    if (device < 0 || device >= 64) return nullptr;
    void *&h = g_thread_slab.host[device];
    if (h == nullptr && cudaHostAlloc(&h, TH_SLAB_BYTES, cudaHostAllocMapped | cudaHostAllocPortable) != cudaSuccess) {
        cudaGetLastError();
        h = nullptr;
        return nullptr;
    }
    if (cudaHostGetDevicePointer(dev_alias, h, 0) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    return h;
}

// ---- ThScratch ---------------------------------------------------------------------------------------
int ThScratch::alloc(void **out, size_t bytes) {
    if (count >= (int)(sizeof(ptrs) / sizeof(ptrs[0]))) {
        th_set_error("ThScratch: too many staging buffers");
        return TH_EINTERNAL;
    }
    if (bytes == 0) bytes = 16;
    void *p = nullptr;
    cudaError_t e = th_malloc_async(&p, bytes, st);
    if (e != cudaSuccess) {
        th_set_error("th_malloc_async(%zu bytes) -> %s", bytes, cudaGetErrorString(e));
        return TH_ENOMEM;
    }
    ptrs[count++] = p;
    *out = p;
    return TH_OK;
}

int ThScratch::in(const void *src, size_t bytes, int mem, const void **out) {
    if (mem == TH_MEM_DEVICE) { *out = src; return TH_OK; }
    void *d = nullptr;
    TH_TRY(alloc(&d, bytes));
    if (bytes) TH_CUDA(cudaMemcpyAsync(d, src, bytes, cudaMemcpyHostToDevice, st));
    *out = d;
    return TH_OK;
}

int ThScratch::out(void *dst, size_t bytes, int mem, void **out_dev) {
    if (dst == nullptr) { *out_dev = nullptr; return TH_OK; }
    if (mem == TH_MEM_DEVICE) { *out_dev = dst; return TH_OK; }
    return alloc(out_dev, bytes);
}

int th_copy_back(void *host_dst, const void *dev_src, size_t bytes, int mem, cudaStream_t st) {
    if (mem == TH_MEM_DEVICE || host_dst == nullptr || bytes == 0) return TH_OK;
    TH_CUDA(cudaMemcpyAsync(host_dst, dev_src, bytes, cudaMemcpyDeviceToHost, st));
    return TH_OK;
}

// ---- handle lifecycle ---------------------------------------------------------------------------------
static void th_snapshot_release(th_snapshot *h) {
    if (h->seg) cudaFree(h->seg);
    if (h->p1) cudaFree(h->p1);
    if (h->p2) cudaFree(h->p2);
    if (h->p3) cudaFree(h->p3);
    if (h->p4) cudaFree(h->p4);
    if (h->uuid) cudaFree(h->uuid);
    if (h->grid) cudaFree(h->grid);
    if (h->ugrid) cudaFree(h->ugrid);
    h->seg = nullptr; h->p1 = nullptr; h->p2 = nullptr; h->p3 = nullptr; h->p4 = nullptr; h->uuid = nullptr; h->grid = nullptr; h->ugrid = nullptr;
    h->loaded = 0; h->n = 0; h->nt = 0; h->n_uuids = 0; h->device_bytes = 0;
    h->cap_nt = 0; h->cap_uuids = 0; h->cap_cells = 0; h->cap_ugrid = 0;
}

extern "C" int th_snapshot_create(int device, th_snapshot **out) {
    TH_REQUIRE(out != nullptr, "th_snapshot_create: out is NULL");
    int count = 0;
    TH_CUDA(cudaGetDeviceCount(&count));
    TH_REQUIRE(device >= 0 && device < count, "th_snapshot_create: device %d not in [0, %d)", device, count);
    cudaDeviceProp prop;
    TH_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10) {
        th_set_error("th_snapshot_create: device %d is sm_%d%d; this library is built for sm_100a only", device,
                     prop.major, prop.minor);
        return TH_ESTATE;
    }
    th_snapshot *h = new th_snapshot();
    memset(h, 0, sizeof(*h));
    h->device = device;
    *out = h;
    return TH_OK;
}

extern "C" int th_snapshot_destroy(th_snapshot *h) {
    if (!h) return TH_OK;
    int prev = 0;
    cudaGetDevice(&prev);
    cudaSetDevice(h->device);
    th_snapshot_release(h);
    cudaSetDevice(prev);
    delete h;
    return TH_OK;
}

extern "C" int th_snapshot_info(const th_snapshot *h, int64_t *n, int64_t *n_uuids, int64_t *device_bytes,
                                int64_t *radix_passes) {
    TH_REQUIRE(h != nullptr, "th_snapshot_info: handle is NULL");
    if (n) *n = h->n;
    if (n_uuids) *n_uuids = h->n_uuids;
    if (device_bytes) *device_bytes = h->device_bytes;
    if (radix_passes) *radix_passes = h->radix_passes;
    return TH_OK;
}

// used by th_snapshot.cu
void th_snapshot_release_internal(th_snapshot *h) { th_snapshot_release(h); }
