// api.cu — context, error channel, device buffer pool and pinned host memory of
// libigloo_cuda.so (include/igloo_cuda.h).  The reference has no device layer;
// the pool and the pinned allocations are the "device-resident batch buffer
// pool" and "pinned-memory batch upload path" BASELINE.json's north_star names.
#include "internal.h"

#include <chrono>
#include <cstdlib>
#include <cstring>

namespace igl {

static thread_local std::string g_last_error;

void set_error(const char* fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_last_error = buf;
}

int32_t fail(int32_t code, const char* fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_last_error = buf;
    return code;
}

// ---- host-time tracing (IGL_TRACE=1) ----------------------------------------------------------
namespace {
struct TraceTable {
    std::mutex mu;
    std::map<std::string, std::pair<double, long>> acc;
    ~TraceTable() {
        if (acc.empty()) return;
        fprintf(stderr, "[igl trace]");
        for (auto& kv : acc) fprintf(stderr, " %s=%.2fms/%ld", kv.first.c_str(), kv.second.first * 1e3, kv.second.second);
        fprintf(stderr, "\n");
    }
};
TraceTable& trace_table() { static TraceTable t; return t; }
}  // namespace
bool trace_on() { static const bool on = getenv("IGL_TRACE") && getenv("IGL_TRACE")[0] == '1'; return on; }
double trace_now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
void trace_add(const char* name, double seconds) {
    TraceTable& t = trace_table();
    std::lock_guard<std::mutex> g(t.mu);
    auto& e = t.acc[name];
    e.first += seconds; e.second++;
}

static size_t bucket_of(size_t bytes) {
    size_t b = 256;
    while (b < bytes) b <<= 1;
    // above 64 MiB round to 2 MiB multiples instead of powers of two: a 32 GB
    // join output must not cost 64 GB of the 180 GB HBM.
    if (b > (64ull << 20)) b = (bytes + (2ull << 20) - 1) & ~((2ull << 20) - 1);
    return b;
}

int32_t pool_alloc(igl_ctx* ctx, size_t bytes, void** out) {
    if (bytes == 0) bytes = 1;
    size_t b = bucket_of(bytes);
    {
        std::lock_guard<std::mutex> g(ctx->mu);
        auto it = ctx->free_lists.find(b);
        if (it != ctx->free_lists.end() && !it->second.empty()) {
            void* p = it->second.back();
            it->second.pop_back();
            ctx->live[p] = b;
            *out = p;
            return IGL_OK;
        }
    }
    void* p = nullptr;
    cudaError_t e = cudaMalloc(&p, b);
    if (e != cudaSuccess) {
        // release cached blocks and retry once
        std::vector<void*> to_free;
        {
            std::lock_guard<std::mutex> g(ctx->mu);
            for (auto& kv : ctx->free_lists) {
                for (void* q : kv.second) { to_free.push_back(q); ctx->pool_bytes -= kv.first; }
                kv.second.clear();
            }
        }
        (void)cudaGetLastError();
        for (void* q : to_free) cudaFree(q);
        e = cudaMalloc(&p, b);
        if (e != cudaSuccess) {
            (void)cudaGetLastError();
            return fail(IGL_ERR_OOM, "device allocation of %zu bytes failed: %s", b, cudaGetErrorString(e));
        }
    }
    std::lock_guard<std::mutex> g(ctx->mu);
    ctx->live[p] = b;
    ctx->pool_bytes += b;
    *out = p;
    return IGL_OK;
}

void pool_free(igl_ctx* ctx, void* p) {
    if (!p) return;
    std::lock_guard<std::mutex> g(ctx->mu);
    auto it = ctx->live.find(p);
    if (it == ctx->live.end()) return;
    size_t b = it->second;
    ctx->live.erase(it);
    if (b > (40ull << 30)) {  // very large blocks go straight back (cached blocks are released on OOM anyway)
        ctx->pool_bytes -= b;
        cudaFree(p);
        return;
    }
    ctx->free_lists[b].push_back(p);
}

}  // namespace igl

using namespace igl;

extern "C" {

const char* igl_last_error(void) { return g_last_error.c_str(); }
int32_t igl_version(void) { return 100; }

int32_t igl_ctx_create(const igl_config* cfg, igl_ctx** out) {
    if (!out) return fail(IGL_ERR_INVALID_ARG, "igl_ctx_create: out is NULL");
    *out = nullptr;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        (void)cudaGetLastError();
        return fail(IGL_ERR_CUDA, "no CUDA device available (%s); igloo-cuda has no CPU fallback",
                    e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
    }
    int dev = cfg ? cfg->device : 0;
    if (dev < 0 || dev >= ndev) return fail(IGL_ERR_INVALID_ARG, "device %d out of range (have %d)", dev, ndev);
    DeviceGuard g(dev);
    cudaDeviceProp prop{};
    IGL_CUDA_OK(cudaGetDeviceProperties(&prop, dev));
    if (prop.major < 10)
        return fail(IGL_ERR_UNSUPPORTED, "device %d is sm_%d%d; this library is built for sm_100a only", dev,
                    prop.major, prop.minor);
    // Experiment knob only: cudaLimitMaxL2FetchGranularity.  Measured on B200: no
    // effect on the ~128 B of HBM traffic each random 16-byte table read costs
    // (profiles/README.md), so nothing is set unless the variable is present.
    if (const char* env = getenv("IGL_L2_FETCH_GRANULARITY")) {
        (void)cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, (size_t)atoi(env));
        (void)cudaGetLastError();
    }
    igl_ctx* c = new igl_ctx();
    c->device = dev;
    c->sm_count = prop.multiProcessorCount;
    if (cfg) c->cfg = *cfg;
    if (c->cfg.coalesce_rows <= 0) c->cfg.coalesce_rows = 4ll << 20;
    if (!(c->cfg.hash_load_factor > 0.05 && c->cfg.hash_load_factor <= 0.9)) c->cfg.hash_load_factor = 0.5;
    if (c->cfg.pinned_ring_bytes <= 0) c->cfg.pinned_ring_bytes = 256ll << 20;
    cudaError_t se = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking);
    if (se != cudaSuccess) {
        delete c;
        return fail(IGL_ERR_CUDA, "cudaStreamCreate failed: %s", cudaGetErrorString(se));
    }
    *out = c;
    return IGL_OK;
}

void igl_ctx_destroy(igl_ctx* ctx) {
    if (!ctx) return;
    DeviceGuard g(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    comm_release(ctx);
    for (auto& kv : ctx->free_lists)
        for (void* p : kv.second) cudaFree(p);
    for (auto& kv : ctx->live) cudaFree(kv.first);
    for (auto& kv : ctx->pinned) cudaFreeHost(kv.first);
    cudaStreamDestroy(ctx->stream);
    delete ctx;
}

int32_t igl_ctx_device(const igl_ctx* ctx) { return ctx ? ctx->device : -1; }
int64_t igl_ctx_launch_count(const igl_ctx* ctx) { return ctx ? ctx->launches.load() : 0; }

int32_t igl_dev_alloc(igl_ctx* ctx, size_t bytes, void** out) {
    if (!ctx || !out) return fail(IGL_ERR_INVALID_ARG, "igl_dev_alloc: NULL argument");
    DeviceGuard g(ctx->device);
    return pool_alloc(ctx, bytes, out);
}

int32_t igl_dev_free(igl_ctx* ctx, void* ptr) {
    if (!ctx) return fail(IGL_ERR_INVALID_ARG, "igl_dev_free: NULL context");
    DeviceGuard g(ctx->device);
    pool_free(ctx, ptr);
    return IGL_OK;
}

int32_t igl_host_alloc_pinned(igl_ctx* ctx, size_t bytes, void** out) {
    if (!ctx || !out) return fail(IGL_ERR_INVALID_ARG, "igl_host_alloc_pinned: NULL argument");
    DeviceGuard g(ctx->device);
    void* p = nullptr;
    cudaError_t e = cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocDefault);
    if (e != cudaSuccess) {
        (void)cudaGetLastError();
        return fail(IGL_ERR_OOM, "pinned host allocation of %zu bytes failed: %s", bytes, cudaGetErrorString(e));
    }
    std::lock_guard<std::mutex> lk(ctx->mu);
    ctx->pinned[p] = bytes;
    *out = p;
    return IGL_OK;
}

int32_t igl_host_free_pinned(igl_ctx* ctx, void* ptr) {
    if (!ctx) return fail(IGL_ERR_INVALID_ARG, "igl_host_free_pinned: NULL context");
    if (!ptr) return IGL_OK;
    {
        std::lock_guard<std::mutex> lk(ctx->mu);
        ctx->pinned.erase(ptr);
    }
    DeviceGuard g(ctx->device);
    IGL_CUDA_OK(cudaFreeHost(ptr));
    return IGL_OK;
}

int32_t igl_memcpy_h2d(igl_ctx* ctx, void* dev_dst, const void* host_src, size_t bytes, void* stream) {
    if (!ctx) return fail(IGL_ERR_INVALID_ARG, "igl_memcpy_h2d: NULL context");
    DeviceGuard g(ctx->device);
    IGL_CUDA_OK(cudaMemcpyAsync(dev_dst, host_src, bytes, cudaMemcpyHostToDevice, pick_stream(ctx, stream)));
    return IGL_OK;
}

int32_t igl_memcpy_d2h(igl_ctx* ctx, void* host_dst, const void* dev_src, size_t bytes, void* stream) {
    if (!ctx) return fail(IGL_ERR_INVALID_ARG, "igl_memcpy_d2h: NULL context");
    DeviceGuard g(ctx->device);
    IGL_CUDA_OK(cudaMemcpyAsync(host_dst, dev_src, bytes, cudaMemcpyDeviceToHost, pick_stream(ctx, stream)));
    return IGL_OK;
}

int32_t igl_stream_sync(igl_ctx* ctx, void* stream) {
    if (!ctx) return fail(IGL_ERR_INVALID_ARG, "igl_stream_sync: NULL context");
    DeviceGuard g(ctx->device);
    IGL_CUDA_OK(cudaStreamSynchronize(pick_stream(ctx, stream)));
    return IGL_OK;
}

int32_t igl_ipc_export(igl_ctx* ctx, void* dev_ptr, uint8_t handle_out[64]) {
    if (!ctx || !dev_ptr || !handle_out) return fail(IGL_ERR_INVALID_ARG, "igl_ipc_export: NULL argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "ipc handle size");
    DeviceGuard g(ctx->device);
    cudaIpcMemHandle_t h;
    IGL_CUDA_OK(cudaIpcGetMemHandle(&h, dev_ptr));
    memcpy(handle_out, &h, 64);
    return IGL_OK;
}

int32_t igl_ipc_open(igl_ctx* ctx, const uint8_t handle[64], void** dev_ptr_out) {
    if (!ctx || !handle || !dev_ptr_out) return fail(IGL_ERR_INVALID_ARG, "igl_ipc_open: NULL argument");
    DeviceGuard g(ctx->device);
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, 64);
    IGL_CUDA_OK(cudaIpcOpenMemHandle(dev_ptr_out, h, cudaIpcMemLazyEnablePeerAccess));
    return IGL_OK;
}

int32_t igl_ipc_close(igl_ctx* ctx, void* dev_ptr) {
    if (!ctx) return fail(IGL_ERR_INVALID_ARG, "igl_ipc_close: NULL context");
    DeviceGuard g(ctx->device);
    IGL_CUDA_OK(cudaIpcCloseMemHandle(dev_ptr));
    return IGL_OK;
}

}  // extern "C"
