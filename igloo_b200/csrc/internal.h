// internal.h — host-side internals shared by the translation units of
// libigloo_cuda.so.  Nothing here crosses the C ABI.
#pragma once

#include <cuda_runtime.h>

#include <atomic>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/igloo_cuda.h"
#include "expr_isa.h"

namespace igl {

// ---- errors ----------------------------------------------------------------
void set_error(const char* fmt, ...);
int32_t fail(int32_t code, const char* fmt, ...);

#define IGL_CUDA_OK(call)                                                                  \
    do {                                                                                   \
        cudaError_t e__ = (call);                                                          \
        if (e__ != cudaSuccess)                                                            \
            return igl::fail(IGL_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), \
                             __FILE__, __LINE__);                                          \
    } while (0)

#define IGL_TRY(call)                   \
    do {                                \
        int32_t r__ = (call);           \
        if (r__ < 0) return r__;        \
    } while (0)

struct ExprNode {
    int op = 0;
    int col = 0;
    int type = 0;        // result type
    bool nullable = false;
    bool lit_null = false;
    uint64_t lit_bits = 0;
    int cast_to = 0;
    int a = -1, b = -1;  // children (indices into nodes)
};

}  // namespace igl

// The opaque handle types of the C ABI.
struct igl_expr {
    std::vector<igl::ExprNode> nodes;  // postorder; root is nodes.back()
    int type = 0;
    bool nullable = false;
    int ncols = 0;
    std::vector<int32_t> col_types;
    std::vector<uint8_t> col_nullable;
    bool is_plain_col() const { return nodes.size() == 1 && nodes[0].op == IGL_OP_COL; }
};

struct igl_ctx {
    int device = 0;
    int sm_count = 148;
    cudaStream_t stream = nullptr;
    igl_config cfg{};
    std::atomic<int64_t> launches{0};
    // device buffer pool: size-class free lists (power-of-two buckets >= 256 B)
    std::mutex mu;
    std::map<size_t, std::vector<void*>> free_lists;
    std::map<void*, size_t> live;      // ptr -> bucket bytes
    size_t pool_bytes = 0;
    std::map<void*, size_t> pinned;
    void* comm = nullptr;              // igl::Comm (shuffle.cu) once igl_comm_init has run
};

namespace igl {
// IGL_TRACE=1: host wall-clock per named section, printed at process exit (api.cu).
void trace_add(const char* name, double seconds);
bool trace_on();
double trace_now();
struct HostTrace {
    const char* name; double t0;
    explicit HostTrace(const char* n) : name(n), t0(trace_on() ? trace_now() : 0.0) {}
    ~HostTrace() { if (trace_on()) trace_add(name, trace_now() - t0); }
};
struct DeviceGuard {
    int prev = -1;
    explicit DeviceGuard(int dev) {
        cudaGetDevice(&prev);
        if (prev != dev) cudaSetDevice(dev);
        else prev = -1;
    }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};
inline cudaStream_t pick_stream(igl_ctx* ctx, void* s) { return s ? (cudaStream_t)s : ctx->stream; }
inline int type_width(int t) { return t == IGL_T_I32 ? 4 : (t == IGL_T_I64 || t == IGL_T_F64) ? 8 : 0; }
int32_t pool_alloc(igl_ctx* ctx, size_t bytes, void** out);
void pool_free(igl_ctx* ctx, void* p);
void comm_release(igl_ctx* ctx);    // shuffle.cu: destroys the NCCL communicator of a context, if any
}  // namespace igl
