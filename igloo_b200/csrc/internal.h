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

// ---- expression IR -----------------------------------------------------------
// Accumulator machine with operand modes.  A leaf (column or literal) is an
// operand of its consumer, never an instruction of its own, so `a < 250` is one
// instruction and `x * 1.5 + 0.25` is two.  An inner result lives in the
// accumulator (registers); when the right subtree is itself an expression the
// left result is parked in a per-warp shared-memory temporary instead.
enum : uint8_t { OPK_NONE = 0, OPK_ACC = 1, OPK_COL = 2, OPK_LIT = 3, OPK_TMP = 4 };
enum : uint8_t { IOP_MOV = 0x50, IOP_ROWID = 0x51 };  // besides the igl_opcode values

constexpr int kMaxTmp = 4;      // shared-memory temporaries per warp
constexpr int kMaxInstr = 48;   // per fused operator (predicate + all outputs)
constexpr int kMaxExprs = 18;   // predicate + 16 outputs + row id
constexpr int kMaxCols = 16;    // staged input columns per operator

struct Instr {            // 32 bytes
    uint8_t op;           // igl_opcode, IOP_MOV or IOP_ROWID
    uint8_t type;         // operand type (igl_type) the operation is carried out in
    uint8_t a_kind, b_kind;
    uint8_t a_idx, b_idx;     // COL: staged slot; TMP: temporary index
    uint8_t dst_kind, dst_idx;  // OPK_ACC or OPK_TMP k
    uint8_t aux;          // CAST: target type; else AUX_*_NULL flags of literal operands
    uint8_t fast;         // 0 = generic step; else 1 + ((pair * 4 + class) * 3 + type): a fully specialised handler
    uint8_t pad[6];
    uint64_t a_lit, b_lit;  // literal bits when *_kind == OPK_LIT
};
static_assert(sizeof(Instr) == 32, "Instr layout");

// Fast handlers (filter_project.cu): operand pair x operator class x type, results in the accumulator.
enum : uint8_t { FP_PAIR_COL_LIT = 0, FP_PAIR_COL_COL = 1, FP_PAIR_ACC_LIT = 2, FP_PAIR_ACC_COL = 3 };
enum : uint8_t { FP_CLS_CMP = 0, FP_CLS_ADD = 1, FP_CLS_SUB = 2, FP_CLS_MUL = 3 };
enum : uint8_t { FP_TY_I32 = 0, FP_TY_I64 = 1, FP_TY_F64 = 2 };
constexpr uint8_t kFastMov32 = 49, kFastMov64 = 50;   // bare column pass-through, by width

// literal flags carried in `aux` for non-CAST ops
enum : uint8_t { AUX_A_NULL = 1, AUX_B_NULL = 2 };

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
};

namespace igl {
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
}  // namespace igl
