// filter_project.cu — K1/K2/K3: fused predicate -> stream compaction ->
// projection over device-resident Arrow columns.
//
// Reference semantics implemented (crates/engine/src/operators/):
//   filter.rs:39-68      mask = predicate(batch); keep rows whose mask is valid
//                        AND true on every column, order preserved
//   projection.rs:48-72  one output column per expression, literals broadcast
// and the planner-side fusion of Projection(Filter(x)) (physical_planner.rs:52-94).
//
// Kernel shape (tile_pipe.cuh: one persistent CTA = compute warps + 1 control warp):
//   * the control warp streams each tile's referenced column slices into a
//     shared-memory ring with 1-D bulk async copies (TMA engine, mbarrier
//     complete_tx) and runs the ordered single-pass scan (decoupled look-back)
//     that turns per-warp survivor counts into global output offsets;
//   * compute warps interpret the fused program (an accumulator machine whose
//     operands are staged columns, literals or per-warp shared-memory
//     temporaries) on 4 rows per lane held in registers — lane i owns rows
//     i, i+32, i+64, i+96 of its warp's 128-row group, so every ballot covers 32
//     consecutive rows and compacted stores are contiguous per instruction —
//     count survivors with __ballot_sync/__popc and write the projected columns
//     straight to their final positions.  The predicate of tile t+1 is evaluated
//     before the outputs of tile t are written, hiding the scan latency.
// HBM traffic = referenced input columns once + surviving output rows once;
// no intermediate array ever reaches HBM.
#include "expr_ir.h"
#include "kernels.h"
#include "fp_device.cuh"
#include "fp_jit.h"

#include <cstring>

namespace igl {

template <int R, int MODE, bool HAS_NULLS>
__global__ void __launch_bounds__(kPipeThreads, kPipeMinCtas) fp_kernel(const __grid_constant__ FpParams P) {
    fp_body<R, MODE, HAS_NULLS>(P);
}

// Between the two passes: popcount of every tile's slice of the survivor bitmap ->
// exclusive offsets (off[ntiles] = surviving rows, also left in the header).  1024 tiles
// per CTA iteration, ordered across CTAs by the block-prefix scan.
constexpr int kToThreads = 256;
constexpr int kToPerThread = 4;
constexpr int kToTile = kToThreads * kToPerThread;

__global__ void __launch_bounds__(kToThreads) fp_tile_offsets_kernel(const uint4* __restrict__ mask, int quads_per_tile,
                                                                     long long ntiles, unsigned long long* __restrict__ off,
                                                                     PipeHeader* hdr, unsigned int* ticket, uint64_t* status) {
    __shared__ unsigned int warp_sums[kToThreads / 32];
    __shared__ unsigned long long s_base;
    __shared__ long long s_tile;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const long long nscan = (ntiles + kToTile - 1) / kToTile;
    while (true) {
        if (threadIdx.x == 0) s_tile = (long long)atomicAdd(ticket, 1u);
        __syncthreads();
        const long long st = s_tile;
        if (st >= nscan) break;
        const long long base_i = st * kToTile + (long long)threadIdx.x * kToPerThread;
        unsigned int x[kToPerThread];
        unsigned int run = 0;
#pragma unroll
        for (int k = 0; k < kToPerThread; ++k) {
            const long long i = base_i + k;
            unsigned int c = 0;
            if (i < ntiles) {
                const uint4* q = mask + i * quads_per_tile;
                for (int j = 0; j < quads_per_tile; ++j) {
                    const uint4 v = q[j];
                    c += __popc(v.x) + __popc(v.y) + __popc(v.z) + __popc(v.w);
                }
            }
            x[k] = run;      // exclusive inside the thread
            run += c;
        }
        unsigned int inc = run;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned int t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) warp_sums[warp] = inc;
        __syncthreads();
        unsigned int wprefix = 0, total = 0;
#pragma unroll
        for (int w = 0; w < kToThreads / 32; ++w) {
            if (w < warp) wprefix += warp_sums[w];
            total += warp_sums[w];
        }
        if (warp == 0) {
            if (lane == 0) scan_publish(status, nscan, st, total);
            __syncwarp();
            const uint64_t b = scan_exclusive(status, nscan, st, total, lane);
            if (lane == 0) {
                s_base = b;
                if (st == nscan - 1) {
                    off[ntiles] = b + total;
                    hdr->out_count = (long long)(b + total);
                    hdr->ticket = 0;      // pass 1 is over; pass 2 claims its tiles from zero again
                }
            }
        }
        __syncthreads();
        const unsigned long long o0 = s_base + wprefix + (inc - run);
#pragma unroll
        for (int k = 0; k < kToPerThread; ++k) {
            const long long i = base_i + k;
            if (i < ntiles) off[i] = o0 + x[k];
        }
        __syncthreads();
    }
}

// byte-per-row {0,1} -> Arrow bitmap; n is read from the device-side row count.
__global__ void pack_bits_kernel(const uint8_t* __restrict__ bytes, uint32_t* __restrict__ bits,
                                 const long long* n_ptr, long long n_fixed) {
    const long long n = n_ptr ? *n_ptr : n_fixed;
    const long long nwords = (n + 31) >> 5;
    const int lane = threadIdx.x & 31;
    for (long long w = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5; w < nwords;
         w += ((long long)gridDim.x * blockDim.x) >> 5) {
        const long long i = (w << 5) + lane;
        const bool b = i < n && bytes[i] != 0;
        const uint32_t word = __ballot_sync(0xffffffffu, b);
        if (lane == 0) bits[w] = word;
    }
}

// status must hold scan_scratch_words(ceil(ntiles / 1024)) zeroed words, *ticket must be zero.
cudaError_t launch_tile_offsets(const void* bitmap, int rows_per_tile, long long ntiles, unsigned long long* off,
                                PipeHeader* hdr, unsigned int* ticket, uint64_t* status, int sm_count, cudaStream_t st) {
    const long long nscan = (ntiles + kToTile - 1) / kToTile;
    const int grid = (int)std::min<long long>(std::max<long long>(nscan, 1), (long long)sm_count * 2);
    fp_tile_offsets_kernel<<<grid, kToThreads, 0, st>>>((const uint4*)bitmap, rows_per_tile / 128, ntiles, off, hdr, ticket, status);
    return cudaGetLastError();
}
long long tile_offsets_scan_tiles(long long ntiles) { return (ntiles + kToTile - 1) / kToTile; }

cudaError_t launch_pack_bits(const uint8_t* bytes, uint32_t* bits, const long long* n_ptr, long long n_fixed,
                             long long cap_rows, int sm_count, cudaStream_t st) {
    long long blocks = (cap_rows + 255) / 256;
    if (blocks < 1) blocks = 1;
    if (blocks > (long long)sm_count * 8) blocks = (long long)sm_count * 8;
    pack_bits_kernel<<<(int)blocks, 256, 0, st>>>(bytes, bits, n_ptr, n_fixed);
    return cudaGetLastError();
}

}  // namespace igl

using namespace igl;

// One compiled launch shape: the lowered program, its staged columns and ring geometry.
struct FpProgram {
    FpParams P{};
    int mode = FP_PROJECT;
    std::vector<int> slot_of_col;            // schema column -> staged slot, -1 = not read
    int rows_per_lane = 4;
    size_t smem_bytes = 0;
    int grid_cap = 0;
    int n_instr = 0, n_exprs = 0, n_cols = 0;
    int64_t rows_seen = 0;                   // rows launched through this program so far
    // run-time specialised kernel (fp_jit.h): 0 not tried, 1 loaded, -1 unavailable
    int jit_state = 0;
    void* jit_kernel = nullptr;
};

struct igl_fp {
    igl_ctx* ctx = nullptr;
    FpProgram one;                           // FP_PROJECT, or FP_SCAN: predicate and outputs in one pass
    FpProgram mask, apply;                   // FP_MASK + FP_APPLY: the two-pass form of the same operator
    bool two_pass_ok = false;
    int two_pass_mode = -1;                  // IGL_FP_TWO_PASS: 0 never, 1 whenever possible, -1 by input size
    bool has_pred = false, has_nulls = false;
    int n_in = 0;
    std::vector<int> in_types;
    std::vector<int> out_types;
    std::vector<uint8_t> out_nullable;
    // scratch (grow-only): header + scan words (+ tile offsets and survivor bitmap); byte planes for bit outputs
    void* scratch = nullptr;
    size_t scratch_bytes = 0;
    std::vector<void*> tmp_bytes;            // per output: boolean byte plane
    std::vector<void*> tmp_valid;            // per output: validity byte plane
    int64_t tmp_rows = 0;
    std::vector<igl_dev_col> last_out;
    int64_t last_nrows = 0;
    PipeHeader* host_hdr = nullptr;          // pinned
    bool launched = false;
};

namespace {

typedef void (*fp_kernel_t)(const FpParams);

template <int R, bool NULLS>
fp_kernel_t pick_mode(int mode) {
    switch (mode) {
        case FP_PROJECT: return fp_kernel<R, FP_PROJECT, NULLS>;
        case FP_SCAN: return fp_kernel<R, FP_SCAN, NULLS>;
        case FP_MASK: return fp_kernel<R, FP_MASK, NULLS>;
        default: return fp_kernel<R, FP_APPLY, NULLS>;
    }
}

fp_kernel_t pick_kernel(int rows_per_lane, int mode, bool nulls) {
    if (rows_per_lane == 4) return nulls ? pick_mode<4, true>(mode) : pick_mode<4, false>(mode);
    return nulls ? pick_mode<1, true>(mode) : pick_mode<1, false>(mode);
}

// Lowers `exprs` (expression 0.. in order; the outputs start at `first_out`) against the
// input schema and lays out the shared-memory ring.  `mask_col` stages one more bit-packed
// column (the survivor bitmap) behind the referenced inputs.  `force_r` = 0 lets the
// geometry pick rows per lane.  Returns IGL_ERR_UNSUPPORTED when the shape does not fit.
int32_t build_program(igl_ctx* ctx, const int32_t* in_types, const uint8_t* in_nullable, int n_in,
                      const std::vector<const igl_expr*>& exprs, size_t first_out, bool with_rowid, bool mask_col,
                      int force_r, bool has_nulls, int mode, FpProgram& prog) {
    prog.mode = mode;
    FpParams& P = prog.P;
    std::vector<bool> used(n_in, false);
    for (auto* e : exprs) referenced_columns(*e, used);
    prog.slot_of_col.assign(n_in, -1);
    uint8_t shifts[kMaxCols + 1] = {};
    int nslots = 0;
    for (int c = 0; c < n_in; ++c) {
        if (!used[c]) continue;
        if (nslots >= kMaxCols) return fail(IGL_ERR_UNSUPPORTED, "more than %d input columns referenced", kMaxCols);
        const int t = in_types[c];
        if (t == IGL_T_UTF8) return fail(IGL_ERR_UNSUPPORTED, "Utf8 columns cannot be read by device expressions");
        shifts[nslots] = t == IGL_T_BOOL ? 0xFF : (t == IGL_T_I32 ? 2 : 3);
        prog.slot_of_col[c] = nslots++;
    }
    (void)in_nullable;
    if (mask_col) {
        if (nslots >= kMaxCols) return fail(IGL_ERR_UNSUPPORTED, "no staged slot left for the survivor bitmap");
        shifts[nslots] = 0xFF;
        P.mask_slot = (uint8_t)nslots++;
    }

    std::vector<Instr> code;
    int n_tmp = 0;
    for (size_t i = 0; i < exprs.size(); ++i) {
        LoweredExpr le;
        IGL_TRY(lower_expr(*exprs[i], code, le));
        P.expr_first[i] = (uint8_t)le.first;
        P.expr_count[i] = (uint8_t)le.count;
        n_tmp = std::max(n_tmp, le.n_tmp);
    }
    const size_t n_out_exprs = exprs.size() - first_out;
    if (with_rowid) {   // extra Int32 output: the input row index of every surviving row
        Instr in{};
        in.op = IOP_ROWID; in.type = IGL_T_I32; in.dst_kind = OPK_ACC;
        P.expr_first[exprs.size()] = (uint8_t)code.size();
        P.expr_count[exprs.size()] = 1;
        code.push_back(in);
    }
    if ((int)code.size() > kMaxInstr) return fail(IGL_ERR_UNSUPPORTED, "fused operator exceeds %d instructions", kMaxInstr);
    for (auto& in : code) {  // schema column index -> staged slot
        if (in.a_kind == OPK_COL) in.a_idx = (uint8_t)prog.slot_of_col[in.a_idx];
        if (in.b_kind == OPK_COL) in.b_idx = (uint8_t)prog.slot_of_col[in.b_idx];
    }
    // fast-handler ids (only used by the no-null kernels; see fast_step)
    for (auto& in : code) {
        in.fast = 0;
        if (in.dst_kind != OPK_ACC) continue;
        int pair = -1;
        if (in.a_kind == OPK_COL && in.b_kind == OPK_LIT) pair = FP_PAIR_COL_LIT;
        else if (in.a_kind == OPK_COL && in.b_kind == OPK_COL) pair = FP_PAIR_COL_COL;
        else if (in.a_kind == OPK_ACC && in.b_kind == OPK_LIT) pair = FP_PAIR_ACC_LIT;
        else if (in.a_kind == OPK_ACC && in.b_kind == OPK_COL) pair = FP_PAIR_ACC_COL;
        const int ty = in.type == IGL_T_I32 ? FP_TY_I32 : in.type == IGL_T_I64 ? FP_TY_I64 : in.type == IGL_T_F64 ? FP_TY_F64 : -1;
        int cls = -1;
        if (in.op >= IGL_OP_EQ && in.op <= IGL_OP_GE) cls = FP_CLS_CMP;
        else if (in.op == IGL_OP_ADD) cls = FP_CLS_ADD;
        else if (in.op == IGL_OP_SUB) cls = FP_CLS_SUB;
        else if (in.op == IGL_OP_MUL) cls = FP_CLS_MUL;
        if (pair >= 0 && ty >= 0 && cls >= 0 && !(in.aux & (AUX_A_NULL | AUX_B_NULL))) in.fast = (uint8_t)(1 + (pair * 4 + cls) * 3 + ty);
        else if (in.op == IOP_MOV && in.a_kind == OPK_COL && in.b_kind == OPK_NONE && !(in.aux & AUX_A_NULL)) {
            const int sh = shifts[in.a_idx];
            if (sh == 2) in.fast = kFastMov32; else if (sh == 3) in.fast = kFastMov64;
        }
    }
    memcpy(P.code, code.data(), code.size() * sizeof(Instr));
    prog.n_instr = (int)code.size();
    prog.n_exprs = (int)exprs.size() + (with_rowid ? 1 : 0);
    prog.n_cols = nslots;
    P.n_out = (uint8_t)(n_out_exprs + (with_rowid ? 1 : 0));
    for (size_t j = 0; j < n_out_exprs; ++j) {
        const int t = exprs[first_out + j]->type;
        P.out_shift[j] = t == IGL_T_BOOL ? 0 : (t == IGL_T_I32 ? 2 : 3);
    }
    if (with_rowid) P.out_shift[n_out_exprs] = 2;

    // shared-memory ring geometry (+ per-warp temporaries behind the ring)
    const uint32_t tmp_bytes = (uint32_t)n_tmp * kPipeComputeWarps * (1024 + 32);
    PipeGeometry geo = pipe_geometry(shifts, nslots, P.T, tmp_bytes, force_r, mode == FP_SCAN ? 1 : 8);
    if (!geo.ok) return fail(IGL_ERR_UNSUPPORTED, "referenced columns need %u bytes per tile; too wide for one pass", geo.stage_bytes);
    prog.rows_per_lane = geo.rows_per_lane;
    P.n_tmp = (uint8_t)n_tmp;
    P.tmp_off = geo.stage_bytes * geo.stages;
    prog.smem_bytes = (size_t)geo.stage_bytes * geo.stages + tmp_bytes;

    fp_kernel_t k = pick_kernel(prog.rows_per_lane, mode, has_nulls);
    IGL_CUDA_OK(cudaFuncSetAttribute((const void*)k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)prog.smem_bytes));
    int per_sm = 0;
    IGL_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, (const void*)k, kPipeThreads, prog.smem_bytes));
    if (per_sm < 1) return fail(IGL_ERR_INTERNAL, "filter/project kernel does not fit on an SM (smem %zu)", prog.smem_bytes);
    prog.grid_cap = per_sm * ctx->sm_count;
    return IGL_OK;
}

// Points a program's staged slots at this call's input columns.
int32_t bind_inputs(const igl_fp* fp, const FpProgram& prog, const igl_dev_col* in, FpParams& P) {
    for (int c = 0; c < fp->n_in; ++c) {
        const int s = prog.slot_of_col[c];
        if (s < 0) continue;
        if (in[c].type != fp->in_types[c]) return fail(IGL_ERR_INVALID_ARG, "input column %d has type %d, plan expects %d", c, in[c].type, fp->in_types[c]);
        if (((uintptr_t)in[c].data & 15) || ((uintptr_t)in[c].validity & 15))
            return fail(IGL_ERR_INVALID_ARG, "input column %d is not 16-byte aligned", c);
        P.T.data[s] = (const uint8_t*)in[c].data;
        P.T.valid[s] = fp->has_nulls ? in[c].validity : nullptr;
    }
    return IGL_OK;
}

FpJitSpec jit_spec(const FpProgram& prog, bool has_nulls) {
    FpJitSpec sp;
    const FpParams& P = prog.P;
    sp.code = P.code; sp.n_instr = prog.n_instr;
    sp.expr_first = P.expr_first; sp.expr_count = P.expr_count; sp.n_exprs = prog.n_exprs;
    sp.n_out = P.n_out; sp.out_shift = P.out_shift;
    sp.smem_off = P.T.smem_off; sp.valid_off = P.T.valid_off; sp.shift = P.T.shift; sp.n_cols = prog.n_cols;
    sp.rows_per_lane = prog.rows_per_lane; sp.mode = prog.mode; sp.has_nulls = has_nulls;
    return sp;
}

// IGL_FP_JIT: 0 never specialise, 1 always; unset: once the plan has been given at least
// IGL_FP_JIT_MIN_ROWS rows (default 64 Mi) over all of its launches — one large launch or a
// stream of 4 Mi-row chunks alike — so the compile time is repaid.
bool want_jit(int64_t nrows) {
    const char* e = getenv("IGL_FP_JIT");
    if (e) return atoi(e) != 0;
    const char* m = getenv("IGL_FP_JIT_MIN_ROWS");
    return nrows >= (m ? atoll(m) : (long long)1 << 26);
}

int32_t launch_program(igl_ctx* ctx, FpProgram& prog, const FpParams& P, bool has_nulls, cudaStream_t st, bool may_compile = true) {
    const int grid = (int)std::min<int64_t>(P.T.ntiles, prog.grid_cap);
    prog.rows_seen += P.T.nrows;
    if (may_compile && prog.jit_state == 0 && want_jit(prog.rows_seen)) {
        void* k = nullptr;
        if (fp_jit_build(jit_spec(prog, has_nulls), &k) >= 0 &&
            cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)prog.smem_bytes) == cudaSuccess) {
            prog.jit_kernel = k;
            prog.jit_state = 1;
        } else {
            cudaGetLastError();
            prog.jit_state = -1;   // keep the interpreter; the reason stays in igl_last_error
        }
    }
    if (prog.jit_state == 1) {
        void* args[] = {(void*)&P};
        IGL_CUDA_OK(cudaLaunchKernel(prog.jit_kernel, dim3(grid), dim3(kPipeThreads), args, prog.smem_bytes, st));
    } else {
        fp_kernel_t k = pick_kernel(prog.rows_per_lane, prog.mode, has_nulls);
        k<<<grid, kPipeThreads, prog.smem_bytes, st>>>(P);
        IGL_CUDA_OK(cudaGetLastError());
    }
    ctx->launches++;
    return IGL_OK;
}

inline size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

}  // namespace

extern "C" {

int32_t igl_fp_create(igl_ctx* ctx, const int32_t* in_types, const uint8_t* in_nullable, int32_t n_in,
                      const igl_expr* pred, const igl_expr* const* proj, int32_t nproj, igl_fp** out) {
    return igl::fp_create_ex(ctx, in_types, in_nullable, n_in, pred, proj, nproj, nullptr, false, out);
}

}  // extern "C"

// `skip` (optional, n_in flags): input columns a pass-through FilterExec must not
// emit on the device (Utf8 columns, gathered afterwards through the row ids).
int32_t igl::fp_create_ex(igl_ctx* ctx, const int32_t* in_types, const uint8_t* in_nullable, int32_t n_in,
                          const igl_expr* pred, const igl_expr* const* proj, int32_t nproj, const uint8_t* skip,
                          bool with_rowid, igl_fp** out) {
    if (!ctx || !out || n_in <= 0 || !in_types) return fail(IGL_ERR_INVALID_ARG, "igl_fp_create: bad arguments");
    *out = nullptr;
    if (nproj < 0 || (nproj > 0 && !proj)) return fail(IGL_ERR_INVALID_ARG, "igl_fp_create: bad projection list");
    if (pred && pred->type != IGL_T_BOOL)
        return fail(IGL_ERR_PREDICATE_NOT_BOOLEAN, "Filter predicate must return boolean array");  // filter.rs:50-55
    auto fp = std::make_unique<igl_fp>();
    fp->ctx = ctx;
    fp->n_in = n_in;
    fp->in_types.assign(in_types, in_types + n_in);
    fp->has_pred = pred != nullptr;

    // FilterExec: nproj == 0 passes every input column through.
    std::vector<std::unique_ptr<igl_expr>> owned;
    std::vector<const igl_expr*> outs;
    if (nproj == 0) {
        for (int c = 0; c < n_in; ++c) {
            if (skip && skip[c]) continue;
            uint8_t bc[3] = {IGL_OP_COL, (uint8_t)(c & 0xFF), (uint8_t)(c >> 8)};
            igl_expr* e = nullptr;
            IGL_TRY(igl_expr_compile(ctx, bc, 3, in_types, in_nullable, n_in, &e));
            owned.emplace_back(e);
            outs.push_back(e);
        }
    } else {
        for (int j = 0; j < nproj; ++j) {
            if (!proj[j]) return fail(IGL_ERR_INVALID_ARG, "igl_fp_create: projection %d is NULL", j);
            outs.push_back(proj[j]);
        }
    }
    if ((int)outs.size() > kMaxOut - 1) return fail(IGL_ERR_UNSUPPORTED, "more than %d output columns", kMaxOut - 1);
    for (auto* e : outs) {
        if (e->ncols != n_in) return fail(IGL_ERR_INVALID_ARG, "expression compiled against a different schema");
        if (e->type == IGL_T_UTF8) return fail(IGL_ERR_UNSUPPORTED, "Utf8 output columns go through the operator API (gather by row index)");
    }
    if (pred && pred->ncols != n_in) return fail(IGL_ERR_INVALID_ARG, "predicate compiled against a different schema");

    // nulls anywhere (a nullable referenced column or a nullable result) select the null-aware kernels
    {
        std::vector<bool> used(n_in, false);
        if (pred) referenced_columns(*pred, used);
        for (auto* e : outs) referenced_columns(*e, used);
        for (int c = 0; c < n_in; ++c) if (used[c] && in_nullable && in_nullable[c]) fp->has_nulls = true;
        if (pred && pred->nullable) fp->has_nulls = true;
        for (auto* e : outs) if (e->nullable) fp->has_nulls = true;
    }
    for (auto* e : outs) {
        fp->out_types.push_back(e->type);
        fp->out_nullable.push_back(e->nullable ? 1 : 0);
    }
    if (with_rowid) {
        fp->out_types.push_back(IGL_T_I32);
        fp->out_nullable.push_back(0);
    }

    DeviceGuard g(ctx->device);
    std::vector<const igl_expr*> all;
    if (pred) all.push_back(pred);
    for (auto* e : outs) all.push_back(e);
    IGL_TRY(build_program(ctx, in_types, in_nullable, n_in, all, pred ? 1 : 0, with_rowid, false, 0, fp->has_nulls,
                          pred ? FP_SCAN : FP_PROJECT, fp->one));
    if (pred) {
        // Two-pass form: the predicate alone writes a survivor bitmap, a small kernel turns the
        // bitmap into tile offsets, the outputs are then written with no ordering between CTAs.
        // Both passes must cut the input into the same tiles.
        if (const char* e = getenv("IGL_FP_TWO_PASS")) fp->two_pass_mode = atoi(e) ? 1 : 0;
        std::vector<const igl_expr*> pe{pred};
        int32_t r1 = build_program(ctx, in_types, in_nullable, n_in, pe, 1, false, false, 0, fp->has_nulls, FP_MASK, fp->mask);
        int32_t r2 = r1 < 0 ? r1 : build_program(ctx, in_types, in_nullable, n_in, outs, 0, with_rowid, true, 0, fp->has_nulls, FP_APPLY, fp->apply);
        if (r1 >= 0 && r2 >= 0 && fp->mask.rows_per_lane != fp->apply.rows_per_lane) {
            FpProgram& wide = fp->mask.rows_per_lane == 4 ? fp->mask : fp->apply;
            wide = FpProgram{};
            const bool is_mask = &wide == &fp->mask;
            r1 = is_mask ? build_program(ctx, in_types, in_nullable, n_in, pe, 1, false, false, 1, fp->has_nulls, FP_MASK, fp->mask)
                         : build_program(ctx, in_types, in_nullable, n_in, outs, 0, with_rowid, true, 1, fp->has_nulls, FP_APPLY, fp->apply);
        }
        fp->two_pass_ok = r1 >= 0 && r2 >= 0 && fp->mask.P.T.sub_rows == fp->apply.P.T.sub_rows;
    }
    IGL_CUDA_OK(cudaHostAlloc((void**)&fp->host_hdr, sizeof(PipeHeader), cudaHostAllocDefault));
    fp->tmp_bytes.assign(outs.size(), nullptr);
    fp->tmp_valid.assign(outs.size(), nullptr);
    *out = fp.release();
    return IGL_OK;
}

extern "C" {

int32_t igl_fp_num_outputs(const igl_fp* fp) { return fp ? (int32_t)fp->out_types.size() : 0; }

int32_t igl_fp_output_type(const igl_fp* fp, int32_t j, int32_t* type, int32_t* nullable) {
    if (!fp || j < 0 || j >= (int)fp->out_types.size()) return fail(IGL_ERR_INVALID_ARG, "igl_fp_output_type: bad index");
    if (type) *type = fp->out_types[j];
    if (nullable) *nullable = fp->out_nullable[j];
    return IGL_OK;
}

int32_t igl_fp_launch(igl_fp* fp, const igl_dev_col* in, int64_t nrows, igl_dev_col* out, void* stream) {
    if (!fp || !in || !out || nrows < 0) return fail(IGL_ERR_INVALID_ARG, "igl_fp_launch: bad arguments");
    igl_ctx* ctx = fp->ctx;
    DeviceGuard g(ctx->device);
    cudaStream_t st = pick_stream(ctx, stream);
    const int tile_rows = fp->one.P.T.tile_rows;
    const long long ntiles1 = (nrows + tile_rows - 1) / tile_rows;
    // One pass keeps every CTA in step with the slowest tile ahead of it; two passes stream at
    // full speed but launch three kernels.  Small inputs stay on one pass.
    bool two = fp->has_pred && fp->two_pass_ok &&
               (fp->two_pass_mode == 1 || (fp->two_pass_mode < 0 && ntiles1 >= 4ll * fp->one.grid_cap));
    FpProgram& body = two ? fp->apply : fp->one;
    FpParams P = body.P;
    P.T.nrows = nrows;
    P.T.ntiles = (nrows + P.T.tile_rows - 1) / P.T.tile_rows;
    IGL_TRY(bind_inputs(fp, body, in, P));
    FpParams PM{};
    if (two) {
        PM = fp->mask.P;
        PM.T.nrows = nrows;
        PM.T.ntiles = (nrows + PM.T.tile_rows - 1) / PM.T.tile_rows;
        IGL_TRY(bind_inputs(fp, fp->mask, in, PM));
    }
    // scratch: [header 256][offsets-kernel header 256][scan words][tile offsets][survivor bitmap]
    // (the bitmap and the offsets are kept per sub-tile: the two passes may stage different numbers of them)
    const long long nsub = (nrows + P.T.sub_rows - 1) / P.T.sub_rows;
    const long long nscan = two ? (nsub + kToTile - 1) / kToTile : P.T.ntiles;
    const size_t scan_bytes = align256((size_t)scan_scratch_words(nscan) * 8);
    const size_t off_bytes = two ? align256((size_t)(nsub + 1) * 8 + 128) : 0;   // + slack: a stage copies up to 9 entries
    const size_t mask_bytes = two ? align256((size_t)nsub * (P.T.sub_rows / 8) + 16) : 0;
    const size_t need = 512 + scan_bytes + off_bytes + mask_bytes;
    if (need > fp->scratch_bytes) {
        if (fp->scratch) pool_free(ctx, fp->scratch);
        fp->scratch = nullptr; fp->scratch_bytes = 0;
        IGL_TRY(pool_alloc(ctx, need, &fp->scratch));
        fp->scratch_bytes = need;
    }
    uint8_t* sb = (uint8_t*)fp->scratch;
    P.T.hdr = (PipeHeader*)sb;
    P.T.status = (uint64_t*)(sb + 512);
    IGL_CUDA_OK(cudaMemsetAsync(fp->scratch, 0, fp->has_pred ? 512 + scan_bytes : 256, st));
    // outputs (+ byte planes for bit-packed results)
    const size_t nout = fp->out_types.size();
    bool any_tmp = false;
    for (size_t j = 0; j < nout; ++j) {
        const bool want_valid = fp->has_nulls && fp->out_nullable[j];
        if (!out[j].data) return fail(IGL_ERR_INVALID_ARG, "output column %zu has no data buffer", j);
        if (want_valid && !out[j].validity) return fail(IGL_ERR_INVALID_ARG, "output column %zu is nullable but has no validity buffer", j);
        if (fp->out_types[j] == IGL_T_BOOL || want_valid) any_tmp = true;
    }
    if (any_tmp && nrows > fp->tmp_rows) {
        for (size_t j = 0; j < nout; ++j) {
            if (fp->tmp_bytes[j]) { pool_free(ctx, fp->tmp_bytes[j]); fp->tmp_bytes[j] = nullptr; }
            if (fp->tmp_valid[j]) { pool_free(ctx, fp->tmp_valid[j]); fp->tmp_valid[j] = nullptr; }
        }
        fp->tmp_rows = 0;
    }
    for (size_t j = 0; j < nout; ++j) {
        const bool want_valid = fp->has_nulls && fp->out_nullable[j];
        if (fp->out_types[j] == IGL_T_BOOL) {
            if (!fp->tmp_bytes[j]) IGL_TRY(pool_alloc(ctx, (size_t)std::max<int64_t>(nrows, 1), &fp->tmp_bytes[j]));
            P.out_data[j] = fp->tmp_bytes[j];
        } else {
            P.out_data[j] = out[j].data;
        }
        if (want_valid) {
            if (!fp->tmp_valid[j]) IGL_TRY(pool_alloc(ctx, (size_t)std::max<int64_t>(nrows, 1), &fp->tmp_valid[j]));
            P.out_valid[j] = (uint8_t*)fp->tmp_valid[j];
        } else {
            P.out_valid[j] = nullptr;
        }
    }
    if (any_tmp) fp->tmp_rows = std::max(fp->tmp_rows, nrows);
    fp->last_out.assign(out, out + nout);
    fp->last_nrows = nrows;
    fp->launched = true;
    if (nrows == 0) return IGL_OK;

    if (two) {
        PipeHeader* hdr2 = (PipeHeader*)(sb + 256);
        unsigned long long* tile_off = (unsigned long long*)(sb + 512 + scan_bytes);
        uint8_t* mask = sb + 512 + scan_bytes + off_bytes;
        PM.T.hdr = P.T.hdr;
        PM.T.status = nullptr;
        PM.mask_words = (uint32_t*)mask;
        IGL_TRY(launch_program(ctx, fp->mask, PM, fp->has_nulls, st));
        IGL_CUDA_OK(launch_tile_offsets(mask, P.T.sub_rows, nsub, tile_off, P.T.hdr, &hdr2->ticket, P.T.status, ctx->sm_count, st));
        ctx->launches++;
        P.T.aux_table = tile_off;
        P.T.data[P.mask_slot] = mask;
        P.T.valid[P.mask_slot] = nullptr;
        P.T.status = nullptr;
    }
    IGL_TRY(launch_program(ctx, body, P, fp->has_nulls, st));
    const long long* n_ptr = fp->has_pred ? &P.T.hdr->out_count : nullptr;
    for (size_t j = 0; j < nout; ++j) {
        const bool want_valid = fp->has_nulls && fp->out_nullable[j];
        if (fp->out_types[j] == IGL_T_BOOL) {
            IGL_CUDA_OK(launch_pack_bits((const uint8_t*)fp->tmp_bytes[j], (uint32_t*)out[j].data, n_ptr, nrows, nrows, ctx->sm_count, st));
            ctx->launches++;
        }
        if (want_valid) {
            IGL_CUDA_OK(launch_pack_bits((const uint8_t*)fp->tmp_valid[j], (uint32_t*)out[j].validity, n_ptr, nrows, nrows, ctx->sm_count, st));
            ctx->launches++;
        }
    }
    return IGL_OK;
}

// 1 when the last launch of `fp` ran on a run-time specialised kernel (all of its passes), else 0.
int32_t igl_fp_is_specialised(const igl_fp* fp) {
    if (!fp) return 0;
    if (fp->one.jit_state == 1) return 1;
    return fp->mask.jit_state == 1 && fp->apply.jit_state == 1 ? 1 : 0;
}

// Compiles every pass of `fp` with NVRTC without loading anything (no GPU needed beyond the
// one the plan was created on).  *cubin_bytes receives the total image size.
int32_t igl_fp_jit_compile_only(igl_fp* fp, int64_t* cubin_bytes) {
    if (!fp) return fail(IGL_ERR_INVALID_ARG, "igl_fp_jit_compile_only: NULL plan");
    int64_t total = 0;
    std::vector<const FpProgram*> progs{&fp->one};
    if (fp->two_pass_ok) { progs.push_back(&fp->mask); progs.push_back(&fp->apply); }
    for (const FpProgram* p : progs) {
        std::string cubin, log;
        IGL_TRY(fp_jit_compile(jit_spec(*p, fp->has_nulls), &cubin, &log));
        total += (int64_t)cubin.size();
    }
    if (cubin_bytes) *cubin_bytes = total;
    return IGL_OK;
}

// Ahead-of-time specialisation: compiles and loads the plan's kernels now (while the query is being
// planned, say), so that no later launch pays the ~1 s compile.  IGL_OK with igl_fp_is_specialised
// still 0 means NVRTC is not available: the interpreter stays in use, nothing else changes.
int32_t igl_fp_prepare(igl_fp* fp) {
    if (!fp) return fail(IGL_ERR_INVALID_ARG, "igl_fp_prepare: NULL plan");
    DeviceGuard g(fp->ctx->device);
    std::vector<FpProgram*> progs{&fp->one};
    if (fp->two_pass_ok) { progs.push_back(&fp->mask); progs.push_back(&fp->apply); }
    for (FpProgram* p : progs) {
        if (p->jit_state != 0) continue;
        void* k = nullptr;
        if (fp_jit_build(jit_spec(*p, fp->has_nulls), &k) >= 0 &&
            cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p->smem_bytes) == cudaSuccess) {
            p->jit_kernel = k;
            p->jit_state = 1;
        } else {
            cudaGetLastError();
            p->jit_state = -1;
        }
    }
    return IGL_OK;
}

int32_t igl_fp_finish(igl_fp* fp, int64_t* out_rows, void* stream) {
    if (!fp || !fp->launched) return fail(IGL_ERR_INVALID_ARG, "igl_fp_finish: nothing launched");
    igl_ctx* ctx = fp->ctx;
    DeviceGuard g(ctx->device);
    cudaStream_t st = pick_stream(ctx, stream);
    fp->launched = false;
    if (fp->last_nrows == 0) {
        IGL_CUDA_OK(cudaStreamSynchronize(st));
        if (out_rows) *out_rows = 0;
        return IGL_OK;
    }
    IGL_CUDA_OK(cudaMemcpyAsync(fp->host_hdr, fp->scratch, sizeof(PipeHeader), cudaMemcpyDeviceToHost, st));
    IGL_CUDA_OK(cudaStreamSynchronize(st));
    const unsigned int ef = fp->host_hdr->err_flags;
    if (ef & ERRF_DIV_ZERO) return fail(IGL_ERR_DIV_BY_ZERO, "Divide by zero error");
    if (ef & ERRF_OVERFLOW) return fail(IGL_ERR_ARITH_OVERFLOW, "Arithmetic overflow: Overflow happened on division");
    if (ef & ERRF_CAST) return fail(IGL_ERR_CAST, "Cast error: Can't cast value to the target integer type");
    if (out_rows) *out_rows = fp->has_pred ? (int64_t)fp->host_hdr->out_count : fp->last_nrows;
    return IGL_OK;
}

int32_t igl_fp_run(igl_fp* fp, const igl_dev_col* in, int64_t nrows, igl_dev_col* out, int64_t* out_rows, void* stream) {
    IGL_TRY(igl_fp_launch(fp, in, nrows, out, stream));
    return igl_fp_finish(fp, out_rows, stream);
}

void igl_fp_destroy(igl_fp* fp) {
    if (!fp) return;
    DeviceGuard g(fp->ctx->device);
    if (fp->scratch) pool_free(fp->ctx, fp->scratch);
    for (void* p : fp->tmp_bytes) if (p) pool_free(fp->ctx, p);
    for (void* p : fp->tmp_valid) if (p) pool_free(fp->ctx, p);
    if (fp->host_hdr) cudaFreeHost(fp->host_hdr);
    delete fp;
}

}  // extern "C"

// ---------------------------------------------------------------------------
// igl_fp_run_static: the same fused Filter/Projection launch as igl_fp_run, for
// callers that are not allowed to allocate device memory while they run.
//
// Who that is: an engine that carves its whole HBM budget up front and treats
// any later cudaMalloc as a bug (a worker sharing its GPU with other tenants, a
// latency-sensitive stream where an allocation's implicit device synchronisation
// would stall unrelated work), or code running at a point where the CUDA
// allocator must not be entered at all.  igl_fp_run grows a per-plan scratch
// block from the context's pool on demand; this entry point never touches the
// pool, cudaMalloc or cudaFree.  The caller still provides inputs and outputs.
//
// The scan scratch (256-byte header + one status word per tile) lives in the
// file-scope device array below instead: kFpStaticScratchBytes (88 KiB) of
// device memory held by every GPU that loads the library, until the process
// exits.  That bounds a call to kFpStaticMaxRows rows.
//
// Rules that follow from sharing one array:
//   * every call clears the header and the status words it is about to use, so
//     a call never sees what the previous one left behind;
//   * two calls whose kernels overlap on the same GPU (different streams or
//     threads) race on the array — serialise them, or use igl_fp_run;
//   * the first call on each device looks the array's device address up
//     (cudaGetSymbolAddress, with that device current) and caches it; make that
//     first call before starting any stream capture.
// Plans with boolean or nullable outputs need per-row byte planes that only the
// pool can provide and are rejected here.
// ---------------------------------------------------------------------------
namespace igl {
constexpr long long kFpStaticMaxRows = 4ll << 20;   // == the default coalesce_rows chunk
constexpr size_t kFpStaticScratchBytes = 88u << 10;  // 256 + 8 * ceil(kFpStaticMaxRows / 384 rows per narrowest tile)
static_assert(256 + 8 * ((kFpStaticMaxRows + 383) / 384) <= (long long)kFpStaticScratchBytes, "static scratch too small");
static __device__ __align__(256) unsigned char g_fp_static_scratch[kFpStaticScratchBytes];

// In host code `g_fp_static_scratch` is only a handle for the runtime, never a
// device pointer: the address a kernel may use comes from cudaGetSymbolAddress.
static std::mutex g_fp_static_mu;
static void* g_fp_static_ptr[64] = {};

static int32_t fp_static_scratch_ptr(int device, void** out) {
    if (device < 0 || device >= 64) return fail(IGL_ERR_UNSUPPORTED, "igl_fp_run_static: device ordinal %d not supported", device);
    std::lock_guard<std::mutex> g(g_fp_static_mu);
    if (!g_fp_static_ptr[device]) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_fp_static_scratch);   // the caller made `device` current
        if (e != cudaSuccess || !p)
            return fail(IGL_ERR_CUDA, "igl_fp_run_static: cudaGetSymbolAddress failed on device %d: %s", device, cudaGetErrorString(e));
        g_fp_static_ptr[device] = p;
    }
    *out = g_fp_static_ptr[device];
    return IGL_OK;
}
}  // namespace igl

extern "C" int32_t igl_fp_run_static(igl_fp* fp, const igl_dev_col* in, int64_t nrows, igl_dev_col* out,
                                     int64_t* out_rows, void* stream) {
    if (!fp || !in || !out || nrows < 0) return fail(IGL_ERR_INVALID_ARG, "igl_fp_run_static: bad arguments");
    if (nrows > kFpStaticMaxRows)
        return fail(IGL_ERR_INVALID_ARG, "igl_fp_run_static: %lld rows exceed the static scratch limit of %lld rows per call; "
                    "split the batch or use igl_fp_run", (long long)nrows, kFpStaticMaxRows);
    const size_t nout = fp->out_types.size();
    for (size_t j = 0; j < nout; ++j)
        if (fp->out_types[j] == IGL_T_BOOL || (fp->has_nulls && fp->out_nullable[j]))
            return fail(IGL_ERR_UNSUPPORTED, "igl_fp_run_static: output %zu is boolean or nullable and needs pooled byte planes; use igl_fp_run", j);
    igl_ctx* ctx = fp->ctx;
    DeviceGuard g(ctx->device);
    cudaStream_t st = pick_stream(ctx, stream);
    void* scratch = nullptr;
    IGL_TRY(fp_static_scratch_ptr(ctx->device, &scratch));
    FpParams P = fp->one.P;   // always the one-pass form: its scratch is bounded by the row limit alone
    P.T.nrows = nrows;
    P.T.ntiles = (nrows + P.T.tile_rows - 1) / P.T.tile_rows;
    IGL_TRY(bind_inputs(fp, fp->one, in, P));
    for (size_t j = 0; j < nout; ++j) {
        if (!out[j].data) return fail(IGL_ERR_INVALID_ARG, "output column %zu has no data buffer", j);
        P.out_data[j] = out[j].data;
        P.out_valid[j] = nullptr;
    }
    if (nrows == 0) { if (out_rows) *out_rows = 0; return IGL_OK; }
    const size_t need = 256 + (size_t)scan_scratch_words(P.T.ntiles) * 8;
    // the row limit bounds this only while no tile is narrower than 384 rows (tile_pipe.cuh): check, never assume
    if (need > kFpStaticScratchBytes)
        return fail(IGL_ERR_INVALID_ARG, "igl_fp_run_static: this plan needs %zu bytes of scan scratch for %lld rows, the static array holds %zu; "
                    "split the batch or use igl_fp_run", need, (long long)nrows, kFpStaticScratchBytes);
    P.T.hdr = (PipeHeader*)scratch;
    P.T.status = (uint64_t*)((uint8_t*)scratch + 256);
    IGL_CUDA_OK(cudaMemsetAsync(scratch, 0, need, st));   // ticket, error flags, count and every status word start at zero
    IGL_TRY(launch_program(ctx, fp->one, P, fp->has_nulls, st, /*may_compile=*/false));   // compiling and loading a kernel allocates
    IGL_CUDA_OK(cudaMemcpyAsync(fp->host_hdr, scratch, sizeof(PipeHeader), cudaMemcpyDeviceToHost, st));
    IGL_CUDA_OK(cudaStreamSynchronize(st));
    const unsigned int ef = fp->host_hdr->err_flags;
    if (ef & ERRF_DIV_ZERO) return fail(IGL_ERR_DIV_BY_ZERO, "Divide by zero error");
    if (ef & ERRF_OVERFLOW) return fail(IGL_ERR_ARITH_OVERFLOW, "Arithmetic overflow: Overflow happened on division");
    if (ef & ERRF_CAST) return fail(IGL_ERR_CAST, "Cast error: Can't cast value to the target integer type");
    if (out_rows) *out_rows = fp->has_pred ? (int64_t)fp->host_hdr->out_count : nrows;
    return IGL_OK;
}

