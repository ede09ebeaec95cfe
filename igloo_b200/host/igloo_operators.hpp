// igloo_operators.hpp — header-only C++ mirror of the reference's operator API
// over the C ABI of libigloo_cuda.so (include/igloo_cuda.h).
//
// The reference is compiled code (Rust); with no Rust toolchain in this image the
// host side above the C ABI is C++.  Names, argument meaning and error behaviour
// follow crates/engine/src:
//   trait ExecutionPlan { execute() -> stream of RecordBatch; schema() }   physical_plan.rs:10-17
//   FilterExec::new(input, predicate)                                       operators/filter.rs:19
//   ProjectionExec::new(input, expressions, schema)                         operators/projection.rs:19-29
//   HashJoinExec::new(left, right, left_keys, right_keys, join_type)        operators/hash_join.rs:35-76
// A "RecordBatch" is an Arrow C Data Interface struct array (ArrowArray) whose
// release callback the consumer calls; a stream is a pull function returning
// false at end of stream.  Errors surface as igloo::Error (DataFusionError).
#pragma once

#include <cstring>
#include <functional>
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "../../include/igloo_cuda.h"

namespace igloo {

struct Error : std::runtime_error {
    int32_t code;
    Error(int32_t c, const std::string& m) : std::runtime_error(m), code(c) {}
};
inline int32_t check(int32_t rc) {
    if (rc < 0) throw Error(rc, igl_last_error());
    return rc;
}

// Owning wrapper of one exported batch.
struct RecordBatch {
    ArrowArray array{};
    RecordBatch() = default;
    RecordBatch(const RecordBatch&) = delete;
    RecordBatch(RecordBatch&& o) noexcept { array = o.array; o.array.release = nullptr; }
    RecordBatch& operator=(RecordBatch&& o) noexcept {
        if (this != &o) { reset(); array = o.array; o.array.release = nullptr; }
        return *this;
    }
    ~RecordBatch() { reset(); }
    void reset() { if (array.release) array.release(&array); array.release = nullptr; }
    int64_t num_rows() const { return array.length; }
};

// `BoxStream<RecordBatch>`: next(out) returns false at end of stream.
using BatchStream = std::function<bool(RecordBatch&)>;

// Postfix expression bytes (include/igloo_expr.h) — what the Rust shim produces
// from an Arc<dyn PhysicalExpr> (physical_planner.rs:60-64).
using ExprBytes = std::vector<uint8_t>;
inline ExprBytes col(uint16_t i) { return {IGL_OP_COL, (uint8_t)(i & 0xFF), (uint8_t)(i >> 8)}; }
inline ExprBytes lit_i32(int32_t v) { ExprBytes b{IGL_OP_LIT_I32, 0, 0, 0, 0}; memcpy(b.data() + 1, &v, 4); return b; }
inline ExprBytes lit_i64(int64_t v) { ExprBytes b(9); b[0] = IGL_OP_LIT_I64; memcpy(b.data() + 1, &v, 8); return b; }
inline ExprBytes lit_f64(double v) { ExprBytes b(9); b[0] = IGL_OP_LIT_F64; memcpy(b.data() + 1, &v, 8); return b; }
inline ExprBytes binary(ExprBytes a, const ExprBytes& b, uint8_t op) { a.insert(a.end(), b.begin(), b.end()); a.push_back(op); return a; }

class Context {
  public:
    explicit Context(int device = 0, int64_t coalesce_rows = 0) {
        igl_config cfg{}; cfg.device = device; cfg.coalesce_rows = coalesce_rows;
        check(igl_ctx_create(&cfg, &ctx_));
    }
    ~Context() { igl_ctx_destroy(ctx_); }
    Context(const Context&) = delete;
    igl_ctx* get() const { return ctx_; }
  private:
    igl_ctx* ctx_ = nullptr;
};

class ExecutionPlan {
  public:
    virtual ~ExecutionPlan() = default;
    virtual BatchStream execute() = 0;
    // Fills `out`; the caller releases it.
    virtual void schema(ArrowSchema* out) = 0;
};

// A scan that replays batches the host already decoded (the upstream neighbour of
// the hot path: ParquetScanExec emits 1024-row batches, parquet_scan.rs:54).
class BatchSourceExec : public ExecutionPlan {
  public:
    using Producer = std::function<bool(ArrowArray*)>;          // exports the next batch, false at the end
    using SchemaFn = std::function<void(ArrowSchema*)>;
    BatchSourceExec(std::function<Producer()> open, SchemaFn schema) : open_(std::move(open)), schema_(std::move(schema)) {}
    BatchStream execute() override {
        auto prod = std::make_shared<Producer>(open_());
        return [prod](RecordBatch& b) { b.reset(); return (*prod)(&b.array); };
    }
    void schema(ArrowSchema* out) override { schema_(out); }
  private:
    std::function<Producer()> open_;
    SchemaFn schema_;
};

namespace detail {
struct OpHandle {
    igl_op* op = nullptr;
    ~OpHandle() { igl_op_destroy(op); }
};
// Feeds `input` into `op` and yields what comes out: push until IGL_AGAIN turns
// into batches, finish the input, drain to IGL_EOS.
inline BatchStream pump(std::shared_ptr<OpHandle> h, BatchStream input) {
    auto done = std::make_shared<bool>(false);
    return [h, input, done](RecordBatch& out) mutable {
        while (true) {
            out.reset();
            const int32_t rc = check(igl_op_pull(h->op, &out.array));
            if (rc == IGL_OK) return true;
            if (rc == IGL_EOS) return false;
            RecordBatch in;                                       // IGL_AGAIN: feed more
            if (!*done && input(in)) check(igl_op_push(h->op, &in.array));
            else { *done = true; check(igl_op_finish_input(h->op)); }
        }
    };
}
}  // namespace detail

class FilterExec : public ExecutionPlan {
  public:
    FilterExec(std::shared_ptr<Context> ctx, std::shared_ptr<ExecutionPlan> input, ExprBytes predicate)
        : ctx_(std::move(ctx)), input_(std::move(input)), predicate_(std::move(predicate)) {}
    BatchStream execute() override {
        ArrowSchema s{};
        input_->schema(&s);
        auto h = std::make_shared<detail::OpHandle>();
        const int32_t rc = igl_op_filter_project_create(ctx_->get(), &s, predicate_.data(), predicate_.size(), nullptr, nullptr, nullptr, 0, &h->op);
        if (s.release) s.release(&s);
        check(rc);
        return detail::pump(h, input_->execute());
    }
    void schema(ArrowSchema* out) override { input_->schema(out); }   // filter.rs:70-72
  private:
    std::shared_ptr<Context> ctx_;
    std::shared_ptr<ExecutionPlan> input_;
    ExprBytes predicate_;
};

class ProjectionExec : public ExecutionPlan {
  public:
    ProjectionExec(std::shared_ptr<Context> ctx, std::shared_ptr<ExecutionPlan> input,
                   std::vector<std::pair<ExprBytes, std::string>> expressions, ExprBytes fused_predicate = {})
        : ctx_(std::move(ctx)), input_(std::move(input)), exprs_(std::move(expressions)), pred_(std::move(fused_predicate)) {}
    BatchStream execute() override { return detail::pump(create(), input_->execute()); }
    void schema(ArrowSchema* out) override { auto h = create(); check(igl_op_schema(h->op, out)); }
  private:
    std::shared_ptr<detail::OpHandle> create() {
        ArrowSchema s{};
        input_->schema(&s);
        std::vector<const uint8_t*> p; std::vector<size_t> l; std::vector<const char*> n;
        for (auto& e : exprs_) { p.push_back(e.first.data()); l.push_back(e.first.size()); n.push_back(e.second.c_str()); }
        auto h = std::make_shared<detail::OpHandle>();
        const int32_t rc = igl_op_filter_project_create(ctx_->get(), &s, pred_.empty() ? nullptr : pred_.data(), pred_.size(),
                                                        p.data(), l.data(), n.data(), (int32_t)p.size(), &h->op);
        if (s.release) s.release(&s);
        check(rc);
        return h;
    }
    std::shared_ptr<Context> ctx_;
    std::shared_ptr<ExecutionPlan> input_;
    std::vector<std::pair<ExprBytes, std::string>> exprs_;
    ExprBytes pred_;   // non-empty: Projection(Filter(x)) fused by the planner
};

enum class JoinType { Inner = IGL_JOIN_INNER, Left = IGL_JOIN_LEFT, Right = IGL_JOIN_RIGHT, Full = IGL_JOIN_FULL };   // hash_join.rs:16-21

class HashJoinExec : public ExecutionPlan {
  public:
    HashJoinExec(std::shared_ptr<Context> ctx, std::shared_ptr<ExecutionPlan> left, std::shared_ptr<ExecutionPlan> right,
                 ExprBytes left_key, ExprBytes right_key, JoinType jt)
        : ctx_(std::move(ctx)), left_(std::move(left)), right_(std::move(right)), lk_(std::move(left_key)), rk_(std::move(right_key)), jt_(jt) {}
    BatchStream execute() override {
        auto h = create();
        BatchStream build = right_->execute();                    // hash_join.rs:100-128: drain the right child first
        RecordBatch b;
        while (build(b)) check(igl_op_push_build(h->op, &b.array));
        check(igl_op_finish_build(h->op));
        return detail::pump(h, left_->execute());                 // :131-213
    }
    void schema(ArrowSchema* out) override { auto h = create(); check(igl_op_schema(h->op, out)); }
  private:
    std::shared_ptr<detail::OpHandle> create() {
        ArrowSchema ls{}, rs{};
        left_->schema(&ls); right_->schema(&rs);
        auto h = std::make_shared<detail::OpHandle>();
        const int32_t rc = igl_op_hash_join_create(ctx_->get(), (int32_t)jt_, &ls, &rs, lk_.data(), lk_.size(), rk_.data(), rk_.size(), &h->op);
        if (ls.release) ls.release(&ls);
        if (rs.release) rs.release(&rs);
        check(rc);
        return h;
    }
    std::shared_ptr<Context> ctx_;
    std::shared_ptr<ExecutionPlan> left_, right_;
    ExprBytes lk_, rk_;
    JoinType jt_;
};

}  // namespace igloo
