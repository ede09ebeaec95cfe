// Compile-only check of the C++ host mirror against the C ABI (no device needed).
#include "../igloo_b200/host/igloo_operators.hpp"
int main(int argc, char**) {
    if (argc > 100) {   // never runs: only proves the calls type-check and link
        auto ctx = std::make_shared<igloo::Context>(0);
        auto src = std::make_shared<igloo::BatchSourceExec>([] { return igloo::BatchSourceExec::Producer([](ArrowArray*) { return false; }); },
                                                            [](ArrowSchema*) {});
        auto f = std::make_shared<igloo::FilterExec>(ctx, src, igloo::binary(igloo::col(0), igloo::lit_i32(25), IGL_OP_GT));
        auto p = std::make_shared<igloo::ProjectionExec>(ctx, f, std::vector<std::pair<igloo::ExprBytes, std::string>>{{igloo::col(0), "a"}});
        igloo::HashJoinExec j(ctx, p, src, igloo::col(0), igloo::col(0), igloo::JoinType::Inner);
        igloo::RecordBatch b;
        auto s = j.execute();
        while (s(b)) {}
    }
    return 0;
}
