// arrow_ops.cpp — streaming operators over host Arrow batches (push / pull).
// PLACEHOLDER: replaced by the real implementation in the next milestone; every
// entry point reports IGL_ERR_UNSUPPORTED so nothing silently falls back.
#include "internal.h"

using namespace igl;

extern "C" {
int32_t igl_op_filter_project_create(igl_ctx*, const struct ArrowSchema*, const uint8_t*, size_t, const uint8_t* const*,
                                     const size_t*, const char* const*, int32_t, igl_op** out) {
    if (out) *out = nullptr;
    return fail(IGL_ERR_UNSUPPORTED, "igl_op_filter_project_create: not built yet");
}
int32_t igl_op_hash_join_create(igl_ctx*, int32_t, const struct ArrowSchema*, const struct ArrowSchema*, const uint8_t*,
                                size_t, const uint8_t*, size_t, igl_op** out) {
    if (out) *out = nullptr;
    return fail(IGL_ERR_UNSUPPORTED, "igl_op_hash_join_create: not built yet");
}
int32_t igl_op_schema(igl_op*, struct ArrowSchema*) { return fail(IGL_ERR_UNSUPPORTED, "igl_op_schema: not built yet"); }
int32_t igl_op_push_build(igl_op*, const struct ArrowArray*) { return fail(IGL_ERR_UNSUPPORTED, "igl_op_push_build: not built yet"); }
int32_t igl_op_finish_build(igl_op*) { return fail(IGL_ERR_UNSUPPORTED, "igl_op_finish_build: not built yet"); }
int32_t igl_op_push(igl_op*, const struct ArrowArray*) { return fail(IGL_ERR_UNSUPPORTED, "igl_op_push: not built yet"); }
int32_t igl_op_finish_input(igl_op*) { return fail(IGL_ERR_UNSUPPORTED, "igl_op_finish_input: not built yet"); }
int32_t igl_op_pull(igl_op*, struct ArrowArray*) { return fail(IGL_ERR_UNSUPPORTED, "igl_op_pull: not built yet"); }
void igl_op_destroy(igl_op*) {}
}
