// dist.cu — multi-GPU block-rotation trainer; filled in below.
#include "common.cuh"

struct DistState { int dummy; };

void ge_dist_free(ge_ctx* ctx) {
    delete ctx->dist;
    ctx->dist = nullptr;
}

extern "C" int ge_dist_unique_id(void*) { return GE_ERR_DIST; }
extern "C" int ge_dist_init(ge_ctx* ctx, int32_t, int32_t, const void*) { if (ctx) ctx->err = "multi-GPU path not built yet"; return GE_ERR_DIST; }
extern "C" int ge_dist_prepare(ge_ctx* ctx, int32_t) { if (ctx) ctx->err = "multi-GPU path not built yet"; return GE_ERR_DIST; }
extern "C" int ge_dist_train_epoch(ge_ctx* ctx, double*) { if (ctx) ctx->err = "multi-GPU path not built yet"; return GE_ERR_DIST; }
