// bh.cu -- Barnes-Hut pipeline (placeholder until the stages land).
#include "common.cuh"

void nb_bh_destroy(nb_context* ctx) { (void)ctx; }

extern "C" {
int nb_bh_accel(nb_context* ctx, const float*, float*, int, float, float, int, void*) {
    return nb_fail(ctx, NB_ERR_INVALID, "Barnes-Hut path not built yet");
}
int nb_bh_accel_groups(nb_context* ctx, const float*, float*, int, float, float, int, int, int, void*) {
    return nb_fail(ctx, NB_ERR_INVALID, "Barnes-Hut path not built yet");
}
int nb_bh_accel_host(nb_context* ctx, const float*, float*, int, float, float, int) {
    return nb_fail(ctx, NB_ERR_INVALID, "Barnes-Hut path not built yet");
}
int nb_bh_stages_host(nb_context* ctx, const float*, int, int, float*, uint32_t*, int32_t*, uint64_t*) {
    return nb_fail(ctx, NB_ERR_INVALID, "Barnes-Hut path not built yet");
}
int nb_bh_export_tree(nb_context* ctx, int*, uint64_t*, int32_t*, int32_t*, double*, double*, double*) {
    return nb_fail(ctx, NB_ERR_INVALID, "Barnes-Hut path not built yet");
}
int nb_bh_last_counters(nb_context* ctx, int64_t*) {
    return nb_fail(ctx, NB_ERR_INVALID, "Barnes-Hut path not built yet");
}
}
