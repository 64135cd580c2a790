// ab_grid.cu — GridSpace / GridSpaceSingle queries and the Schelling rule.
#include "ab_internal.cuh"

int ab_grid_build(ab_handle* h) { return ab_fail(h, AB_ERR_INVALID, "grid spaces: not implemented yet"); }
int ab_grid_walk(ab_handle* h, const void*, const int64_t*, int64_t, int) { return ab_fail(h, AB_ERR_INVALID, "grid spaces: not implemented yet"); }
int ab_grid_nearby(ab_handle* h, const void*, const int64_t*, int64_t, double, int, int64_t*, const int64_t*, int64_t*) {
    return ab_fail(h, AB_ERR_INVALID, "grid spaces: not implemented yet");
}
extern "C" int ab_step_schelling(ab_handle* h, const ab_schelling_params* p, int64_t nsteps, int64_t* n_moved) {
    return ab_fail(h, AB_ERR_INVALID, "ab_step_schelling: not implemented yet");
}
