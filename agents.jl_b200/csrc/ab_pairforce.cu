// ab_pairforce.cu — fused pair-force step (examples/celllistmap.jl:143-191).
#include "ab_internal.cuh"

extern "C" int ab_step_pairforce(ab_handle* h, const ab_pairforce_params* p, int64_t nsteps) {
    return ab_fail(h, AB_ERR_INVALID, "ab_step_pairforce: not implemented yet");
}
