// ab_flock_tiled.cuh — shared-memory tiled flock sweep (placeholder until the
// tiled kernel lands; the generic kernel covers every configuration).
#pragma once
struct TiledPlan { int dummy; };
template <class T, int D>
static bool plan_tiled(ab_handle* h, double visual, double separation, bool exact, TiledPlan* plan) { return false; }
template <class T, int D>
static int launch_tiled(ab_handle* h, const SpaceDev<T, D>& sp, const FlockU<T>& u, bool exact, const TiledPlan& plan,
                        int32_t* ncount) { return ab_fail(h, AB_ERR_INVALID, "tiled kernel unavailable"); }
