// ab_flock.cu — fused flock step (examples/flock.jl:79-108) and bulk movement
// (src/spaces/continuous.jl:157-194, src/spaces/walk.jl:58-91).
//
// One step of `ab_step_flock` = [sweep] + [k_scan, k_bucket, k_place] (ab_core.cu):
// the sweep reads the current cell-sorted state, computes every bird's new
// velocity and position from the OLD state (synchronous update), and already
// emits the next step's cell key and arrival rank, so the re-sort that follows
// never touches positions again before the final placement.
//
// Two sweeps:
//   k_flock_generic  any space the reference accepts (mixed periodicity, tiny
//                    grids, clipped offset tables, per-agent parameters).  One
//                    thread per bird, walks the reference's cell offsets as runs
//                    of the sorted arrays through L1/L2.
//   k_flock_fine     (ab_flock_fine.cuh) fine-grid sweep for large 2-D/3-D spaces
//                    with uniform parameters — the throughput path.
#include <algorithm>
#include <cmath>
#include <limits>

#include "ab_internal.cuh"

template <class T> struct FlockU {
    T speed, cohere_factor, separation, separate_factor, match_factor;
    double visual_distance;
    T thr;      // exact_threshold<T>(visual_distance), computed once on the host for uniform parameters
    double gr_t1, gr_t2;   // (spacing - r)², (2 spacing - r)² (negative: never) — grid_radius_12
    int reach;             // fine-cell reach of the separation sweep on the fine path (FinePlan::R)
    const uint8_t* active; // scheduled agents by id (Schedulers.Partially / an id list); nullptr: every agent steps
};

// host twin of exact_threshold (ab_internal.cuh): IEEE sqrt on both sides, same result
template <class T> static T exact_threshold_host(double r) {
    if (!(r >= 0.0)) return T(-1);
    if (std::isinf(r)) return (T)r;
    T s = (T)(r * r);
    const T inf = std::numeric_limits<T>::infinity();
    while ((double)std::sqrt(s) > r) s = std::nextafter(s, -inf);
    for (;;) {
        T nx = std::nextafter(s, inf);
        if (nx == s || !((double)std::sqrt(nx) <= r)) break;
        s = nx;
    }
    return s;
}

template <class T, int D>
__device__ __forceinline__ void flock_epilogue(const SpaceDev<T, D>& sp, const T* p, const T* v, const T* cohere,
                                               const T* separate, const T* match, int N, const FlockU<T>& fp,
                                               T* np, T* nv, int* err) {
    // flock.jl:100-105
    // (x / N and v / norm(v) are evaluated as multiplications by 1/N and 1/sqrt(Σv²): results are
    //  tolerance-checked, not bit-compared — they differ from the reference by an ulp or two)
    T s = 0;
    const T rdenom = T(1) / (T)(N > 1 ? N : 1);
#pragma unroll
    for (int d = 0; d < D; ++d) {
        T c = cohere[d] * fp.cohere_factor;
        T sp_ = separate[d] * fp.separate_factor;
        T m = match[d] * fp.match_factor;
        nv[d] = v[d] + ((c + sp_) + m) * rdenom;
        T sq = nv[d] * nv[d];
        s = (d == 0) ? sq : s + sq;
    }
    const T rnrm = ab_rsqrt(s);
    bool bad = false;
#pragma unroll
    for (int d = 0; d < D; ++d) {
        nv[d] = nv[d] * rnrm;
        // move_agent!(bird, model, speed): continuous.jl:186-194, walk.jl:58-66
        T t = p[d] + fp.speed * nv[d];
        t = normalize_coord(t, sp.extent[d], d == D - 1 ? sp.gper_last : sp.periodic[d]);
        if (!(t >= T(0) && t <= sp.extent[d])) bad = true;    // continuous.jl:160 (NaN lands here too)
        np[d] = t;
    }
    if (bad) atomicOr(err, AB_EF_OUTSIDE);
}

// The reference walk for one bird: nearby_agents (continuous.jl:212-239, grid_multi.jl:135-232) over
// the cell-sorted arrays in global memory, accumulating flock.jl:86-99.
template <class T, int D, bool EXACT>
__device__ __noinline__ void flock_accumulate_generic(const SpaceDev<T, D>& sp, const FlockU<T>& fp, int s, const T* p,
                                                      const int* c, int gr, const T* __restrict__ pos,
                                                      const T* __restrict__ vel, const uint32_t* __restrict__ cell_start,
                                                      int64_t cap, int* Nout, T* cohere, T* separate, T* match) {
    const T thr = fp.thr;
    const T sep2 = fp.separation * fp.separation;
    // min-image shortcut: inside the box and far from wrap, heading == direct
    int permask = 0, slowmask = 0;
#pragma unroll
    for (int d = 0; d < D; ++d) {
        if (sp.periodic[d]) {
            permask |= 1 << d;
            if (!((long long)sp.dims[d] >= 2LL * gr + 3)) slowmask |= 1 << d;
        }
    }
    if (sp.W) slowmask |= 1 << (D - 1);     // slab window: local wrap says nothing about the global boundary — always minimum image
    int N = 0;
    for_each_run<D>(sp.dims, sp.periodic, AB_METRIC_EUCLIDEAN, c, gr, cell_start, sp.kb, false,
                    [&](uint32_t b, uint32_t e, int wrap) {
        const int full = (wrap | slowmask) & permask;
        for (uint32_t j = b; j < e; ++j) {
            if ((int)j == s) continue;       // nearby_ids(agent, ...) drops the agent's own id
            T h[D];
            T d2 = 0;
#pragma unroll
            for (int d = 0; d < D; ++d) {
                T direct = pos[(size_t)d * cap + j] - p[d];
                if (full & (1 << d)) direct = min_image(direct, sp.extent[d]);   // utilities.jl:185-193
                h[d] = direct;
                T sq = direct * direct;
                d2 = (d == 0) ? sq : d2 + sq;
            }
            if (EXACT && !(d2 <= thr)) continue;    // continuous.jl:237-238
            ++N;
            const bool close = d2 < sep2;           // flock.jl:93
#pragma unroll
            for (int d = 0; d < D; ++d) {
                cohere[d] += h[d];
                match[d] += vel[(size_t)d * cap + j];
                if (close) separate[d] -= h[d];
            }
        }
    });
    *Nout = N;
}

template <class T, int D, bool EXACT>
__global__ void __launch_bounds__(128)
k_flock_generic(SpaceDev<T, D> sp, FlockU<T> uni, const T* __restrict__ per_agent, const T* __restrict__ pos,
                const T* __restrict__ vel, const uint32_t* __restrict__ key, const uint32_t* __restrict__ cell_start,
                int64_t cap, int n, T* __restrict__ npos, T* __restrict__ nvel, uint32_t* __restrict__ newkey,
                uint32_t* __restrict__ rank, uint32_t* __restrict__ cell_count, const int32_t* __restrict__ ids,
                int32_t* __restrict__ ncount, int* __restrict__ err) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    FlockU<T> fp = uni;
    if (per_agent) {
        fp.speed = per_agent[(size_t)0 * cap + s];
        fp.cohere_factor = per_agent[(size_t)1 * cap + s];
        fp.separation = per_agent[(size_t)2 * cap + s];
        fp.separate_factor = per_agent[(size_t)3 * cap + s];
        fp.match_factor = per_agent[(size_t)4 * cap + s];
        fp.visual_distance = (double)per_agent[(size_t)5 * cap + s];
        if (EXACT) fp.thr = exact_threshold<T>(fp.visual_distance);
    }
    T p[D], v[D];
    int c[D];
#pragma unroll
    for (int d = 0; d < D; ++d) { p[d] = pos[(size_t)d * cap + s]; v[d] = vel[(size_t)d * cap + s]; }
    if (uni.active && !uni.active[ids[s]]) {
        // not scheduled this tick: the agent keeps its state and stays a neighbour of the others
#pragma unroll
        for (int d = 0; d < D; ++d) { npos[(size_t)d * cap + s] = p[d]; nvel[(size_t)d * cap + s] = v[d]; }
        const uint32_t k0 = key[s];
        newkey[s] = k0;
        rank[s] = atomicAdd(&cell_count[k0], 1u);
        if (ncount) ncount[ids[s]] = 0;
        return;
    }
    key_to_cell<T, D>(sp, key[s], c);
    const int gr = grid_radius<T, D>(sp, p, c, fp.visual_distance);
    T cohere[D], separate[D], match[D];
#pragma unroll
    for (int d = 0; d < D; ++d) cohere[d] = separate[d] = match[d] = T(0);
    int N = 0;
    flock_accumulate_generic<T, D, EXACT>(sp, fp, s, p, c, gr, pos, vel, cell_start, cap, &N, cohere, separate, match);
    T np[D], nv[D];
    flock_epilogue<T, D>(sp, p, v, cohere, separate, match, N, fp, np, nv, err);
    int nc[D];
    bool ok = pos2cell0<T, D>(sp, np, nc);
    if (!ok) atomicOr(err, AB_EF_NUMERIC);
    const uint32_t k = device_key<T, D>(sp, np, nc);
#pragma unroll
    for (int d = 0; d < D; ++d) { npos[(size_t)d * cap + s] = np[d]; nvel[(size_t)d * cap + s] = nv[d]; }
    newkey[s] = k;
    rank[s] = atomicAdd(&cell_count[k], 1u);
    if (ncount) ncount[ids[s]] = N;     // diagnostic, last step of a call only
}

#include "ab_flock_fine.cuh"

// Bulk Euler move / walk for a set of slots (NULL = all):
// move_agent!(agent, model, dt) continuous.jl:186-194 and walk! walk.jl:58-66.
template <class T, int D>
__global__ void k_move(SpaceDev<T, D> sp, T* __restrict__ pos, const T* __restrict__ vel, int64_t cap, int n,
                       const int32_t* __restrict__ slots, int nactive, T dt, const T* __restrict__ directions,
                       int* __restrict__ err) {
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= nactive) return;
    int s = slots ? slots[t] : t;
    if (s < 0 || s >= n) return;
    bool bad = false;
#pragma unroll
    for (int d = 0; d < D; ++d) {
        T dir = directions ? directions[(size_t)t * D + d] : dt * vel[(size_t)d * cap + s];
        T x = pos[(size_t)d * cap + s] + dir;
        x = normalize_coord(x, sp.extent[d], sp.periodic[d]);
        if (!(x >= T(0) && x <= sp.extent[d])) bad = true;
        pos[(size_t)d * cap + s] = x;
    }
    if (bad) atomicOr(err, AB_EF_OUTSIDE);
}

__global__ void k_ids_to_slots(const int64_t* __restrict__ ids, const int32_t* __restrict__ slot_of_id, int n,
                               int nactive, int32_t* __restrict__ slots, int* __restrict__ err) {
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= nactive) return;
    int64_t id = ids[t];
    if (id < 1 || id > n) { slots[t] = -1; atomicOr(err, AB_EF_OVERFLOW); return; }
    slots[t] = slot_of_id[id - 1];
}

// uniform_randomwalk!(agent, model, r) — src/spaces/walk.jl:389-408: v = D standard normals, direction =
// v ./ norm(v) .* r (r = norm(agent.vel) when not given), or (±1, …) * r / sqrt(D) if the normals are all zero;
// agent.vel = direction; then walk!(agent, direction, model).  The normals are Box–Muller pairs on the
// counter stream of ab_internal.cuh (draws 0..3), the signs of the degenerate branch draws 8..8+D-1.
struct RwArgs { double r; uint64_t seed; const double* polar = nullptr; const double* azim = nullptr; };

// randomwalk!(agent, model, r; polar, azimuthal) with angle distributions — src/spaces/walk.jl:236-330.  The host
// draws the angles (anything that supports `rand` in the reference); the device rotates the velocity:
//   2-D: direction = Angle2d(θ) * vel;   3-D: rotate(w, θ, ϕ) of walk.jl:347-357 — w turned by θ about a vector u
//   normal to it (u[n] = w[m], u[m] = -w[n], m the first nonzero component, n = m % 3 + 1), the result turned by ϕ
//   about w.  Rotations.jl (not in the tree; Project.toml compat "1") supplies Angle2d / AngleAxis: an AngleAxis
//   normalises its axis and acts as the Rodrigues rotation v cosθ + (k × v) sinθ + k (k · v)(1 - cosθ).
// Then direction .*= r / norm(vel) when r is given, vel = direction, walk!.  Arithmetic in f64, stored as T.
__device__ __forceinline__ void rodrigues(const double* k_unnorm, double ang, const double* v, double* out) {
    const double kn = sqrt(k_unnorm[0] * k_unnorm[0] + k_unnorm[1] * k_unnorm[1] + k_unnorm[2] * k_unnorm[2]);
    const double k[3] = {k_unnorm[0] / kn, k_unnorm[1] / kn, k_unnorm[2] / kn};
    const double c = cos(ang), s = sin(ang);
    const double kv = k[0] * v[0] + k[1] * v[1] + k[2] * v[2];
    const double cr[3] = {k[1] * v[2] - k[2] * v[1], k[2] * v[0] - k[0] * v[2], k[0] * v[1] - k[1] * v[0]};
#pragma unroll
    for (int d = 0; d < 3; ++d) out[d] = v[d] * c + cr[d] * s + k[d] * kv * (1.0 - c);
}

template <class T, int D>
__global__ void k_rw_rotate(T* __restrict__ vel, int64_t cap, int n, const int32_t* __restrict__ slots, int nactive,
                            double r, const double* __restrict__ polar, const double* __restrict__ azim,
                            T* __restrict__ dirs, int* __restrict__ err) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= nactive) return;
    const int s = slots ? slots[t] : t;
    double w[3] = {0.0, 0.0, 0.0}, out[3] = {0.0, 0.0, 0.0};
    bool ok = s >= 0 && s < n;
    if (ok) {
#pragma unroll
        for (int d = 0; d < D; ++d) w[d] = (double)vel[(size_t)d * cap + s];
        const double nv = sqrt(w[0] * w[0] + w[1] * w[1] + w[2] * w[2]);
        if (!(nv > 0.0)) { atomicOr(err, AB_EF_ARG); ok = false; }
        else {
            if (D == 2) {
                const double c = cos(polar[t]), sn = sin(polar[t]);
                out[0] = c * w[0] - sn * w[1];
                out[1] = sn * w[0] + c * w[1];
            } else {
                const int m = w[0] != 0.0 ? 0 : (w[1] != 0.0 ? 1 : 2), nn = (m + 1) % 3;
                double u[3] = {0.0, 0.0, 0.0}, a[3];
                u[nn] = w[m];
                u[m] = -w[nn];
                rodrigues(u, polar[t], w, a);
                rodrigues(w, azim[t], a, out);
            }
            if (r > 0.0) {
                const double rel = r / nv;
#pragma unroll
                for (int d = 0; d < D; ++d) out[d] *= rel;
            }
        }
    }
#pragma unroll
    for (int d = 0; d < D; ++d) {
        const T dir = ok ? (T)out[d] : T(0);
        if (ok) vel[(size_t)d * cap + s] = dir;
        dirs[(size_t)t * D + d] = dir;
    }
}

template <class T, int D>
__global__ void k_rw_dirs(T* __restrict__ vel, int64_t cap, int n, const int32_t* __restrict__ slots,
                          const int64_t* __restrict__ ids, int nactive, RwArgs rw, T* __restrict__ dirs, int* __restrict__ err) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= nactive) return;
    const int s = slots ? slots[t] : t;
    const int64_t id = ids ? ids[t] : (int64_t)t + 1;
    if (s < 0 || s >= n) {
#pragma unroll
        for (int d = 0; d < D; ++d) dirs[(size_t)t * D + d] = T(0);
        return;
    }
    double r = rw.r;
    if (r < 0.0) {                                            // default r = sqrt(sum(abs2.(agent.vel)))
        T q = T(0);
#pragma unroll
        for (int d = 0; d < D; ++d) { const T v = vel[(size_t)d * cap + s]; q += v * v; }
        r = (double)sqrt(q);
    }
    if (!(r > 0.0)) {
        atomicOr(err, AB_EF_ARG);
#pragma unroll
        for (int d = 0; d < D; ++d) dirs[(size_t)t * D + d] = T(0);
        return;
    }
    T v[D];
#pragma unroll
    for (int pr = 0; pr < (D + 1) / 2; ++pr) {
        const double u1 = ab_rw_u01(rw.seed, id, 2 * pr), u2 = ab_rw_u01(rw.seed, id, 2 * pr + 1);
        const double R = sqrt(-2.0 * log(1.0 - u1)), a = 6.283185307179586 * u2;
        v[2 * pr] = (T)(R * cos(a));
        if (2 * pr + 1 < D) v[2 * pr + 1] = (T)(R * sin(a));
    }
    T q = T(0);
#pragma unroll
    for (int d = 0; d < D; ++d) q += v[d] * v[d];
    const T nv = sqrt(q);
#pragma unroll
    for (int d = 0; d < D; ++d) {
        T dir;
        if (nv != T(0)) dir = (T)((double)(v[d] / nv) * r);
        else dir = (T)((ab_rw_u01(rw.seed, id, 8 + d) < 0.5 ? -1.0 : 1.0) * r / sqrt((double)D));
        vel[(size_t)d * cap + s] = dir;
        dirs[(size_t)t * D + d] = dir;
    }
}

template <class T, int D>
static int move_t(ab_handle* h, double dt, const void* directions_dev, const int32_t* slots, int nactive) {
    if (nactive == 0) return AB_OK;
    AB_LAUNCH(h, "k_move");
    k_move<T, D><<<ab_blocks(nactive, 256), 256, 0, h->stream>>>(make_space<T, D>(h), (T*)h->fields[0].buf[h->cur],
                                                                 (const T*)h->fields[1].buf[h->cur], h->cap, (int)h->n,
                                                                 slots, nactive, (T)dt, (const T*)directions_dev, h->d_err);
    return AB_OK;
}

static int move_common(ab_handle* h, double dt, const void* directions, const int64_t* active_ids, int64_t nactive,
                       const RwArgs* rw = nullptr) {
    if (h->desc.kind != AB_SPACE_CONTINUOUS) return ab_fail(h, AB_ERR_INVALID, "continuous-space call on a grid handle");
    if (h->win_W) return ab_fail(h, AB_ERR_STATE, "bulk moves are not available on a sharded (slab-window) handle");
    AB_CUDA(h, cudaSetDevice(h->device));
    if (h->n == 0) return AB_OK;
    int32_t* slots = nullptr;
    void* dir_dev = nullptr;
    int na = active_ids ? (int)nactive : (int)h->n;
    if (active_ids && nactive < 0) return ab_fail(h, AB_ERR_INVALID, "negative nactive");
    if (!active_ids && directions && nactive != 0 && nactive != h->n)
        return ab_fail(h, AB_ERR_INVALID, "directions for all agents: nactive must be 0 or the agent count");
    size_t dir_bytes = (directions || rw) ? (size_t)na * h->D * h->tsize : 0;
    const int64_t* ids_dev = nullptr;
    const double *d_polar = nullptr, *d_azim = nullptr;
    if (active_ids || directions || rw) {
        size_t ids_bytes = active_ids ? (size_t)na * sizeof(int64_t) : 0;
        size_t slots_bytes = active_ids ? (size_t)na * sizeof(int32_t) : 0;
        size_t o1 = (ids_bytes + 255) / 256 * 256, o2 = o1 + (slots_bytes + 255) / 256 * 256;
        char* base;
        const size_t o3 = o2 + (dir_bytes + 255) / 256 * 256;
        const size_t ang_bytes = (rw && rw->polar) ? (size_t)na * sizeof(double) : 0, ang_stride = (ang_bytes + 255) / 256 * 256;
        AB_TRY(ab_scratch(h, o3 + 2 * ang_stride + 256, (void**)&base));
        if (ang_bytes && na > 0) {
            d_polar = (const double*)(base + o3);
            AB_CUDA(h, cudaMemcpyAsync(base + o3, rw->polar, ang_bytes, cudaMemcpyHostToDevice, h->stream));
            if (rw->azim) {
                d_azim = (const double*)(base + o3 + ang_stride);
                AB_CUDA(h, cudaMemcpyAsync(base + o3 + ang_stride, rw->azim, ang_bytes, cudaMemcpyHostToDevice, h->stream));
            }
        }
        if (active_ids && na > 0) {
            AB_CUDA(h, cudaMemcpyAsync(base, active_ids, ids_bytes, cudaMemcpyHostToDevice, h->stream));
            AB_TRY(ab_ensure_slot_map(h));
            slots = (int32_t*)(base + o1);
            ids_dev = (const int64_t*)base;
            AB_LAUNCH(h, "k_ids_to_slots");
            k_ids_to_slots<<<ab_blocks(na, 256), 256, 0, h->stream>>>((const int64_t*)base, h->slot_of_id, (int)h->n, na, slots, h->d_err);
        }
        if (directions && na > 0) {
            dir_dev = base + o2;
            AB_CUDA(h, cudaMemcpyAsync(dir_dev, directions, dir_bytes, cudaMemcpyHostToDevice, h->stream));
        }
        if (rw && na > 0) dir_dev = base + o2;
    }
    if ((directions || rw) && !active_ids) {
        // directions are given in id order: address slots through the id→slot map
        AB_TRY(ab_ensure_slot_map(h));
        slots = h->slot_of_id;
    }
    if (rw && na > 0) {
        AB_LAUNCH(h, "k_rw_dirs");
        const int nb = (int)ab_blocks(na, 256);
        void* vel = h->fields[1].buf[h->cur];
#define RWD(T, DD) do { if (d_polar) k_rw_rotate<T, DD><<<nb, 256, 0, h->stream>>>((T*)vel, h->cap, (int)h->n, slots, na, rw->r, d_polar, d_azim, (T*)dir_dev, h->d_err); \
                        else k_rw_dirs<T, DD><<<nb, 256, 0, h->stream>>>((T*)vel, h->cap, (int)h->n, slots, ids_dev, na, *rw, (T*)dir_dev, h->d_err); } while (0)
        if (h->dtype == AB_F64) { if (h->D == 2) RWD(double, 2); else RWD(double, 3); }
        else { if (h->D == 2) RWD(float, 2); else RWD(float, 3); }
#undef RWD
        h->agg_valid = false;                                 // Σ vel per cell changed
    }
    int rc;
    if (h->dtype == AB_F64) rc = h->D == 2 ? move_t<double, 2>(h, dt, dir_dev, slots, na) : move_t<double, 3>(h, dt, dir_dev, slots, na);
    else rc = h->D == 2 ? move_t<float, 2>(h, dt, dir_dev, slots, na) : move_t<float, 3>(h, dt, dir_dev, slots, na);
    AB_TRY(rc);
    h->index_valid = false;
    h->pairs_n = -1;
    int e = ab_check_device_errors_now(h);                    // directions / ids belong to the caller: synchronous in every mode
    if (e == AB_ERR_INVALID && !rw) return ab_fail(h, AB_ERR_INVALID, "agent id out of range");
    return e;
}

// part: 0 = all cell rows; 1 = the rows of the interior layers [2, nl - 2) of a slab window (they depend on owned
// aggregates only); 2 = all the other rows (after the halo aggregates are in place).
template <class T, int D>
static int launch_cell_M(ab_handle* h, const SpaceDev<T, D>& sp, int part = 0) {
    const size_t need = (size_t)2 * FineCfg<D>::NC * h->ncells * sizeof(T);
    if (h->cell_M_bytes < need) {
        if (h->cell_M) { AB_CUDA(h, cudaStreamSynchronize(h->stream)); AB_CUDA(h, cudaFree(h->cell_M)); h->cell_M = nullptr; }
        AB_CUDA(h, cudaMalloc(&h->cell_M, need));
        h->cell_M_bytes = need;
    }
    const int wpr = (h->dims[0] + CM_OUT - 1) / CM_OUT;                  // warps per reference cell row
    const int rows_all = (int)(h->ncells / h->dims[0]);
    const int rpl = D == 3 ? h->dims[1] : 1;                             // cell rows per layer
    int skip0 = 0x7fffffff, skiplen = 0, rows = rows_all;
    if (part != 0) {
        const int r0 = 2 * rpl, r1 = std::max(r0, (h->win_nl - 2) * rpl);  // interior rows [r0, r1)
        if (part == 1) { skip0 = 0; skiplen = r0; rows = r1 - r0; }
        else { skip0 = r0; skiplen = r1 - r0; rows = rows_all - (r1 - r0); }
    }
    if (part != 1) h->M_epoch = h->agg_epoch;                            // complete after this launch
    if (rows <= 0) return AB_OK;
    int nfr = D == 2 ? h->K : h->K * h->K;
    const T* aggp = (const T*)h->cell_agg;
    if (nfr == 4) {
        // 4 partials per cell: collapse them once instead of reading all four from every neighbour row
        const int64_t total = (int64_t)FineCfg<D>::NC * h->ncells;
        const size_t need1 = (size_t)total * sizeof(T);
        if (h->cell_agg1_bytes < need1) {
            if (h->cell_agg1) { AB_CUDA(h, cudaStreamSynchronize(h->stream)); AB_CUDA(h, cudaFree(h->cell_agg1)); h->cell_agg1 = nullptr; }
            AB_CUDA(h, cudaMalloc(&h->cell_agg1, need1));
            h->cell_agg1_bytes = need1;
        }
        {
            // all cells; on a slab window with the split schedule: the owned cells with part 1 (they are final after
            // the re-sort), the halo cells with part 2 (after the install)
            const int64_t own = h->win_W ? (h->ncells / h->dims[D - 1]) * h->win_nl : h->ncells;
            const int64_t c0 = part == 2 ? own : 0, nc = part == 1 ? own : h->ncells - c0;
            if (nc > 0) {
                AB_LAUNCH(h, "k_agg_collapse");
                AB_CUDA(h, ab_launch_chain(h->pdl, k_agg_collapse<T>, dim3(ab_blocks(FineCfg<D>::NC * nc, 256)), dim3(256), h->stream, aggp,
                                           (int64_t)h->ncells, c0, nc, (int)FineCfg<D>::NC, (T*)h->cell_agg1));
            }
        }
        aggp = (const T*)h->cell_agg1;
        nfr = 1;
    }
    AB_LAUNCH(h, "k_cell_M");
    if constexpr (D == 2) {
        if (!h->cellm_old) {
            // rolling window: 4 output rows per warp when every row is launched, single rows for the row subsets of
            // the overlapped sharded step (same association: bitwise equal)
            if (part == 0) {
                const int64_t nw = (int64_t)((rows + 3) / 4) * wpr;
                AB_CUDA(h, ab_launch_chain(h->pdl, k_cell_M2<T, 4>, dim3(ab_blocks(nw * 32, 128)), dim3(128), h->stream, sp, aggp,
                                           (int64_t)h->ncells, wpr, (int)nw, (T*)h->cell_M, 0x7fffffff, 0, nfr, rows));
            } else {
                const int64_t nw = (int64_t)rows * wpr;
                AB_CUDA(h, ab_launch_chain(h->pdl, k_cell_M2<T, 1>, dim3(ab_blocks(nw * 32, 128)), dim3(128), h->stream, sp, aggp,
                                           (int64_t)h->ncells, wpr, (int)nw, (T*)h->cell_M, skip0, skiplen, nfr, rows));
            }
            return AB_OK;
        }
    }
    const int64_t nwarps = (int64_t)rows * wpr;
    AB_CUDA(h, ab_launch_chain(h->pdl, k_cell_M<T, D>, dim3(ab_blocks(nwarps * 32, 128)), dim3(128), h->stream, sp, aggp,
                               (int64_t)h->ncells, wpr, (int)nwarps, (T*)h->cell_M, skip0, skiplen, nfr));
    return AB_OK;
}

template <class T, int D, bool EXACT>
static int launch_fine(ab_handle* h, const SpaceDev<T, D>& sp, const FlockU<T>& u, int32_t* ncount, int phase = 0) {
    const int cur = h->cur;
    const int n = h->win_W ? (int)h->cap : (int)h->n;      // slab window: grid over the capacity, live count read on the device
    const int TB = FINE_TB;
    AB_LAUNCH(h, EXACT ? "k_flock_fine_exact" : "k_flock_fine");
    const uint32_t own_keys = h->win_W ? (uint32_t)((h->nfine / h->dims[D - 1]) * h->win_nl) : 0u;
#define FINE_ARGS sp, u, (const T*)h->fields[0].buf[cur], (const T*)h->fields[1].buf[cur], h->key[cur], h->cell_start, \
        (const T*)h->cell_M, (int64_t)h->ncells, h->cap, n, (T*)h->tmp_pos, (T*)h->tmp_vel, h->newkey, make_tilegeo(h), \
        h->id[cur], ncount, h->d_err, h->win_W ? h->cell_start + h->nfine : nullptr, phase, own_keys
    bool done = false;
    if constexpr (!EXACT && D == 2) {
        if (u.reach == 2) {
            AB_CUDA(h, ab_launch_chain(h->pdl, k_flock_fine<T, D, false, true>, dim3(ab_blocks(n, TB)), dim3(TB), h->stream, FINE_ARGS));
            done = true;
        }
    }
    if (!done) AB_CUDA(h, ab_launch_chain(h->pdl, k_flock_fine<T, D, EXACT, false>, dim3(ab_blocks(n, TB)), dim3(TB), h->stream, FINE_ARGS));
#undef FINE_ARGS
    return AB_OK;
}

// One flock step is split in two halves so that several slabs (one process, ab_step_flock_group) can be
// stepped in lockstep:  A = sweep + re-sort (+ pack of the boundary layers on a sharded handle),
// then the boundary messages move,  B = install the halo (sharded) + per-cell sums for the next sweep.
template <class T, int D> struct FlockCtx {
    FlockU<T> u;
    SpaceDev<T, D> sp;
    bool exact = false, fine = false;
    int pa_field = -1;
};

template <class T, int D>
static int flock_prepare(ab_handle* h, const ab_flock_params* p, FlockCtx<T, D>* cx) {
    FlockU<T>& u = cx->u;
    u.speed = (T)p->speed; u.cohere_factor = (T)p->cohere_factor; u.separation = (T)p->separation;
    u.separate_factor = (T)p->separate_factor; u.match_factor = (T)p->match_factor;
    u.visual_distance = p->visual_distance;
    const bool exact = cx->exact = p->search == AB_SEARCH_EXACT;
    u.thr = exact ? exact_threshold_host<T>(p->visual_distance) : T(0);
    {
        const double sd = (double)(T)h->desc.spacing, d1 = sd - p->visual_distance, d2 = 2.0 * sd - p->visual_distance;
        u.gr_t1 = d1 >= 0 ? d1 * d1 : -1.0;
        u.gr_t2 = d2 >= 0 ? d2 * d2 : -1.0;
    }
    cx->pa_field = p->per_agent_field;
    if (cx->pa_field >= 0) {
        if (cx->pa_field >= (int)h->fields.size() || h->fields[cx->pa_field].ncomp != 6 || h->fields[cx->pa_field].dtype != h->dtype)
            return ab_fail(h, AB_ERR_INVALID, "per_agent_field must be a 6-component column of the space's float type");
    }
    // kernel choice
    bool fine = false;
    FinePlan plan;
    if (p->kernel != 1) fine = plan_fine<T, D>(h, p, exact, &plan);
    if (p->kernel == 2 && !fine) return ab_fail(h, AB_ERR_INVALID, "fine-grid flock kernel not applicable to this configuration");
    if (h->win_W && !fine) return ab_fail(h, AB_ERR_INVALID, "a sharded handle runs the fine-grid flock step only (uniform parameters, visual_distance <= spacing, >= 6 cells per axis)");
    cx->fine = fine;
    u.reach = fine ? plan.R : 1;
    u.active = h->step_mask_cur;
    if (fine) AB_TRY(ab_set_refine(h, plan.K));
    AB_TRY(ab_ensure_index(h));
    if (fine) AB_TRY(ab_resort_plan(h));
    cx->sp = make_space<T, D>(h);
    if (fine && !exact) {
        AB_TRY(ab_ensure_agg(h));
        if (h->M_epoch != h->agg_epoch || !h->cell_M) AB_TRY((launch_cell_M<T, D>(h, cx->sp)));   // else: left valid by the previous call
    }
    return AB_OK;
}

template <class T, int D>
static int flock_half_A(ab_handle* h, const FlockCtx<T, D>& cx, bool last) {
    const int n = (int)h->n;
    const int cur = h->cur;
    int32_t* ncount = (last && !h->win_W) ? h->ncount : nullptr;   // neighbour counts by id (diagnostic)
    if (cx.fine) {
        if (cx.exact) AB_TRY((launch_fine<T, D, true>(h, cx.sp, cx.u, ncount)));
        else AB_TRY((launch_fine<T, D, false>(h, cx.sp, cx.u, ncount)));
        AB_TRY(ab_resort_tiled(h, h->tmp_pos, h->tmp_vel, !cx.exact));
        if (h->win_W) AB_TRY(ab_comm_pack(h));
        return AB_OK;
    }
    const T* pos = (const T*)h->fields[0].buf[cur];
    const T* vel = (const T*)h->fields[1].buf[cur];
    const T* pa = cx.pa_field >= 0 ? (const T*)h->fields[cx.pa_field].buf[cur] : nullptr;
    {
        AB_LAUNCH(h, "k_flock_generic");
        const int TB = 128;
        if (cx.exact)
            k_flock_generic<T, D, true><<<ab_blocks(n, TB), TB, 0, h->stream>>>(
                cx.sp, cx.u, pa, pos, vel, h->key[cur], h->cell_start, h->cap, n, (T*)h->tmp_pos, (T*)h->tmp_vel,
                h->newkey, h->rank, h->cell_count, h->id[cur], ncount, h->d_err);
        else
            k_flock_generic<T, D, false><<<ab_blocks(n, TB), TB, 0, h->stream>>>(
                cx.sp, cx.u, pa, pos, vel, h->key[cur], h->cell_start, h->cap, n, (T*)h->tmp_pos, (T*)h->tmp_vel,
                h->newkey, h->rank, h->cell_count, h->id[cur], ncount, h->d_err);
    }
    return ab_reindex_from(h, h->tmp_pos, h->tmp_vel, true);
}

template <class T, int D>
static int flock_half_B(ab_handle* h, const FlockCtx<T, D>& cx) {
    if (!cx.fine) return AB_OK;
    if (h->win_W) AB_TRY(ab_comm_unpack(h, !cx.exact));            // halo refresh + migration (ab_comm.cu)
    if (!cx.exact) AB_TRY((launch_cell_M<T, D>(h, cx.sp)));
    return AB_OK;
}

template <class T, int D>
static int flock_steps_t(ab_handle* h, const ab_flock_params* p, int64_t nsteps) {
    FlockCtx<T, D> cx;
    AB_TRY((flock_prepare<T, D>(h, p, &cx)));
    if (ab_comm_local(h)) return ab_fail(h, AB_ERR_STATE, "in-process slabs are stepped together: ab_step_flock_group");
    AB_CUDA(h, cudaEventRecord(h->ev0, h->stream));
    if (h->win_W && cx.fine && h->overlap_halo) {
        // Sharded handle: the boundary messages of a step travel while the NEXT step's interior layers are swept.
        //   sweep(interior)  -  install the halo sent at the end of the previous step, halo aggregates, M of the
        //   boundary rows  -  sweep(boundary + first halo layers)  -  scan, re-sort, pack + send  -  M of the interior rows
        // The interior birds read no halo data (aggregates within 2 layers, pairs within 1), so the only thing that
        // waits for the neighbours is the thin boundary sweep.  The call ends with the state complete.
        // The halo install and the boundary sweep are short, latency-bound kernels (a few hundred warps): they run on
        // a side stream CONCURRENTLY with the interior sweep (disjoint slots, disjoint cell rows, shared histogram).
        auto install = [&]() -> int {
            AB_TRY(ab_comm_transfer(h));
            AB_TRY(ab_comm_unpack(h, !cx.exact));
            if (!cx.exact) AB_TRY((launch_cell_M<T, D>(h, cx.sp, 2)));
            return AB_OK;
        };
        if (!h->stream2) {
            // highest priority: its CTAs are dispatched as soon as interior-sweep CTAs retire, not after the last one
            int prio_lo = 0, prio_hi = 0;
            AB_CUDA(h, cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
            AB_CUDA(h, cudaStreamCreateWithPriority(&h->stream2, cudaStreamNonBlocking, prio_hi));
            AB_CUDA(h, cudaEventCreateWithFlags(&h->ev_fork, cudaEventDisableTiming));
            AB_CUDA(h, cudaEventCreateWithFlags(&h->ev_join, cudaEventDisableTiming));
        }
        for (int64_t st = 0; st < nsteps; ++st) {
            AB_CUDA(h, cudaEventRecord(h->ev_fork, h->stream));
            AB_CUDA(h, cudaStreamWaitEvent(h->stream2, h->ev_fork, 0));
            if (cx.exact) AB_TRY((launch_fine<T, D, true>(h, cx.sp, cx.u, nullptr, 1)));
            else AB_TRY((launch_fine<T, D, false>(h, cx.sp, cx.u, nullptr, 1)));
            std::swap(h->stream, h->stream2);                    // everything queued below goes to the side stream
            int rc = st > 0 ? install() : AB_OK;
            if (rc == AB_OK) rc = cx.exact ? launch_fine<T, D, true>(h, cx.sp, cx.u, nullptr, 2) : launch_fine<T, D, false>(h, cx.sp, cx.u, nullptr, 2);
            cudaError_t ce = cudaEventRecord(h->ev_join, h->stream);
            std::swap(h->stream, h->stream2);
            AB_TRY(rc);
            AB_CUDA(h, ce);
            AB_CUDA(h, cudaStreamWaitEvent(h->stream, h->ev_join, 0));
            AB_TRY(ab_resort_tiled(h, h->tmp_pos, h->tmp_vel, !cx.exact));
            AB_TRY(ab_comm_pack(h));
            if (!cx.exact) AB_TRY((launch_cell_M<T, D>(h, cx.sp, 1)));
        }
        AB_TRY(install());
        AB_CUDA(h, cudaEventRecord(h->ev1, h->stream));
        return AB_OK;
    }
    // Direct exchange: the halo install of a step runs on a side stream CONCURRENTLY with the step's own re-sort.
    // It needs (i) the neighbours' messages — their boundary tiles run first in THEIR re-sort —, (ii) the owned count,
    // which the step's per-row counters give since the sweep (scan-free spaces) or the scanned tile totals since the
    // small scan in front of the re-sort (the event is recorded between the two), and it writes only halo slots,
    // halo cell_start entries and halo aggregates, none of which the re-sort touches.  k_cell_M joins both.
    const bool side_unpack = h->win_W && cx.fine && ab_comm_direct_on(h) && !h->no_side_unpack;
    if (side_unpack && !h->stream2) {
        AB_CUDA(h, cudaStreamCreateWithFlags(&h->stream2, cudaStreamNonBlocking));
        AB_CUDA(h, cudaEventCreateWithFlags(&h->ev_fork, cudaEventDisableTiming));
        AB_CUDA(h, cudaEventCreateWithFlags(&h->ev_join, cudaEventDisableTiming));
    }
    for (int64_t st = 0; st < nsteps; ++st) {
        if (side_unpack) {
            int32_t* ncount = nullptr;
            if (cx.exact) AB_TRY((launch_fine<T, D, true>(h, cx.sp, cx.u, ncount)));
            else AB_TRY((launch_fine<T, D, false>(h, cx.sp, cx.u, ncount)));
            if (h->rs_scanfree) AB_CUDA(h, cudaEventRecord(h->ev_fork, h->stream));   // the counters of this step are complete
            AB_TRY(ab_resort_tiled(h, h->tmp_pos, h->tmp_vel, !cx.exact, h->rs_scanfree ? nullptr : h->ev_fork));   // delivers small boundary messages too
            // Large messages (3-D) are packed by a launch of their own; the install then cannot hide behind the re-sort,
            // so it hides behind the cell sums of the INTERIOR rows instead (they need no halo aggregate):
            //   main: pack - M(interior rows)            side: wait for the flags - install
            //   join - M(boundary rows)
            const bool split_M = !h->halo_packed && !cx.exact && !h->no_split_M;
            AB_TRY(ab_comm_pack(h));
            AB_TRY(ab_comm_transfer(h));
            AB_CUDA(h, cudaStreamWaitEvent(h->stream2, h->ev_fork, 0));
            std::swap(h->stream, h->stream2);                                // the install is queued on the side stream
            h->side_unpack_now = true;
            const bool have_counters = h->unpack_rows || h->unpack_tb;       // what the install sums the owned count from
            const int rc = have_counters ? ab_comm_unpack(h, !cx.exact) : AB_OK;
            h->side_unpack_now = false;
            const cudaError_t ce = cudaEventRecord(h->ev_join, h->stream);
            std::swap(h->stream, h->stream2);
            if (!have_counters) return ab_fail(h, AB_ERR_STATE, "side-stream halo install without row counters");
            AB_TRY(rc);
            AB_CUDA(h, ce);
            if (split_M) AB_TRY((launch_cell_M<T, D>(h, cx.sp, 1)));
            AB_CUDA(h, cudaStreamWaitEvent(h->stream, h->ev_join, 0));
            if (!cx.exact) AB_TRY((launch_cell_M<T, D>(h, cx.sp, split_M ? 2 : 0)));
            continue;
        }
        AB_TRY((flock_half_A<T, D>(h, cx, st == nsteps - 1)));
        if (h->win_W) AB_TRY(ab_comm_transfer(h));
        AB_TRY((flock_half_B<T, D>(h, cx)));
    }
    AB_CUDA(h, cudaEventRecord(h->ev1, h->stream));
    return AB_OK;
}

// All slabs of one process in lockstep (handles in rank order, linked by ab_comm_link_local).
template <class T, int D>
static int flock_group_t(ab_handle** hs, int nh, const ab_flock_params* p, int64_t nsteps) {
    std::vector<FlockCtx<T, D>> cx(nh);
    for (int r = 0; r < nh; ++r) {
        AB_CUDA(hs[r], cudaSetDevice(hs[r]->device));
        AB_TRY((flock_prepare<T, D>(hs[r], p, &cx[r])));
        if (cx[r].exact != cx[0].exact || hs[r]->K != hs[0]->K) return ab_fail(hs[r], AB_ERR_INVALID, "slabs disagree on the step plan");
        AB_CUDA(hs[r], cudaEventRecord(hs[r]->ev0, hs[r]->stream));
    }
    for (int64_t st = 0; st < nsteps; ++st) {
        for (int r = 0; r < nh; ++r) {
            AB_CUDA(hs[r], cudaSetDevice(hs[r]->device));
            AB_TRY((flock_half_A<T, D>(hs[r], cx[r], false)));
        }
        for (int r = 0; r < nh; ++r) {
            AB_CUDA(hs[r], cudaSetDevice(hs[r]->device));
            AB_TRY(ab_comm_transfer(hs[r]));
            AB_TRY((flock_half_B<T, D>(hs[r], cx[r])));
        }
    }
    for (int r = 0; r < nh; ++r) {
        AB_CUDA(hs[r], cudaSetDevice(hs[r]->device));
        AB_CUDA(hs[r], cudaEventRecord(hs[r]->ev1, hs[r]->stream));
    }
    return AB_OK;
}

// By-id mask of the agents a scheduler returned (Schedulers.Partially, a custom id list — schedulers.jl:68-120):
// h->step_mask[id - 1] = 1 for the scheduled ones.  Shared by the flock, pair-force and Schelling steps.
int ab_set_step_mask(ab_handle* h, const int64_t* active_ids, int64_t nactive) {
    if (h->win_W) return ab_fail(h, AB_ERR_STATE, "a scheduled subset is not available on a sharded handle");
    if (nactive < 0) return ab_fail(h, AB_ERR_INVALID, "negative nactive");
    AB_CUDA(h, cudaSetDevice(h->device));
    std::vector<uint8_t> mask((size_t)h->n, 0);
    for (int64_t k = 0; k < nactive; ++k) {
        if (active_ids[k] < 1 || active_ids[k] > h->n) return ab_fail(h, AB_ERR_INVALID, "active id out of range");
        mask[(size_t)active_ids[k] - 1] = 1;
    }
    if (h->step_mask_bytes < (size_t)h->n) {
        if (h->step_mask) { AB_CUDA(h, cudaStreamSynchronize(h->stream)); AB_CUDA(h, cudaFree(h->step_mask)); h->step_mask = nullptr; }
        AB_CUDA(h, cudaMalloc(&h->step_mask, (size_t)h->cap));
        h->step_mask_bytes = (size_t)h->cap;
    }
    // pageable source: the copy has left `mask` when the call returns; stream order protects the device buffer
    AB_CUDA(h, cudaMemcpyAsync(h->step_mask, mask.data(), (size_t)h->n, cudaMemcpyHostToDevice, h->stream));
    return AB_OK;
}

extern "C" {

int ab_move(ab_handle* h, double dt, const int64_t* active_ids, int64_t nactive) {
    return move_common(h, dt, nullptr, active_ids, nactive);
}

int ab_walk(ab_handle* h, const void* directions, const int64_t* active_ids, int64_t nactive, int ifempty) {
    if (h->desc.kind != AB_SPACE_CONTINUOUS) {
        return ab_grid_walk(h, directions, active_ids, nactive, ifempty);
    }
    if (!directions) return ab_fail(h, AB_ERR_INVALID, "null directions");
    return move_common(h, 0.0, directions, active_ids, nactive);
}

int ab_randomwalk(ab_handle* h, double r, uint64_t seed, const int64_t* active_ids, int64_t nactive, int ifempty,
                  int force_motion) {
    if (h->desc.kind != AB_SPACE_CONTINUOUS) return ab_grid_randomwalk(h, r, seed, active_ids, nactive, ifempty, force_motion);
    if (r == 0.0) return ab_fail(h, AB_ERR_INVALID, "ArgumentError: The displacement must be larger than 0.");
    RwArgs rw{r, seed};                                       // r < 0: the agent's own speed
    return move_common(h, 0.0, nullptr, active_ids, nactive, &rw);
}

int ab_randomwalk_angles(ab_handle* h, double r, const double* polar, const double* azimuthal, const int64_t* active_ids,
                         int64_t nactive) {
    if (h->desc.kind != AB_SPACE_CONTINUOUS) return ab_fail(h, AB_ERR_INVALID, "angle distributions are for ContinuousSpace random walks");
    if (!polar) return ab_fail(h, AB_ERR_INVALID, "null polar angles");
    if (h->D == 3 && !azimuthal) return ab_fail(h, AB_ERR_INVALID, "3-D random walks need azimuthal angles");
    if (r == 0.0) return ab_fail(h, AB_ERR_INVALID, "ArgumentError: The displacement must be larger than 0.");
    RwArgs rw{r, 0, polar, h->D == 3 ? azimuthal : nullptr};
    const bool was_async = h->async_mode;
    h->async_mode = false;                                    // the angle arrays belong to the caller: finish before returning
    const int rc = move_common(h, 0.0, nullptr, active_ids, nactive, &rw);
    h->async_mode = was_async;
    return rc;
}

int ab_step_flock(ab_handle* h, const ab_flock_params* p, int64_t nsteps) {
    if (!p) return ab_fail(h, AB_ERR_INVALID, "null params");
    if (h->desc.kind != AB_SPACE_CONTINUOUS) return ab_fail(h, AB_ERR_INVALID, "flock rule needs a ContinuousSpace handle");
    if (nsteps < 0) return ab_fail(h, AB_ERR_INVALID, "negative nsteps");
    AB_CUDA(h, cudaSetDevice(h->device));
    if (h->n == 0 || nsteps == 0) return AB_OK;
    int rc;
    if (h->dtype == AB_F64) rc = h->D == 2 ? flock_steps_t<double, 2>(h, p, nsteps) : flock_steps_t<double, 3>(h, p, nsteps);
    else rc = h->D == 2 ? flock_steps_t<float, 2>(h, p, nsteps) : flock_steps_t<float, 3>(h, p, nsteps);
    if (rc != AB_OK) return ab_step_failed(h, rc);
    if (h->async_mode) { h->step_pending = true; return AB_OK; }     // errors and timing: ab_sync
    if ((rc = ab_check_device_errors(h)) != AB_OK) return ab_step_state_lost(h, rc);
    if (h->win_W) AB_TRY(ab_comm_refresh_counts(h));
    float ms = 0;
    AB_CUDA(h, cudaEventElapsedTime(&ms, h->ev0, h->ev1));
    h->last_step_ms = ms;
    return AB_OK;
}

int ab_step_flock_active(ab_handle* h, const ab_flock_params* p, int64_t nsteps, const int64_t* active_ids, int64_t nactive) {
    if (!active_ids) return ab_step_flock(h, p, nsteps);
    if (h->n == 0 || nsteps == 0) return AB_OK;
    AB_TRY(ab_set_step_mask(h, active_ids, nactive));
    h->step_mask_cur = h->step_mask;
    const int rc = ab_step_flock(h, p, nsteps);
    h->step_mask_cur = nullptr;
    return rc;
}

int ab_step_flock_group(ab_handle** hs, int nh, const ab_flock_params* p, int64_t nsteps) {
    if (!hs || nh < 2 || !p) return AB_ERR_INVALID;
    ab_handle* h0 = hs[0];
    if (nsteps < 0) return ab_fail(h0, AB_ERR_INVALID, "negative nsteps");
    for (int r = 0; r < nh; ++r) {
        if (!hs[r] || !ab_comm_local(hs[r])) return ab_fail(hs[r] ? hs[r] : h0, AB_ERR_STATE, "link the slabs first (ab_comm_link_local)");
        if (hs[r]->dtype != h0->dtype || hs[r]->D != h0->D) return ab_fail(hs[r], AB_ERR_INVALID, "slabs of one space must share D and eltype");
    }
    if (nsteps == 0) return AB_OK;
    int rc;
    if (h0->dtype == AB_F64) rc = h0->D == 2 ? flock_group_t<double, 2>(hs, nh, p, nsteps) : flock_group_t<double, 3>(hs, nh, p, nsteps);
    else rc = h0->D == 2 ? flock_group_t<float, 2>(hs, nh, p, nsteps) : flock_group_t<float, 3>(hs, nh, p, nsteps);
    if (rc != AB_OK) for (int r = 0; r < nh; ++r) ab_step_failed(hs[r], rc);
    if (rc != AB_OK) { if (h0->err.empty()) for (int r = 0; r < nh; ++r) if (!hs[r]->err.empty()) { h0->err = hs[r]->err; break; } return rc; }
    double worst = 0;
    for (int r = 0; r < nh; ++r) {
        ab_handle* h = hs[r];
        AB_CUDA(h, cudaSetDevice(h->device));
        int e = ab_check_device_errors(h);
        if (e != AB_OK) { h0->err = h->err; return e; }
        AB_TRY(ab_comm_refresh_counts(h));
        float ms = 0;
        AB_CUDA(h, cudaEventElapsedTime(&ms, h->ev0, h->ev1));
        h->last_step_ms = ms;
        worst = std::max(worst, (double)ms);
    }
    h0->last_step_ms = worst;
    return AB_OK;
}

int ab_flock_last_counts(ab_handle* h, int32_t* counts, int64_t n) {
    if (h->win_W) return ab_fail(h, AB_ERR_STATE, "not available on a sharded handle");
    if (n != h->n) return ab_fail(h, AB_ERR_INVALID, "length must equal the agent count");
    if (n == 0) return AB_OK;
    AB_CUDA(h, cudaSetDevice(h->device));
    AB_CUDA(h, cudaMemcpyAsync(counts, h->ncount, (size_t)n * 4, cudaMemcpyDeviceToHost, h->stream));
    AB_CUDA(h, cudaStreamSynchronize(h->stream));
    return AB_OK;
}

}  // extern "C"
