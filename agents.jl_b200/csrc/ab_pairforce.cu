// ab_pairforce.cu — fused pair-force step: force law examples/celllistmap.jl:143-155, integrator
// :176-191, synchronous by construction (forces first, then every agent moves).
//
// Pair enumeration follows the reference's own interacting_pairs(model, cutoff, :all; search = :exact)
// (src/spaces/continuous.jl:418-452): the pair (i, j) exists when i finds j OR j finds i in its
// nearby_ids walk with the exact distance filter d <= cutoff.  (The example itself delegates the
// enumeration to CellListMap.jl, which is not part of the reference tree.)
//
// Two paths:
//   fine     cutoff <= spacing on a grid of >= 3 cells per dimension: every pair within the cutoff
//            lies in the 3^D block of reference cells around either agent, so each agent sums its
//            own force over that block (every pair is evaluated from both sides — no atomics, no
//            force array, one canonical summation order) and integrates in the same kernel.
//            Followed by k_scan + k_resort; a device flag says whether some agent moved further than one
//            cell, and such a step is re-indexed by the generic bucket/place kernels instead (the flag is read
//            back once per step: one 4-byte copy + stream synchronisation).
//   generic  anything else: one walk per agent over the reference's offset table; pairs that only one
//            side sees hand the reaction to the other side through a floating-point atomicAdd
//            (the only non-deterministic sum in the library, and only for such one-sided pairs).
#include <cmath>
#include <limits>
#include <type_traits>

#include "ab_internal.cuh"

template <class T> struct PfU {
    T dt, half_dt2;
    double cutoff;
    T thr;          // exact_threshold<T>(cutoff): d <= cutoff  <=>  d² <= thr
    const uint8_t* active;   // scheduled agents by id (Schedulers.Partially / an id list); nullptr: every agent moves
};

template <class T> static T pf_threshold_host(double r) {
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

// force on i from the pair (i, j): fij = 2 (k_i k_j) (d² - (r_i + r_j)²) (dr / d) when d <= r_i + r_j
// (celllistmap.jl:143-155).  `d <= rs` is decided on d² with a guard band; inside the band the
// reference's own comparison sqrt(d²) <= rs is evaluated.  dr/d is computed as dr * rsqrt(d²)
// (tolerance-checked result).
template <class T, int D>
__device__ __forceinline__ void pf_pair(const T* dr, T d2, T ri, T ki, T rj, T kj, T* f) {
    const T rs = ri + rj;
    const T rs2 = rs * rs;
    const T band = sizeof(T) == 8 ? T(1e-14) : T(1e-5);
    bool in;
    if (d2 < rs2 * (T(1) - band)) in = true;
    else if (d2 > rs2 * (T(1) + band)) in = false;
    else in = ab_sqrt(d2) <= rs;
    if (!in) return;
    const T pref = (T(2) * (ki * kj)) * (d2 - rs2);
    const T rinv = ab_rsqrt(d2);
#pragma unroll
    for (int d = 0; d < D; ++d) f[d] += pref * (dr[d] * rinv);
}

// integrator celllistmap.jl:176-191 + normalize_position walk.jl:77-91 + move_agent! continuous.jl:157-169
template <class T, int D>
__device__ __forceinline__ bool pf_integrate(const SpaceDev<T, D>& sp, const PfU<T>& u, const T* p, const T* v, const T* f,
                                             T mass, T* np, T* nv) {
    bool bad = false;
#pragma unroll
    for (int d = 0; d < D; ++d) {
        const T a = f[d] / mass;
        nv[d] = v[d] + a * u.dt;
        T x = (p[d] + nv[d] * u.dt) + (a / T(2)) * (u.dt * u.dt);
        x = normalize_coord(x, sp.extent[d], d == D - 1 ? sp.gper_last : sp.periodic[d]);
        if (!(x >= T(0) && x <= sp.extent[d])) bad = true;
        np[d] = x;
    }
    return !bad;
}

// Own force of one agent by the reference walk (nearby_ids with the exact filter, continuous.jl:212-239):
// used by the fine kernel for agents within rounding distance of a cell boundary, whose partners may sit
// (in floating point) one cell beyond the 3^D block.
template <class T, int D>
__device__ __noinline__ void pf_accumulate_walk(const SpaceDev<T, D>& sp, const PfU<T>& u, int s, const T* p, const int* c,
                                                const int* cg, T ri, T ki, const T* __restrict__ pos, const T* __restrict__ rad,
                                                const T* __restrict__ kk, const uint32_t* __restrict__ cs, int64_t cap, T* f) {
    // c: the agent's cell in this handle's (window) frame — the walk; cg: its global cell — the cell centre of grid_r
    const int gr = grid_radius<T, D>(sp, p, cg, u.cutoff);
    for_each_run<D>(sp.dims, sp.periodic, AB_METRIC_EUCLIDEAN, c, gr, cs, sp.kb, false, [&](uint32_t b, uint32_t e, int) {
        for (uint32_t j = b; j < e; ++j) {
            if ((int)j == s) continue;
            T dr[D];
            T dd = 0;
#pragma unroll
            for (int d = 0; d < D; ++d) {
                T direct = pos[(size_t)d * cap + j] - p[d];
                if (sp.periodic[d]) direct = min_image(direct, sp.extent[d]);
                dr[d] = direct;
                const T sq = direct * direct;
                dd = (d == 0) ? sq : dd + sq;
            }
            if (!(dd <= u.thr)) continue;
            pf_pair<T, D>(dr, dd, ri, ki, rad[j], kk[j], f);
        }
    });
}

#ifndef PF_UNROLL
#define PF_UNROLL 2
#endif
#ifndef PF_MINB
#define PF_MINB 8
#endif
constexpr int kPfUnroll = PF_UNROLL;
template <class T, int D>
__global__ void __launch_bounds__(128, PF_MINB)
k_pf_fine(SpaceDev<T, D> sp, PfU<T> u, const T* __restrict__ pos, const T* __restrict__ vel, const T* __restrict__ rad,
          const T* __restrict__ kk, const T* __restrict__ mass, const uint32_t* __restrict__ key,
          const uint32_t* __restrict__ cs, int64_t cap, int n, T* __restrict__ npos, T* __restrict__ nvel,
          uint32_t* __restrict__ newkey, uint32_t* __restrict__ rank, TileGeo tg, const int32_t* __restrict__ ids,
          int* __restrict__ far, int* __restrict__ err, const uint32_t* __restrict__ nslots) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= (nslots ? (int)*nslots : n)) return;            // slab window: the slot count lives on the device
    T p[D], v[D], f[D];
    int c[D];
#pragma unroll
    for (int d = 0; d < D; ++d) { p[d] = pos[(size_t)d * cap + s]; v[d] = vel[(size_t)d * cap + s]; f[d] = T(0); }
    if (u.active && !u.active[ids[s]]) {
        // not scheduled this tick (agent_step! of celllistmap.jl:176-191 does not run for it): the particle keeps its
        // state and remains a force partner of the others
        const uint32_t k0 = key[s];
#pragma unroll
        for (int d = 0; d < D; ++d) { npos[(size_t)d * cap + s] = p[d]; nvel[(size_t)d * cap + s] = v[d]; }
        newkey[s] = k0;
        const uint32_t fr0 = k0 / (uint32_t)sp.dims[0];
        if (fr0 >= tg.nrows) return;
        rank[s] = atomicAdd(&tg.cell_count[k0], 1u);
        count_tile(tg, fr0, k0 - fr0 * (uint32_t)sp.dims[0]);
        return;
    }
    const T ri = rad[s], ki = kk[s], mi = mass[s];
    key_to_cell<T, D>(sp, key[s], c);        // K == 1 on this path: fine cell == reference cell
    int cg[D];
#pragma unroll
    for (int d = 0; d < D; ++d) cg[d] = c[d];
    if (sp.W) {
        // slab window: owned layers plus the first halo layer on each side (its particles may enter the owned layers
        // in this step: they are this step's immigrants, computed here instead of being received)
        const int l = c[D - 1];
        if (!(l <= sp.nl || l == sp.nl + 2 * sp.W - 1)) return;
        cg[D - 1] = local_to_global_last<T, D>(sp, l);
    }
    const int d0 = sp.dims[0], d1 = sp.dims[1], d2n = D == 3 ? sp.dims[2] : 1;
    // cutoff may equal the spacing: a partner at distance ~cutoff can then sit (by rounding) two cells away,
    // but only if BOTH agents are within rounding distance of a cell boundary — those take the reference walk
    bool near = false;
    {
        const T eps = sizeof(T) == 8 ? T(1.4210854715202004e-14) : T(7.62939453125e-06);     // 2^-46, 2^-17
#pragma unroll
        for (int d = 0; d < D; ++d) {
            const T tau = sp.extent[d] * eps;
            const T a = p[d] - sp.spacing * (T)cg[d];
            if (a < tau || sp.spacing - a < tau) near = true;
        }
    }
    auto sweep_t = [&](uint32_t jb, uint32_t je, int wrap, auto wrapped) {
        constexpr bool WRAP = decltype(wrapped)::value;
#pragma unroll kPfUnroll
        for (uint32_t j = jb; j < je; ++j) {
            T dr[D];
            T dd = 0;
#pragma unroll
            for (int d = 0; d < D; ++d) {
                T direct = pos[(size_t)d * cap + j] - p[d];
                if (WRAP && (wrap & (1 << d))) direct = min_image(direct, sp.extent[d]);     // utilities.jl:185-193
                dr[d] = direct;
                const T sq = direct * direct;
                dd = (d == 0) ? sq : dd + sq;
            }
            // Branch-free form of "pair within the cutoff (continuous.jl:237-238) and within r_i + r_j
            // (celllistmap.jl:147): add fij" — the lanes of a warp disagree on almost every candidate, so a
            // branchy body runs most of its paths for most candidates with half the lanes idle (measured: 89 warp
            // instructions per candidate, 15.6 of 32 lanes active).  Excluded pairs get the weight 0 and a
            // harmless d² = 1 for the rsqrt; included pairs are evaluated exactly as pf_pair does.
            const T rj = rad[j], kj = kk[j];
            const T rs = ri + rj;
            const T rs2 = rs * rs;
            const T band = sizeof(T) == 8 ? T(1e-14) : T(1e-5);
            bool in = dd < rs2 * (T(1) - band);
            if (!in && !(dd > rs2 * (T(1) + band))) in = ab_sqrt(dd) <= rs;                   // inside the rounding band: the reference's own comparison
            in = in && (int)j != s && dd <= u.thr;
            const T dsafe = in ? dd : T(1);
            const T pref = in ? (T(2) * (ki * kj)) * (dd - rs2) : T(0);
            const T rinv = ab_rsqrt(dsafe);
#pragma unroll
            for (int d = 0; d < D; ++d) f[d] += pref * (dr[d] * rinv);
        }
    };
    auto sweep = [&](uint32_t jb, uint32_t je, int wrap) {
        if (wrap == 0) sweep_t(jb, je, 0, std::false_type{});
        else sweep_t(jb, je, wrap, std::true_type{});
    };
    if (near) pf_accumulate_walk<T, D>(sp, u, s, p, c, cg, ri, ki, pos, rad, kk, cs, cap, f);
    else
    for (int rz = (D == 3 ? -1 : 0); rz <= (D == 3 ? 1 : 0); ++rz) {
        int cz = 0, wz = 0;
        if (D == 3) {
            cz = c[2] + rz;
            if (cz < 0 || cz >= d2n) {
                if (!sp.periodic[2]) continue;
                cz = cz < 0 ? cz + d2n : cz - d2n;
                wz = 4;
            }
            // slab window: whether the row lies across the space's periodic boundary is decided on GLOBAL layers
            if (sp.W) { const int g = cg[2] + rz; wz = (g < 0 || g >= sp.gdl) ? 4 : 0; }
        }
        uint32_t rbv[3], jb[3], je[3];
        int wrv[3];
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            wrv[i] = -1; jb[i] = je[i] = 0;
            int cy = c[1] + i - 1, wy = 0;
            if (cy < 0 || cy >= d1) {
                if (!sp.periodic[1]) continue;
                cy = cy < 0 ? cy + d1 : cy - d1;
                wy = 2;
            }
            if (D == 2 && sp.W) { const int g = cg[1] + i - 1; wy = (g < 0 || g >= sp.gdl) ? 2 : 0; }
            const uint32_t rb = (D == 3 ? ((uint32_t)cz * (uint32_t)d1 + (uint32_t)cy) : (uint32_t)cy) * (uint32_t)d0;
            rbv[i] = rb; wrv[i] = wy | wz;
            const int a0 = max(c[0] - 1, 0), a1 = min(c[0] + 1, d0 - 1);
            jb[i] = cs[rb + (uint32_t)a0]; je[i] = cs[rb + (uint32_t)a1 + 1];
        }
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            if (wrv[i] < 0) continue;
            sweep(jb[i], je[i], wrv[i]);
            if (sp.periodic[0]) {
                if (c[0] == 0) sweep(cs[rbv[i] + (uint32_t)(d0 - 1)], cs[rbv[i] + (uint32_t)d0], wrv[i] | 1);
                if (c[0] == d0 - 1) sweep(cs[rbv[i]], cs[rbv[i] + 1u], wrv[i] | 1);
            }
        }
    }
    T np_[D], nv[D];
    if (!pf_integrate<T, D>(sp, u, p, v, f, mi, np_, nv)) atomicOr(err, AB_EF_OUTSIDE);
    int nc[D];
    if (!pos2cell0<T, D>(sp, np_, nc)) atomicOr(err, AB_EF_NUMERIC);
    bool moved_far = false;
#pragma unroll
    for (int d = 0; d < D; ++d) {
        int dc = nc[d] - cg[d];
        if (dc < 0) dc = -dc;
        const int lim = d == D - 1 ? sp.gdl : sp.dims[d];
        if (sp.periodic[d] && dc == lim - 1) dc = 1;
        if (dc > 1) moved_far = true;
    }
    // further than one cell: the tiled re-sort cannot find it — the step falls back to the generic re-index
    // (a slab window has no such fallback: device error)
    if (moved_far) { if (sp.W) atomicOr(err, AB_EF_OVERFLOW); else atomicOr(far, 1); }
    uint32_t nfrow = 0;
    bool inwin = true;
    uint32_t nk = device_key<T, D>(sp, np_, nc, &inwin, &nfrow);
    if (!inwin) { atomicOr(err, AB_EF_OVERFLOW); nk = 0; nfrow = 0; }
#pragma unroll
    for (int d = 0; d < D; ++d) { npos[(size_t)d * cap + s] = np_[d]; nvel[(size_t)d * cap + s] = nv[d]; }
    newkey[s] = nk;
    // arrival rank (used only when the step falls back to the generic re-index) + per-tile histogram of the tiled re-sort
    if (nfrow >= tg.nrows) return;                           // a row this handle does not re-sort (halo rows of a slab window)
    rank[s] = atomicAdd(&tg.cell_count[nk], 1u);
    count_tile(tg, nfrow, nk - nfrow * (uint32_t)sp.dims[0]);
}

// generic path, kernel 1: own force into fown, one-sided reactions into freact (atomics)
template <class T, int D>
__global__ void __launch_bounds__(128)
k_pf_generic(SpaceDev<T, D> sp, PfU<T> u, const T* __restrict__ pos, const T* __restrict__ rad, const T* __restrict__ kk,
             const uint32_t* __restrict__ key, const uint32_t* __restrict__ cs, int64_t cap, int n, T* __restrict__ fown,
             T* __restrict__ freact) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    T p[D], f[D];
    int c[D];
#pragma unroll
    for (int d = 0; d < D; ++d) { p[d] = pos[(size_t)d * cap + s]; f[d] = T(0); }
    const T ri = rad[s], ki = kk[s];
    key_to_cell<T, D>(sp, key[s], c);
    const int gr = grid_radius<T, D>(sp, p, c, u.cutoff);
    for_each_run<D>(sp.dims, sp.periodic, AB_METRIC_EUCLIDEAN, c, gr, cs, sp.kb, false, [&](uint32_t b, uint32_t e, int) {
        for (uint32_t j = b; j < e; ++j) {
            if ((int)j == s) continue;
            T o[D], dr[D];
            T dd = 0;
#pragma unroll
            for (int d = 0; d < D; ++d) {
                o[d] = pos[(size_t)d * cap + j];
                T direct = o[d] - p[d];
                if (sp.periodic[d]) direct = min_image(direct, sp.extent[d]);
                dr[d] = direct;
                const T sq = direct * direct;
                dd = (d == 0) ? sq : dd + sq;
            }
            if (!(dd <= u.thr)) continue;
            T fij[D];
#pragma unroll
            for (int d = 0; d < D; ++d) fij[d] = T(0);
            pf_pair<T, D>(dr, dd, ri, ki, rad[j], kk[j], fij);
#pragma unroll
            for (int d = 0; d < D; ++d) f[d] += fij[d];
            // does j see me in its own walk?  if not, it gets the reaction from here
            int cj[D];
            key_to_cell<T, D>(sp, key[j], cj);
            const int grj = grid_radius<T, D>(sp, o, cj, u.cutoff);
            if (!cell_reachable<D>(sp.dims, sp.periodic, AB_METRIC_EUCLIDEAN, cj, c, grj)) {
#pragma unroll
                for (int d = 0; d < D; ++d) if (fij[d] != T(0)) atomicAdd(&freact[(size_t)d * cap + j], -fij[d]);
            }
        }
    });
#pragma unroll
    for (int d = 0; d < D; ++d) fown[(size_t)d * cap + s] = f[d];
}

// generic path, kernel 2: integrate in place (forces in -> new pos / vel out, same slot)
template <class T, int D>
__global__ void k_pf_integrate(SpaceDev<T, D> sp, PfU<T> u, const T* __restrict__ pos, const T* __restrict__ vel,
                               const T* __restrict__ mass, int64_t cap, int n, T* __restrict__ a_pos /*fown -> npos*/,
                               T* __restrict__ a_vel /*freact -> nvel*/, const int32_t* __restrict__ ids, int* __restrict__ err) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    T p[D], v[D], f[D], np_[D], nv[D];
#pragma unroll
    for (int d = 0; d < D; ++d) {
        p[d] = pos[(size_t)d * cap + s]; v[d] = vel[(size_t)d * cap + s];
        f[d] = a_pos[(size_t)d * cap + s] + a_vel[(size_t)d * cap + s];
    }
    if (u.active && !u.active[ids[s]]) {                  // not scheduled: keeps its state
#pragma unroll
        for (int d = 0; d < D; ++d) { a_pos[(size_t)d * cap + s] = p[d]; a_vel[(size_t)d * cap + s] = v[d]; }
        return;
    }
    if (!pf_integrate<T, D>(sp, u, p, v, f, mass[s], np_, nv)) atomicOr(err, AB_EF_OUTSIDE);
#pragma unroll
    for (int d = 0; d < D; ++d) { a_pos[(size_t)d * cap + s] = np_[d]; a_vel[(size_t)d * cap + s] = nv[d]; }
}

// ---------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------
template <class T, int D>
static int pf_steps_t(ab_handle* h, const ab_pairforce_params* p, int64_t nsteps) {
    const int n = h->win_W ? (int)h->cap : (int)h->n, TB = 128;     // slab window: grid over the capacity, live count on the device
    PfU<T> u;
    u.dt = (T)p->dt;
    u.half_dt2 = T(0);
    u.cutoff = p->cutoff;
    u.thr = pf_threshold_host<T>(p->cutoff);
    u.active = h->step_mask_cur;
    const double s = (double)(T)h->desc.spacing;
    bool fine = p->cutoff > 0 && p->cutoff <= s && ab_resort_applicable(h);
    for (int d = 0; d < D; ++d) if (h->dims[d] < 3) fine = false;
    if (h->win_W && !fine) return ab_fail(h, AB_ERR_INVALID, "a sharded handle runs the fine pair-force step only (cutoff <= spacing, >= 3 cells per axis)");
    if (h->win_W && ab_comm_local(h)) return ab_fail(h, AB_ERR_STATE, "in-process slabs step the flock rule only (ab_step_flock_group)");
    if (fine) AB_TRY(ab_set_refine(h, 1));
    AB_TRY(ab_ensure_index(h));
    if (fine) AB_TRY(ab_resort_plan(h));
    SpaceDev<T, D> sp = make_space<T, D>(h);
    AB_CUDA(h, cudaEventRecord(h->ev0, h->stream));
    for (int64_t st = 0; st < nsteps; ++st) {
        const int cur = h->cur;
        const T* pos = (const T*)h->fields[0].buf[cur];
        const T* vel = (const T*)h->fields[1].buf[cur];
        const T* rad = (const T*)h->fields[p->field_r].buf[cur];
        const T* kk = (const T*)h->fields[p->field_k].buf[cur];
        const T* mass = (const T*)h->fields[p->field_mass].buf[cur];
        if (fine) {
            {
                AB_LAUNCH(h, "k_pf_fine");
                k_pf_fine<T, D><<<ab_blocks(n, TB), TB, 0, h->stream>>>(sp, u, pos, vel, rad, kk, mass, h->key[cur], h->cell_start, h->cap, n,
                                                                        (T*)h->tmp_pos, (T*)h->tmp_vel, h->newkey, h->rank, make_tilegeo(h), h->id[cur],
                                                                        h->d_err + 1, h->d_err, h->win_W ? h->cell_start + h->nfine : nullptr);
            }
            if (h->win_W) {
                // slab window: re-sort of the owned rows, then the boundary layers (with the r, k, mass columns) replace
                // the neighbours' halo layers — refresh and migration in one message, as in the flock step
                AB_TRY(ab_resort_tiled(h, h->tmp_pos, h->tmp_vel, false));
                AB_TRY(ab_comm_pack(h));
                AB_TRY(ab_comm_transfer(h));
                AB_TRY(ab_comm_unpack(h, false));
                continue;
            }
            // did anybody move further than one cell?  (then the tiled re-sort cannot find it)
            AB_CUDA(h, cudaMemcpyAsync(h->h_err + 1, h->d_err + 1, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
            AB_CUDA(h, cudaStreamSynchronize(h->stream));
            if (h->h_err[1]) {
                AB_CUDA(h, cudaMemsetAsync(h->d_err + 1, 0, sizeof(int), h->stream));
                // the per-tile / per-row histograms are not consumed by a tiled re-sort this time
                AB_CUDA(h, cudaMemsetAsync(h->tile_count, 0, 2 * h->tile_cap * sizeof(uint32_t), h->stream));
                if (h->row_count) AB_CUDA(h, cudaMemsetAsync(h->row_count, 0, 2 * h->row_cap * sizeof(uint32_t), h->stream));
                AB_TRY(ab_reindex_from(h, h->tmp_pos, h->tmp_vel, true));
            } else {
                AB_TRY(ab_resort_tiled(h, h->tmp_pos, h->tmp_vel, false));
            }
        } else {
            AB_CUDA(h, cudaMemsetAsync(h->tmp_vel, 0, (size_t)D * h->cap * sizeof(T), h->stream));
            {
                AB_LAUNCH(h, "k_pf_generic");
                k_pf_generic<T, D><<<ab_blocks(n, TB), TB, 0, h->stream>>>(sp, u, pos, rad, kk, h->key[cur], h->cell_start, h->cap, n,
                                                                           (T*)h->tmp_pos, (T*)h->tmp_vel);
            }
            {
                AB_LAUNCH(h, "k_pf_integrate");
                k_pf_integrate<T, D><<<ab_blocks(n, 256), 256, 0, h->stream>>>(sp, u, pos, vel, mass, h->cap, n, (T*)h->tmp_pos,
                                                                             (T*)h->tmp_vel, h->id[cur], h->d_err);
            }
            AB_TRY(ab_reindex_from(h, h->tmp_pos, h->tmp_vel, false));
        }
    }
    AB_CUDA(h, cudaEventRecord(h->ev1, h->stream));
    return AB_OK;
}

extern "C" int ab_step_pairforce(ab_handle* h, const ab_pairforce_params* p, int64_t nsteps);
// The pair-force step for a scheduled subset (Schedulers.Partially / an id list, schedulers.jl:68-120): model_step!
// computes the forces of every particle, agent_step! (the integrator, celllistmap.jl:176-191) runs for the scheduled ones.
extern "C" int ab_step_pairforce_active(ab_handle* h, const ab_pairforce_params* p, int64_t nsteps, const int64_t* active_ids,
                                        int64_t nactive) {
    if (!active_ids) return ab_step_pairforce(h, p, nsteps);
    if (h->n == 0 || nsteps == 0) return AB_OK;
    AB_TRY(ab_set_step_mask(h, active_ids, nactive));
    h->step_mask_cur = h->step_mask;
    const int rc = ab_step_pairforce(h, p, nsteps);
    h->step_mask_cur = nullptr;
    return rc;
}

extern "C" int ab_step_pairforce(ab_handle* h, const ab_pairforce_params* p, int64_t nsteps) {
    if (!p) return ab_fail(h, AB_ERR_INVALID, "null params");
    if (h->desc.kind != AB_SPACE_CONTINUOUS) return ab_fail(h, AB_ERR_INVALID, "pair-force rule needs a ContinuousSpace handle");
    if (nsteps < 0) return ab_fail(h, AB_ERR_INVALID, "negative nsteps");
    const int fs[3] = {p->field_r, p->field_k, p->field_mass};
    for (int f : fs)
        if (f < 2 || f >= (int)h->fields.size() || h->fields[f].ncomp != 1 || h->fields[f].dtype != h->dtype)
            return ab_fail(h, AB_ERR_INVALID, "r, k, mass must be registered 1-component columns of the space's float type");
    if (!(p->cutoff >= 0)) return ab_fail(h, AB_ERR_INVALID, "bad cutoff");
    AB_CUDA(h, cudaSetDevice(h->device));
    if (h->n == 0 || nsteps == 0) return AB_OK;
    int rc;
    if (h->dtype == AB_F64) rc = h->D == 2 ? pf_steps_t<double, 2>(h, p, nsteps) : pf_steps_t<double, 3>(h, p, nsteps);
    else rc = h->D == 2 ? pf_steps_t<float, 2>(h, p, nsteps) : pf_steps_t<float, 3>(h, p, nsteps);
    if (rc != AB_OK) return ab_step_failed(h, rc);
    if ((rc = ab_check_device_errors_now(h)) != AB_OK) return ab_step_state_lost(h, rc);   // synchronous call (also in asynchronous mode): the events below must have completed
    if (h->win_W) AB_TRY(ab_comm_refresh_counts(h));
    float ms = 0;
    AB_CUDA(h, cudaEventElapsedTime(&ms, h->ev0, h->ev1));
    h->last_step_ms = ms;
    return AB_OK;
}
