// ab_flock_fine.cuh — the throughput path of the fused flock step
// (examples/flock.jl:79-108) for large spaces with uniform rule parameters.
//
// One thread per bird, 32 consecutive slots of the cell-sorted state per warp.
// The state is sorted by FINE cell in fine-row-major order (ab_internal.cuh), so
// the short-range part of the rule reads a few contiguous slot runs straight
// from the sorted arrays (L1/L2) — no restaging, no shared memory, no barriers —
// and the lanes of a warp, being neighbours in space, walk (nearly) the same
// runs: in a crowded region every lane of a warp sits in the same fine cell and
// every candidate load is a single broadcast transaction.
//
//   `:approximate` search (the reference's default).  nearby_ids returns EVERY
//      agent of the cells of the offset table of radius grid_r around the
//      bird's cell (continuous.jl:221-230), so cohere = Σ heading and
//      match = Σ vel of flock.jl:88-91 are sums over whole cells:
//         cohere_i = M.C - M.N * (pos_i - centre_i),  match_i = M.V - vel_i,
//         N_i = M.N - 1,
//      with M = per-reference-cell sums over the offset tables of radius 1 and 2
//      (k_cell_M, built from the aggregates the re-sort leaves behind).  Only
//      `separate` (flock.jl:93-96, pairs closer than `separation`) needs a pair
//      sweep, over the fine cells fx-1..fx+1 of the fine rows fy-1..fy+1
//      (fine cell size >= separation).
//   `:exact` search.  The pair sweep covers the cells of the bird's offset
//      table (grid_r in {1,2}), clipped to the fine cells within K of the
//      bird's (fine cell size * K = spacing >= visual_distance), and applies
//      the reference's distance filter (continuous.jl:232-238).
// Birds whose grid_r is not 1 or 2 take the reference walk (flock_accumulate_generic),
// so the kernel is correct for every state; only its speed depends on the state.
#pragma once
#include <type_traits>

// launch shape of the sweep (measured on B200, F2D-1M: unroll 4 + 8 resident CTAs/SM is 15 % faster than
// unroll 2 without an occupancy hint; unroll 8 and 256-thread CTAs are slower)
#ifndef FINE_UNROLL
#define FINE_UNROLL 4
#endif
#ifndef FINE_TB
#define FINE_TB 128
#endif
#ifndef FINE_MINB
#define FINE_MINB 8
#endif
constexpr int kFineUnroll = FINE_UNROLL;
template <int D> struct FineCfg { static constexpr int NC = 1 + 2 * D; };   // count, Σ(pos-centre)[D], Σvel[D]

// row offsets (dy, dz) with dy² + dz² ≤ 4 (3-D) — summation order of k_cell_M
__constant__ int8_t c_ro3[13][2] = {{0, -2}, {-1, -1}, {0, -1}, {1, -1}, {-2, 0}, {-1, 0}, {0, 0},
                                    {1, 0},  {2, 0},   {-1, 1}, {0, 1},  {1, 1},  {0, 2}};
template <int D> __device__ __forceinline__ void ro_offset(int ro, int* dy, int* dz) {
    if (D == 2) { *dy = ro - 2; *dz = 0; }
    else { *dy = c_ro3[ro][0]; *dz = c_ro3[ro][1]; }
}

struct FinePlan {
    int K;
    int R;      // fine-cell reach of the `separation` sweep (approximate search): R * spacing / K >= separation
};

// wrap (periodic) or invalidate (-1) a cell coordinate
__device__ __forceinline__ int wrap_coord(int c, int n, bool periodic) {
    if (c >= 0 && c < n) return c;
    if (!periodic) return -1;
    return c < 0 ? c + n : c - n;
}

// M[g-1][c][cell] = Σ over the offset table of radius g (Σβ² ≤ g², grid_general.jl:138-151) of the
// per-cell aggregates, with neighbour positions re-expressed relative to THIS cell's centre
// (+ β·spacing per neighbour agent: Σ_nb count_nb·β_nb is accumulated as an exact integer).
// aggp holds one partial per fine row of a reference cell (written by k_resort / k_cell_agg);
// they are added in fixed order.
// A warp covers CM_OUT consecutive cells of one reference cell row plus 2 halo cells per side:
// each lane loads the aggregates of ONE cell per neighbour row and the ±1, ±2 x-neighbours
// come from warp shuffles, so every aggregate is read once per row offset, not once per offset.
// One aggregate per reference cell out of its 4 fine-row partials (3-D K = 2, 2-D K = 4), summed in the order
// k_cell_M would use — so that its 13 (5) neighbour-row reads per cell touch a quarter of the data.
template <class T>
__global__ void k_agg_collapse(const T* __restrict__ aggp, int64_t ncells, int64_t c0, int64_t nc, int ncomp, T* __restrict__ out) {
    ab_dep_wait();
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;       // (component, cell of [c0, c0 + nc))
    if (t >= (int64_t)ncomp * nc) return;
    const int64_t i = t / nc;
    const int64_t idx = i * ncells + c0 + (t - i * nc);
    const int64_t total = (int64_t)ncomp * ncells;
    T v = aggp[idx];
    v += aggp[total + idx];
    v += aggp[2 * total + idx];
    v += aggp[3 * total + idx];
    out[idx] = v;
}

#define CM_OUT 28
template <class T> __device__ __forceinline__ T shfl_T(T v, int src) { return __shfl_sync(0xffffffffu, v, src); }
template <class T, int D>
__global__ void __launch_bounds__(128)
k_cell_M(SpaceDev<T, D> sp, const T* __restrict__ aggp, int64_t ncells, int wpr, int nwarps, T* __restrict__ M,
         int skip0, int skiplen, int nfr) {
    constexpr int NC = FineCfg<D>::NC;
    constexpr int NRO = D == 2 ? 5 : 13;
    ab_dep_wait();
    const int warp = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    if (warp >= nwarps) return;
    const int lane = threadIdx.x & 31;
    // launched rows r map to cell rows r (r < skip0) or r + skiplen: all rows, the interior rows only, or all but them
    const int r_ = warp / wpr, seg = warp - r_ * wpr;
    const int row = r_ < skip0 ? r_ : r_ + skiplen;
    const int cy = D == 2 ? row : row % sp.dims[1];
    const int cz = D == 2 ? 0 : row / sp.dims[1];
    const int xout = seg * CM_OUT + lane - 2;                                  // this lane's output cell (lanes 2..29)
    const int xl = xout > sp.dims[0] + 1 ? -1 : wrap_coord(xout, sp.dims[0], sp.periodic[0]);   // cell whose aggregates this lane loads
    // nfr = partials per cell in aggp: K^(D-1) as written by the re-sort, or 1 after k_agg_collapse
    T acc1[NC], acc2[NC];
    int cb1[D], cb2[D];
#pragma unroll
    for (int i = 0; i < NC; ++i) acc1[i] = acc2[i] = T(0);
#pragma unroll
    for (int d = 0; d < D; ++d) cb1[d] = cb2[d] = 0;
#pragma unroll
    for (int ro = 0; ro < NRO; ++ro) {
        int dy, dz;
        ro_offset<D>(ro, &dy, &dz);
        const int rem2 = 4 - dy * dy - dz * dz;
        const int hw2 = rem2 >= 4 ? 2 : (rem2 >= 1 ? 1 : 0);
        const int rem1 = 1 - dy * dy - dz * dz;
        const int y = wrap_coord(cy + dy, sp.dims[1], sp.periodic[1]);
        const int z = D == 3 ? wrap_coord(cz + dz, sp.dims[2], sp.periodic[2]) : 0;
        T a[NC];
#pragma unroll
        for (int i = 0; i < NC; ++i) a[i] = T(0);
        if (y >= 0 && z >= 0 && xl >= 0) {                                       // warp-uniform except xl
            const size_t nb = (size_t)(D == 3 ? ((uint32_t)z * sp.dims[1] + y) : (uint32_t)y) * (uint32_t)sp.dims[0] + (uint32_t)xl;
#pragma unroll
            for (int i = 0; i < NC; ++i) {
                T v = aggp[(size_t)i * ncells + nb];
                if (nfr >= 2) v += aggp[((size_t)NC + i) * ncells + nb];
                if (nfr == 4) {
                    v += aggp[((size_t)2 * NC + i) * ncells + nb];
                    v += aggp[((size_t)3 * NC + i) * ncells + nb];
                }
                a[i] = v;
            }
        }
#pragma unroll
        for (int bx = -2; bx <= 2; ++bx) {
            if (bx < -hw2 || bx > hw2) continue;                                  // compile-time after unrolling
            T b[NC];
#pragma unroll
            for (int i = 0; i < NC; ++i) b[i] = bx == 0 ? a[i] : shfl_T(a[i], lane + bx);
            const int cnt = (int)b[0];
            const bool in1 = rem1 >= 0 && (bx == 0 || (rem1 >= 1 && (bx == 1 || bx == -1)));
#pragma unroll
            for (int i = 0; i < NC; ++i) { acc2[i] += b[i]; if (in1) acc1[i] += b[i]; }
#pragma unroll
            for (int d = 0; d < D; ++d) {
                const int beta = d == 0 ? bx : (d == 1 ? dy : dz);
                cb2[d] += cnt * beta;
                if (in1) cb1[d] += cnt * beta;
            }
        }
    }
    if (lane >= 2 && lane < 2 + CM_OUT && xout < sp.dims[0]) {
        const size_t cell = (size_t)(D == 3 ? ((uint32_t)cz * sp.dims[1] + cy) : (uint32_t)cy) * (uint32_t)sp.dims[0] + (uint32_t)xout;
#pragma unroll
        for (int d = 0; d < D; ++d) {
            acc1[1 + d] += (T)cb1[d] * sp.spacing;
            acc2[1 + d] += (T)cb2[d] * sp.spacing;
        }
#pragma unroll
        for (int i = 0; i < NC; ++i) {
            M[(size_t)i * ncells + cell] = acc1[i];
            M[((size_t)NC + i) * ncells + cell] = acc2[i];
        }
    }
}

// 2-D variant with a rolling window: a warp produces R consecutive cell rows of one 28-cell strip.  Every input row
// is loaded ONCE per warp; its 3-wide and 5-wide horizontal sums (H1, H2) are built once with four shuffles per
// component and serve all the output rows the input row belongs to:
//     M1(y) = (A(y-1) + H1(y)) + A(y+1)          M2(y) = (((A(y-2) + H1(y-1)) + H2(y)) + H1(y+1)) + A(y+2)
// with H1 = (a[x-1] + a[x]) + a[x+1], H2 = (a[x-2] + H1) + a[x+2].  The association is the same for every R, so
// launches with different R (whole grid: R = 4; row subsets of the overlapped sharded step: R = 1) agree bit for bit.
template <class T, int R>
__global__ void __launch_bounds__(128)
k_cell_M2(SpaceDev<T, 2> sp, const T* __restrict__ aggp, int64_t ncells, int wpr, int nwarps, T* __restrict__ M,
          int skip0, int skiplen, int nfr, int nrows_launch) {
    constexpr int NC = 5;
    ab_dep_wait();
    const int warp = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    if (warp >= nwarps) return;
    const int lane = threadIdx.x & 31;
    const int rb = warp / wpr, seg = warp - rb * wpr;
    const int r0_ = rb * R;                                           // first launched row of this warp
    const int row0 = r0_ < skip0 ? r0_ : r0_ + skiplen;               // R > 1 is only launched with skiplen == 0
    const int xout = seg * CM_OUT + lane - 2;
    const int xl = xout > sp.dims[0] + 1 ? -1 : wrap_coord(xout, sp.dims[0], sp.periodic[0]);
    T acc1[R][NC], acc2[R][NC];
    int c1x[R], c1y[R], c2x[R], c2y[R];
#pragma unroll
    for (int j = 0; j < R; ++j) {
        c1x[j] = c1y[j] = c2x[j] = c2y[j] = 0;
#pragma unroll
        for (int i = 0; i < NC; ++i) acc1[j][i] = acc2[j][i] = T(0);
    }
#pragma unroll
    for (int k = 0; k < R + 4; ++k) {
        const int y = wrap_coord(row0 - 2 + k, sp.dims[1], sp.periodic[1]);
        T a[NC], h1[NC], h2[NC];
#pragma unroll
        for (int i = 0; i < NC; ++i) a[i] = T(0);
        if (y >= 0 && xl >= 0) {
            const size_t nb = (size_t)(uint32_t)y * (uint32_t)sp.dims[0] + (uint32_t)xl;
#pragma unroll
            for (int i = 0; i < NC; ++i) {
                T v = aggp[(size_t)i * ncells + nb];
                if (nfr >= 2) v += aggp[((size_t)NC + i) * ncells + nb];
                if (nfr == 4) {
                    v += aggp[((size_t)2 * NC + i) * ncells + nb];
                    v += aggp[((size_t)3 * NC + i) * ncells + nb];
                }
                a[i] = v;
            }
        }
        int cl1, cr1, cl2, cr2;
#pragma unroll
        for (int i = 0; i < NC; ++i) {
            const T l1 = shfl_T(a[i], lane - 1), r1 = shfl_T(a[i], lane + 1);
            const T l2 = shfl_T(a[i], lane - 2), r2 = shfl_T(a[i], lane + 2);
            h1[i] = (l1 + a[i]) + r1;
            h2[i] = (l2 + h1[i]) + r2;
            if (i == 0) { cl1 = (int)l1; cr1 = (int)r1; cl2 = (int)l2; cr2 = (int)r2; }
        }
        const int ca = (int)a[0], n1 = cl1 + ca + cr1;
        const int cx1 = cr1 - cl1, cx2 = cx1 + 2 * (cr2 - cl2);
#pragma unroll
        for (int j = 0; j < R; ++j) {
            const int dy = k - 2 - j;                                  // input row relative to output row j (compile time)
            if (dy == -2 || dy == 2) {
#pragma unroll
                for (int i = 0; i < NC; ++i) acc2[j][i] += a[i];
                c2y[j] += dy * ca;
            } else if (dy == -1 || dy == 1) {
#pragma unroll
                for (int i = 0; i < NC; ++i) { acc2[j][i] += h1[i]; acc1[j][i] += a[i]; }
                c2x[j] += cx1; c2y[j] += dy * n1; c1y[j] += dy * ca;
            } else if (dy == 0) {
#pragma unroll
                for (int i = 0; i < NC; ++i) { acc2[j][i] += h2[i]; acc1[j][i] += h1[i]; }
                c2x[j] += cx2; c1x[j] += cx1;
            }
        }
    }
    if (lane >= 2 && lane < 2 + CM_OUT && xout < sp.dims[0]) {
#pragma unroll
        for (int j = 0; j < R; ++j) {
            if (r0_ + j >= nrows_launch) break;
            const int cy = row0 + j;
            const size_t cell = (size_t)(uint32_t)cy * (uint32_t)sp.dims[0] + (uint32_t)xout;
            acc1[j][1] += (T)c1x[j] * sp.spacing; acc1[j][2] += (T)c1y[j] * sp.spacing;
            acc2[j][1] += (T)c2x[j] * sp.spacing; acc2[j][2] += (T)c2y[j] * sp.spacing;
#pragma unroll
            for (int i = 0; i < NC; ++i) {
                M[(size_t)i * ncells + cell] = acc1[j][i];
                M[((size_t)NC + i) * ncells + cell] = acc2[j][i];
            }
        }
    }
}

// grid_r (continuous.jl:221-225) when the plan guarantees it is 1 or 2: ceil((r + δ)/spacing) is 1 iff
// δ <= spacing - r and 2 iff δ <= 2 spacing - r, so δ² is compared with the squared thresholds; a bird
// whose δ² lies within 1e-9 (relative) of a threshold takes the reference arithmetic instead.
template <class T, int D>
__device__ __forceinline__ int grid_radius_12(const SpaceDev<T, D>& sp, const T* p, const int* c0, const FlockU<T>& fp) {
    double s2 = 0;
#pragma unroll
    for (int d = 0; d < D; ++d) {
        const double centre = sp.spacing_d * ((double)(c0[d] + 1) - 0.5);
        // c0 is the bird's own cell: direct <= spacing / 2 (+ rounding) and the plan guarantees >= 6 cells per axis,
        // so min(direct, extent - direct) of continuous.jl:119-121 is `direct`, bit for bit
        const double m = fabs((double)p[d] - centre);
        const double sq = m * m;
        s2 = (d == 0) ? sq : s2 + sq;
    }
    const double lo = 1.0 - 1e-9, hi = 1.0 + 1e-9;
    if (s2 < fp.gr_t1 * lo) return 1;
    if (s2 > fp.gr_t1 * hi && s2 < fp.gr_t2 * lo) return 2;
    return grid_radius<T, D>(sp, p, c0, fp.visual_distance);
}

// pos2cell0 with the division replaced by a multiplication whenever the quotient is provably not
// within rounding distance of an integer (then ceil of either is the same integer).
template <class T, int D>
__device__ __forceinline__ bool pos2cell0_fast(const SpaceDev<T, D>& sp, const T* p, int* c) {
    const T eps8 = sizeof(T) == 8 ? T(1.7763568394002505e-15) : T(9.5367431640625e-07);   // 8 ulp
    bool fastok = true;
#pragma unroll
    for (int d = 0; d < D; ++d) {
        const T qa = p[d] * sp.inv_spacing;
        const T cq = ceil(qa);
        const T margin = qa * eps8;
        // (cq - qa) in (margin, 1 - margin): not near an integer; 0 < cq <= dims
        const int lim = d == D - 1 ? sp.gdl : sp.dims[d];
        if (!(cq - qa > margin && cq - qa < T(1) - margin && cq >= T(1) && cq <= (T)lim)) fastok = false;
        c[d] = (int)cq - 1;
    }
    if (fastok) return true;
    return pos2cell0<T, D>(sp, p, c);
}

// REACH2: quarter cells (K = 4) with a reach-2 separation sweep — a separate instantiation, so that the common
// reach-1 sweep keeps its compile-time trip counts (a run-time reach cost the 3-D sweep 13 %).
template <class T, int D, bool EXACT, bool REACH2 = false>
__global__ void __launch_bounds__(FINE_TB, FINE_MINB)
k_flock_fine(const __grid_constant__ SpaceDev<T, D> sp, const __grid_constant__ FlockU<T> fp, const T* __restrict__ pos, const T* __restrict__ vel,
             const uint32_t* __restrict__ key, const uint32_t* __restrict__ cs, const T* __restrict__ M,
             int64_t ncells, int64_t cap, int n, T* __restrict__ npos, T* __restrict__ nvel,
             uint32_t* __restrict__ newkey, TileGeo tg, const int32_t* __restrict__ ids,
             int32_t* __restrict__ ncount, int* __restrict__ err, const uint32_t* __restrict__ nslots, int phase,
             uint32_t own_keys) {
    constexpr int NC = FineCfg<D>::NC;
    ab_dep_wait();
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    // slab window: the slot count lives on the device (migration changes it without a host round trip).
    // phase 1 (interior layers, swept while the halo of this state may still be arriving): owned slots only —
    // the halo slots and the halo part of cell_start are not valid yet.
    if (s >= (phase == 1 ? (int)cs[own_keys] : (nslots ? (int)*nslots : n))) return;
    const int K = sp.K, kb = sp.kb;
    int f[D], c[D], cg[D];
    const uint32_t k = key[s];
    key_to_fine_fast<T, D>(sp, k, f);
#pragma unroll
    for (int d = 0; d < D; ++d) cg[d] = c[d] = f[d] >> kb;
    if (sp.W) {
        // owned layers plus the first halo layer on each side (recomputed here instead of being
        // received: its birds may enter the owned layers in this step); outer halo slots are skipped
        const int l = c[D - 1];
        if (!(l <= sp.nl || l == sp.nl + 2 * sp.W - 1)) return;
        // interior = layers whose birds read no halo data: aggregates within 2 layers, pair sweep within 1
        const bool interior = l >= 2 && l < sp.nl - 2;
        if ((phase == 1 && !interior) || (phase == 2 && interior)) return;
        cg[D - 1] = local_to_global_last<T, D>(sp, l);
    }
    T p[D], v[D];
#pragma unroll
    for (int d = 0; d < D; ++d) { p[d] = pos[(size_t)d * cap + s]; v[d] = vel[(size_t)d * cap + s]; }
    if (fp.active && !fp.active[ids[s]]) {
        // not scheduled this tick: the agent keeps its state and stays a neighbour of the others
#pragma unroll
        for (int d = 0; d < D; ++d) { npos[(size_t)d * cap + s] = p[d]; nvel[(size_t)d * cap + s] = v[d]; }
        newkey[s] = k;
        count_new_key(tg, k, D == 3 ? (uint32_t)f[D - 1] * (uint32_t)(sp.dims[1] << kb) + (uint32_t)f[1] : (uint32_t)f[1], (uint32_t)f[0]);
        if (ncount) ncount[ids[s]] = 0;
        return;
    }
    const int gr = grid_radius_12<T, D>(sp, p, cg, fp);
    const T sep2 = fp.separation * fp.separation;
    T cohere[D], separate[D], match[D];
#pragma unroll
    for (int d = 0; d < D; ++d) cohere[d] = separate[d] = match[d] = T(0);
    int N = 0;
    bool fast = gr == 1 || gr == 2;
    T q[D];
#pragma unroll
    for (int d = 0; d < D; ++d) q[d] = p[d] - sp.spacing * ((T)(cg[d] + 1) - T(0.5));
    if (EXACT && fast) {
        // birds within tau of a fine-cell boundary could see (in floating point) a neighbour that the
        // clipped sweep below does not visit: they take the reference walk.
        const T eps = sizeof(T) == 8 ? T(1.4210854715202004e-14) : T(7.62939453125e-06);     // 2^-46, 2^-17
#pragma unroll
        for (int d = 0; d < D; ++d) {
            const T tau = sp.extent[d] * eps;
            const T a = fabs(q[d]);
            if (fabs(a - sp.spacing * T(0.5)) < tau || (K == 2 && a < tau)) fast = false;
        }
    }
    if (fast) {
        if (!EXACT) {
            const uint32_t refc = lin_cell<D>(sp.dims, c);
            const T* Mg = M + (size_t)(gr - 1) * NC * ncells + refc;
            const T Nt = Mg[0];
            N = (int)Nt - 1;
#pragma unroll
            for (int d = 0; d < D; ++d) {
                cohere[d] = Mg[(size_t)(1 + d) * ncells] - Nt * q[d];
                match[d] = Mg[(size_t)(1 + D + d) * ncells] - v[d];
            }
        }
        const T thr = fp.thr;
        const int R = EXACT ? K : (REACH2 ? 2 : 1);    // fine-cell reach of the sweep
        const int fd0 = sp.dims[0] << kb, fd1 = sp.dims[1] << kb, fd2 = D == 3 ? (sp.dims[2] << kb) : 1;
        // slab window: rows are indexed in the local (window-periodic) frame, but whether a row lies across
        // the space's periodic boundary — i.e. needs the minimum image — is decided on GLOBAL fine rows
        const int gf_last = (cg[D - 1] << kb) | (f[D - 1] & (K - 1)), gfdl = sp.gdl << kb;
        // two instantiations of the pair loop: runs reached across a periodic boundary apply the
        // reference's minimum-image rule (utilities.jl:185-193); all other runs are plain differences
        auto sweep_t = [&](uint32_t jb, uint32_t je, int wrap, auto wrapped) {
            constexpr bool WRAP = decltype(wrapped)::value;
#pragma unroll kFineUnroll
            for (uint32_t j = jb; j < je; ++j) {
                T h[D];
                T d2 = 0;
#pragma unroll
                for (int d = 0; d < D; ++d) {
                    T direct = pos[(size_t)d * cap + j] - p[d];
                    if (WRAP && (wrap & (1 << d))) direct = min_image(direct, sp.extent[d]);
                    h[d] = direct;
                    const T sq = direct * direct;
                    d2 = (d == 0) ? sq : d2 + sq;
                }
                if (EXACT) {
                    if ((int)j == s || !(d2 <= thr)) continue;                            // continuous.jl:237-238
                    ++N;
#pragma unroll
                    for (int d = 0; d < D; ++d) { cohere[d] += h[d]; match[d] += vel[(size_t)d * cap + j]; }
                }
                // the bird itself (h == 0) passes this test and contributes exactly nothing.
                // separate -= h  as  fma(-w, h, separate), w = 1 or 0: one rounding, the same value as the subtraction
                // (w * h is exact), without the select pairs a predicated update compiles to.  With a non-finite h
                // (a NaN / Inf position) 0 * h is NaN where the predicated form skipped the bird; such a step has
                // already raised AB_EF_NUMERIC and its result is discarded (AB_ERR_NUMERIC), so nothing visible differs
                const T wgt = d2 < sep2 ? T(1) : T(0);                                   // flock.jl:93
#pragma unroll
                for (int d = 0; d < D; ++d) separate[d] = fma(-wgt, h[d], separate[d]);
            }
        };
        auto sweep = [&](uint32_t jb, uint32_t je, int wrap) {
            if (wrap == 0) sweep_t(jb, je, 0, std::false_type{});
            else sweep_t(jb, je, wrap, std::true_type{});
        };
        constexpr int NROW = (EXACT || REACH2) ? 5 : 3; // fine rows per z-slab: 2R+1
        // REACH2 (2-D quarter cells): every row is trimmed to the cells that can hold a partner closer than
        // `separation` — a superset of the disc, from the bird's offset inside its quarter cell in f32 with a margin
        float tq_h = 0.f, tq_x = 0.f, tq_y = 0.f, tq_s = 0.f;
        if (REACH2) {
            tq_h = (float)sp.spacing * 0.25f;
            tq_x = (float)(p[0] - (sp.spacing * T(0.25)) * (T)f[0]);
            tq_y = (float)(p[D - 1] - (sp.spacing * T(0.25)) * (T)gf_last);
            tq_s = (float)fp.separation * 1.00001f + 1e-4f * tq_h;
        }
        for (int rz = (D == 3 ? -R : 0); rz <= (D == 3 ? R : 0); ++rz) {
            int fz = 0, wz = 0, dz = 0;
            if (D == 3) {
                fz = f[2] + rz;
                dz = (fz >> kb) - c[2];            // arithmetic shift: floor for fz = -1, -2
                if (fz < 0 || fz >= fd2) {
                    if (!sp.periodic[2]) continue;
                    fz = fz < 0 ? fz + fd2 : fz - fd2;
                    wz = 4;
                }
                if (sp.W) { const int g = gf_last + rz; wz = (g < 0 || g >= gfdl) ? 4 : 0; }
            }
            // phase 1: slot bounds of the main run of every row — independent loads, all in flight together
            uint32_t rbv[NROW], jb[NROW], je[NROW];
            int ulov[NROW], uhiv[NROW], wrv[NROW];
#pragma unroll
            for (int i = 0; i < NROW; ++i) {
                wrv[i] = -1;                                 // row not visited
                jb[i] = je[i] = 0;
                const int ry = i - R;
                if (i > 2 * R) continue;
                int fy = f[1] + ry;
                const int dy = (fy >> kb) - c[1];
                int ulo = f[0] - R, uhi = f[0] + R;          // inclusive fine cells, unwrapped
                if (REACH2) {
                    const float dym = ry > 0 ? (float)ry * tq_h - tq_y : (ry < 0 ? tq_y + (float)(-ry - 1) * tq_h : 0.f);
                    if (dym >= tq_s) continue;               // the whole row is farther than `separation`
                    const float wq = sqrtf(tq_s * tq_s - fmaxf(dym, 0.f) * fmaxf(dym, 0.f));
                    ulo = f[0] + max((int)floorf((tq_x - wq) / tq_h), -R);
                    uhi = f[0] + min((int)floorf((tq_x + wq) / tq_h), R);
                }
                if (EXACT) {
                    // only cells of the bird's offset table (Σβ² ≤ grid_r²)
                    const int rem2 = gr * gr - dy * dy - dz * dz;
                    if (rem2 < 0) continue;
                    const int hw = rem2 >= 4 ? 2 : (rem2 >= 1 ? 1 : 0);
                    ulo = max(ulo, (c[0] - hw) * K);
                    uhi = min(uhi, (c[0] + hw + 1) * K - 1);
                }
                int wy = 0;
                if (fy < 0 || fy >= fd1) {
                    if (!sp.periodic[1]) continue;
                    fy = fy < 0 ? fy + fd1 : fy - fd1;
                    wy = 2;
                }
                if (D == 2 && sp.W) { const int g = gf_last + ry; wy = (g < 0 || g >= gfdl) ? 2 : 0; }
                const uint32_t rb = (D == 3 ? ((uint32_t)fz * (uint32_t)fd1 + (uint32_t)fy) : (uint32_t)fy) * (uint32_t)fd0;
                rbv[i] = rb; ulov[i] = ulo; uhiv[i] = uhi; wrv[i] = wy | wz;
                const int a0 = max(ulo, 0), a1 = min(uhi, fd0 - 1);
                if (a0 <= a1) { jb[i] = cs[rb + (uint32_t)a0]; je[i] = cs[rb + (uint32_t)a1 + 1]; }
            }
            // phase 2: the pair loops
#pragma unroll
            for (int i = 0; i < NROW; ++i) {
                if (wrv[i] < 0) continue;
                sweep(jb[i], je[i], wrv[i]);
                if (sp.periodic[0]) {                        // parts reached across the x boundary (rare)
                    if (ulov[i] < 0) sweep(cs[rbv[i] + (uint32_t)(ulov[i] + fd0)], cs[rbv[i] + (uint32_t)fd0], wrv[i] | 1);
                    if (uhiv[i] >= fd0) sweep(cs[rbv[i]], cs[rbv[i] + (uint32_t)(uhiv[i] - fd0) + 1], wrv[i] | 1);
                }
            }
        }
    } else {
        flock_accumulate_generic<T, D, EXACT>(sp, fp, s, p, c, gr, pos, vel, cs, cap, &N, cohere, separate, match);
    }
    T np_[D], nv[D];
    flock_epilogue<T, D>(sp, p, v, cohere, separate, match, N, fp, np_, nv, err);
    int nc[D];
    if (!pos2cell0_fast<T, D>(sp, np_, nc)) atomicOr(err, AB_EF_NUMERIC);
    bool inwin = true;
    uint32_t nfrow = 0;
    uint32_t nk = device_key<T, D>(sp, np_, nc, &inwin, &nfrow);
    if (!inwin) { atomicOr(err, AB_EF_OVERFLOW); nk = 0; nfrow = 0; }
#pragma unroll
    for (int d = 0; d < D; ++d) { npos[(size_t)d * cap + s] = np_[d]; nvel[(size_t)d * cap + s] = nv[d]; }
    newkey[s] = nk;
    // histogram of the new keys per fine cell and per re-sort tile (the re-sort turns them into the next cell_start)
    count_new_key(tg, nk, nfrow, nk - nfrow * (uint32_t)(sp.dims[0] << kb));
    if (ncount) ncount[ids[s]] = N;
}

// Can the fine-grid step (sweep + tiled re-sort) run this configuration?
template <class T, int D>
static bool plan_fine(ab_handle* h, const ab_flock_params* p, bool exact, FinePlan* plan) {
    if (h->desc.kind != AB_SPACE_CONTINUOUS || p->per_agent_field >= 0 || h->n < 1) return false;
    const double s = (double)(T)h->desc.spacing, r = p->visual_distance, sep = p->separation, margin = 1.0 + 1e-9;
    for (int d = 0; d < D; ++d) if (h->dims[d] < 6) return false;     // radius-2 offset table unclipped (grid_general.jl:49)
    if (!(r > 0) || !(sep >= 0) || !(p->speed >= 0)) return false;
    if (!((r + s * sqrt((double)D) * 0.5) * margin <= 2.0 * s)) return false;   // grid_r ∈ {1, 2} for every bird
    if (exact && !(r <= s)) return false;                              // ±1 reference cell covers the radius
    int K = 0, R = 1;
    // fine cell >= separation / R (±R fine cells cover it) and the per-step displacement stays below one fine cell
    // (the re-sort collects from the 3^D neighbouring fine cells).  2-D approximate search: quarter cells with
    // reach 2 — 25 cells of (s/4)² instead of 9 of (s/2)², 31 % fewer candidates for the pair sweep.
    const bool fits4 = D == 2 && !exact && (h->ncells << 4) <= h->nfine_cap && h->fine4;
    if (fits4 && sep * margin <= s * 0.5 && sep * margin > s * 0.25 && p->speed * margin < s * 0.25) { K = 4; R = 2; }
    else if (sep * margin <= s * 0.5 && p->speed * margin < s * 0.5 && (h->ncells << D) <= h->nfine_cap) K = 2;
    else if (sep * margin <= s && p->speed * margin < s) K = 1;
    if (!K) return false;
    if (!ab_resort_applicable(h)) return false;
    plan->K = K;
    plan->R = R;
    return true;
}
