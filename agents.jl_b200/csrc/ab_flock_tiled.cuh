// ab_flock_tiled.cuh — the throughput path of the fused flock step
// (examples/flock.jl:79-108) for large spaces with uniform rule parameters.
//
// CTA = SWT consecutive slots of the cell-sorted state (= a run of reference
// cells of one or two cell rows, a "piece" each).  Per CTA:
//   1. A-strips: per-reference-cell aggregates (count, Σ(pos-centre), Σvel —
//      written by the re-sort, ab_resort.cu) of the rows within 2 cells of the
//      piece are loaded into shared memory.
//   2. M: for every reference cell of the piece, the sums of those aggregates
//      over the reference's offset tables of radius 1 and 2
//      (offsets_within_radius, src/spaces/grid_general.jl:98-151; Σβ² ≤ g²).
//      In `:approximate` search nearby_ids returns EVERY agent of those cells
//      (continuous.jl:221-230), so cohere = Σ heading and match = Σ vel of
//      flock.jl:88-91 are cell sums:  cohere_i = M.C - M.N·(pos_i - centre_i),
//      match_i = M.V - vel_i,  N_i = M.N - 1.
//   3. candidates: positions (and, for `:exact`, velocities) of the agents of
//      the 3^D neighbouring reference cells are staged in shared memory in
//      sub-cell-row-major order, so that the sub-cells an agent must test for
//      `separation` (flock.jl:93-96) — or for the exact distance filter
//      (continuous.jl:232-238) — are a few contiguous runs.
//   4. one thread per bird: separation / exact sweep over its runs, the
//      epilogue of the generic kernel (same arithmetic), new device key.
// CTAs whose slot range does not fit this shape (sparse rows, crowded cells)
// and birds whose grid_r is not 1 or 2 run the generic walk instead, so the
// kernel is correct for every state; only its speed depends on the shape.
#pragma once

template <int D> struct TiledCfg;
template <> struct TiledCfg<2> {
    static constexpr int SWT = 256, WTOT = 136, NQ = 3, NRO = 5, NC = 5, CAPC = 1280, FOFF = 1664;
};
template <> struct TiledCfg<3> {
    static constexpr int SWT = 128, WTOT = 72, NQ = 9, NRO = 13, NC = 7, CAPC = 1536, FOFF = 5120;
};

// row offsets (dy, dz) with dy² + dz² ≤ 4, and their inverse table
__constant__ int8_t c_ro3[13][2] = {{0, -2}, {-1, -1}, {0, -1}, {1, -1}, {-2, 0}, {-1, 0}, {0, 0},
                                    {1, 0},  {2, 0},   {-1, 1}, {0, 1},  {1, 1},  {0, 2}};

template <int D> __device__ __forceinline__ void ro_offset(int ro, int* dy, int* dz) {
    if (D == 2) { *dy = ro - 2; *dz = 0; }
    else { *dy = c_ro3[ro][0]; *dz = c_ro3[ro][1]; }
}

struct TiledPlan {
    int K;
    size_t smem;
};

template <class T, int D, bool EXACT> struct TiledSmem {
    using C = TiledCfg<D>;
    static constexpr size_t M_BYTES = sizeof(T) * 2 * C::NC * C::WTOT;
    static constexpr size_t A_BYTES = sizeof(T) * C::NRO * C::NC * C::WTOT;
    static constexpr size_t CAND_BYTES = sizeof(T) * C::CAPC * D * (EXACT ? 2 : 1);
    static constexpr size_t F_BYTES = sizeof(uint32_t) * C::FOFF;
    static constexpr size_t X_BYTES = A_BYTES > (CAND_BYTES + F_BYTES) ? A_BYTES : (CAND_BYTES + F_BYTES);
    static constexpr size_t TOTAL = M_BYTES + X_BYTES;
};

// wrap (periodic) or invalidate (-1) a cell coordinate
__device__ __forceinline__ int wrap_coord(int c, int n, bool periodic) {
    if (c >= 0 && c < n) return c;
    if (!periodic) return -1;
    return c < 0 ? c + n : c - n;
}

template <class T, int D, bool EXACT>
__global__ void __launch_bounds__(TiledCfg<D>::SWT)
k_flock_tiled(SpaceDev<T, D> sp, FlockU<T> fp, const T* __restrict__ pos, const T* __restrict__ vel,
              const uint32_t* __restrict__ key, const uint32_t* __restrict__ cs, const T* __restrict__ agg,
              int64_t agg_stride, int64_t cap, int n, T* __restrict__ npos, T* __restrict__ nvel,
              uint32_t* __restrict__ newkey, const int32_t* __restrict__ ids, int32_t* __restrict__ ncount,
              int* __restrict__ err) {
    using C = TiledCfg<D>;
    using SM = TiledSmem<T, D, EXACT>;
    constexpr int SWT = C::SWT, NQ = C::NQ, NRO = C::NRO, NC = C::NC, WTOT = C::WTOT;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    T* sM = reinterpret_cast<T*>(smem_raw);                          // [2][NC][WTOT]
    T* sA = reinterpret_cast<T*>(smem_raw + SM::M_BYTES);            // [NRO][NC][WTOT]     (phase 1-2)
    uint32_t* sF = reinterpret_cast<uint32_t*>(smem_raw + SM::M_BYTES);                  // (phase 3-4)
    T* sC = reinterpret_cast<T*>(smem_raw + SM::M_BYTES + SM::F_BYTES);                  // candidates, AoS
    __shared__ uint32_t s_warp[SWT / 32];
    __shared__ uint32_t s_total;

    const int tid = threadIdx.x;
    const int s0 = blockIdx.x * SWT;
    const int cnt = min(SWT, n - s0);
    const int K = sp.K, ksh = sp.ksh, KR = (D == 2) ? K : K * K;
    const int dims0 = sp.dims[0];

    // ---- pieces ------------------------------------------------------------------------
    const uint32_t rf = key[s0] >> ksh, rl = key[s0 + cnt - 1] >> ksh;
    const int rowf = (int)(rf / (uint32_t)dims0), rowl = (int)(rl / (uint32_t)dims0);
    const int np = rowl - rowf + 1;
    int xa[2], w[2], prow[2];
    prow[0] = rowf; prow[1] = rowl;
    xa[0] = (int)(rf - (uint32_t)rowf * (uint32_t)dims0);
    const int xl = (int)(rl - (uint32_t)rowl * (uint32_t)dims0);
    w[0] = (np == 1 ? xl : dims0 - 1) - xa[0] + 1;
    xa[1] = 0;
    w[1] = np == 2 ? xl + 1 : 0;
    bool tiled = np <= 2 && (w[0] + 4 + (np == 2 ? w[1] + 4 : 0)) <= WTOT;
    const int aoff[2] = {0, w[0] + 4};       // A-strip x offset of each piece
    const int moff[2] = {0, w[0]};           // M x offset of each piece
    // flat sub-cell-offset table: line (p, q, t) has L_p + 1 entries
    const int L[2] = {K * (w[0] + 2), K * (w[1] + 2)};
    const int fbase1 = NQ * KR * (L[0] + 1);
    const int nf = fbase1 + (np == 2 ? NQ * KR * (L[1] + 1) : 0);
    tiled = tiled && nf <= C::FOFF;

    if (tiled && !EXACT) {
        // ---- 1. A-strips ------------------------------------------------------------------
        for (int p = 0; p < np; ++p) {
            const int wa = w[p] + 4;
            const int cy = D == 2 ? prow[p] : prow[p] % sp.dims[1];
            const int cz = D == 2 ? 0 : prow[p] / sp.dims[1];
            for (int i = tid; i < NRO * wa; i += SWT) {
                const int ro = i / wa, xi = i - ro * wa;
                int dy, dz;
                ro_offset<D>(ro, &dy, &dz);
                const int x = wrap_coord(xa[p] - 2 + xi, dims0, sp.periodic[0]);
                const int y = wrap_coord(cy + dy, sp.dims[1], sp.periodic[1]);
                const int z = D == 3 ? wrap_coord(cz + dz, sp.dims[2], sp.periodic[2]) : 0;
                T* dst = sA + (size_t)ro * NC * WTOT + aoff[p] + xi;
                if (x < 0 || y < 0 || z < 0) {
#pragma unroll
                    for (int c = 0; c < NC; ++c) dst[c * WTOT] = T(0);
                } else {
                    const uint32_t cell = (D == 3 ? ((uint32_t)z * sp.dims[1] + y) : (uint32_t)y) * (uint32_t)dims0 + x;
                    dst[0] = (T)(cs[(size_t)(cell + 1) << ksh] - cs[(size_t)cell << ksh]);
#pragma unroll
                    for (int c = 1; c < NC; ++c) dst[c * WTOT] = agg[(size_t)(c - 1) * agg_stride + cell];
                }
            }
        }
        __syncthreads();
        // ---- 2. M = offset-table sums, radius 1 and 2 -------------------------------------
        {
            const int wsum = w[0] + w[1];
            for (int i = tid; i < 2 * NC * wsum; i += SWT) {
                const int g = i / (NC * wsum) + 1;
                const int rem = i - (g - 1) * NC * wsum;
                const int c = rem / wsum, xm = rem - c * wsum;
                const int p = xm >= w[0] ? 1 : 0;
                const int x = xm - moff[p];                 // cell of the piece; A index = aoff + x + 2 + bx
                T acc = T(0);
                for (int ro = 0; ro < NRO; ++ro) {
                    int dy, dz;
                    ro_offset<D>(ro, &dy, &dz);
                    const int rem2 = g * g - dy * dy - dz * dz;
                    if (rem2 < 0) continue;
                    const int hw = rem2 >= 4 ? 2 : (rem2 >= 1 ? 1 : 0);
                    const T* row = sA + (size_t)ro * NC * WTOT + aoff[p] + x + 2;
                    for (int bx = -hw; bx <= hw; ++bx) {
                        if (c == 0 || c > D) acc += row[c * WTOT + bx];
                        else {
                            const int beta = c == 1 ? bx : (c == 2 ? dy : dz);
                            acc += row[c * WTOT + bx] + row[bx] * ((T)beta * sp.spacing);
                        }
                    }
                }
                sM[((size_t)(g - 1) * NC + c) * WTOT + xm] = acc;
            }
        }
        __syncthreads();
    }
    if (tiled) {
        // ---- 3. sub-cell offsets of the candidate strips ------------------------------------
        for (int i = tid; i < nf; i += SWT) {
            const int p = i >= fbase1 ? 1 : 0;
            const int j = i - (p ? fbase1 : 0);
            const int line = j / (L[p] + 1), u = j - line * (L[p] + 1);
            uint32_t c = 0;
            if (u < L[p]) {
                const int q = line / KR, t = line - q * KR;
                const int qy = q % 3, qz = q / 3;
                const int cy = D == 2 ? prow[p] : prow[p] % sp.dims[1];
                const int cz = D == 2 ? 0 : prow[p] / sp.dims[1];
                const int x = wrap_coord(xa[p] - 1 + u / K, dims0, sp.periodic[0]);
                const int y = wrap_coord(cy + qy - 1, sp.dims[1], sp.periodic[1]);
                const int z = D == 3 ? wrap_coord(cz + qz - 1, sp.dims[2], sp.periodic[2]) : 0;
                if (x >= 0 && y >= 0 && z >= 0) {
                    const uint32_t cell = (D == 3 ? ((uint32_t)z * sp.dims[1] + y) : (uint32_t)y) * (uint32_t)dims0 + x;
                    const uint32_t sub = K == 2 ? (uint32_t)((u & 1) | (t << 1)) : 0u;
                    const size_t k = ((size_t)cell << ksh) | sub;
                    c = cs[k + 1] - cs[k];
                }
            }
            sF[i] = c;
        }
        __syncthreads();
        {   // block exclusive scan of sF[0..nf)
            const int per = (nf + SWT - 1) / SWT;
            const int b = tid * per, e = min(nf, b + per);
            uint32_t tsum = 0;
            for (int i = b; i < e; ++i) tsum += sF[i];
            const int lane = tid & 31, wid = tid >> 5;
            uint32_t inc = tsum;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { uint32_t t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += t; }
            if (lane == 31) s_warp[wid] = inc;
            __syncthreads();
            uint32_t woff = 0, total = 0;
#pragma unroll
            for (int ww = 0; ww < SWT / 32; ++ww) { uint32_t t = s_warp[ww]; if (ww < wid) woff += t; total += t; }
            uint32_t run = woff + inc - tsum;
            for (int i = b; i < e; ++i) { uint32_t t = sF[i]; sF[i] = run; run += t; }
            if (tid == 0) s_total = total;
            __syncthreads();
        }
        tiled = s_total <= (uint32_t)C::CAPC;
    }
    if (tiled) {
        // ---- 3b. stage candidates (coalesced reads of the sorted arrays) ----------------------
        for (int p = 0; p < np; ++p) {
            const int cy = D == 2 ? prow[p] : prow[p] % sp.dims[1];
            const int cz = D == 2 ? 0 : prow[p] / sp.dims[1];
            for (int q = 0; q < NQ; ++q) {
                const int qy = q % 3, qz = q / 3;
                const int y = wrap_coord(cy + qy - 1, sp.dims[1], sp.periodic[1]);
                const int z = D == 3 ? wrap_coord(cz + qz - 1, sp.dims[2], sp.periodic[2]) : 0;
                if (y < 0 || z < 0) continue;
                const uint32_t rb = (D == 3 ? ((uint32_t)z * sp.dims[1] + y) : (uint32_t)y) * (uint32_t)dims0;
                const int lbase = (p ? fbase1 : 0) + q * KR * (L[p] + 1);
                // up to three runs of real cells: wrapped left neighbour, main range, wrapped right neighbour
                for (int part = 0; part < 3; ++part) {
                    int ra, rb2, ux0;              // real cell range [ra, rb2], unwrapped x of ra
                    if (part == 0) {
                        if (!(xa[p] == 0 && sp.periodic[0])) continue;
                        ra = rb2 = dims0 - 1; ux0 = -1;
                    } else if (part == 1) {
                        ra = max(xa[p] - 1, 0); rb2 = min(xa[p] + w[p], dims0 - 1); ux0 = ra;
                    } else {
                        if (!(xa[p] + w[p] == dims0 && sp.periodic[0])) continue;
                        ra = rb2 = 0; ux0 = dims0;
                    }
                    const uint32_t sb = cs[(size_t)(rb + ra) << ksh], se = cs[(size_t)(rb + rb2 + 1) << ksh];
                    for (uint32_t s = sb + tid; s < se; s += SWT) {
                        const uint32_t k = key[s];
                        const int rx = (int)((k >> ksh) - rb);
                        const int ux = ux0 + (rx - ra);
                        const uint32_t sub = k & ((1u << ksh) - 1u);
                        const int u = K * (ux - (xa[p] - 1)) + (K == 2 ? (int)(sub & 1u) : 0);
                        const int t = K == 2 ? (int)(sub >> 1) : 0;
                        const uint32_t dsti = sF[lbase + t * (L[p] + 1) + u] + (s - cs[k]);
                        T* dst = sC + (size_t)dsti * (EXACT ? 2 * D : D);
#pragma unroll
                        for (int d = 0; d < D; ++d) dst[d] = pos[(size_t)d * cap + s];
                        if (EXACT) {
#pragma unroll
                            for (int d = 0; d < D; ++d) dst[D + d] = vel[(size_t)d * cap + s];
                        }
                    }
                }
            }
        }
        __syncthreads();
    }

    // ---- 4. one thread per bird ---------------------------------------------------------------
    const int s = s0 + tid;
    if (tid >= cnt) return;
    T p[D], v[D];
    int c[D];
#pragma unroll
    for (int d = 0; d < D; ++d) { p[d] = pos[(size_t)d * cap + s]; v[d] = vel[(size_t)d * cap + s]; }
    const uint32_t k = key[s];
    const uint32_t refc = k >> ksh;
    unlin_cell<D>(sp.dims, refc, c);
    const int gr = grid_radius<T, D>(sp, p, c, fp.visual_distance);
    const T sep2 = fp.separation * fp.separation;
    T cohere[D], separate[D], match[D];
#pragma unroll
    for (int d = 0; d < D; ++d) cohere[d] = separate[d] = match[d] = T(0);
    int N = 0;
    bool fast = tiled && (gr == 1 || gr == 2);
    T q[D];
#pragma unroll
    for (int d = 0; d < D; ++d) q[d] = p[d] - sp.spacing * ((T)(c[d] + 1) - T(0.5));
    if (EXACT && fast) {
        // birds within tau of a sub-cell boundary could see (in floating point) a neighbour that the
        // block sweep below does not visit: they take the reference walk.
        const T eps = sizeof(T) == 8 ? T(1.4210854715202004e-14) : T(7.62939453125e-06);     // 2^-46, 2^-17
#pragma unroll
        for (int d = 0; d < D; ++d) {
            const T tau = sp.extent[d] * eps;
            const T a = fabs(q[d]);
            if (fabs(a - sp.spacing * T(0.5)) < tau || (K == 2 && a < tau)) fast = false;
        }
    }
    if (fast) {
        const int row = (int)(refc / (uint32_t)dims0);
        const int pi = (np == 2 && row == rowl) ? 1 : 0;
        const int xin = c[0] - xa[pi];
        const int sx = K == 2 ? (int)(k & 1u) : 0, sy = K == 2 ? (int)((k >> 1) & 1u) : 0, sz = (K == 2 && D == 3) ? (int)((k >> 2) & 1u) : 0;
        const int pbase = pi ? fbase1 : 0;
        const int Lp1 = L[pi] + 1;
        int wrapx = (sp.periodic[0] && (c[0] == 0 || c[0] == dims0 - 1)) ? 1 : 0;
        if (!EXACT) {
            const T* Mg = sM + (size_t)(gr - 1) * NC * WTOT + moff[pi] + xin;
            const T Nt = Mg[0];
            N = (int)Nt - 1;
#pragma unroll
            for (int d = 0; d < D; ++d) {
                cohere[d] = Mg[(1 + d) * WTOT] - Nt * q[d];
                match[d] = Mg[(1 + D + d) * WTOT] - v[d];
            }
        }
        const int R = EXACT ? K : 1;                   // sub-cell reach of the sweep
        uint32_t selfj = 0xffffffffu;
        if (EXACT) {
            const int line = (1 + (D == 3 ? 3 : 0)) * KR + (K == 2 ? (sy | (sz << 1)) : 0);     // q = centre row, t = own sub-row
            selfj = sF[pbase + line * Lp1 + K * (xin + 1) + sx] + ((uint32_t)s - cs[k]);
        }
        const T thr = EXACT ? exact_threshold<T>(fp.visual_distance) : T(0);
        for (int rz = (D == 3 ? -R : 0); rz <= (D == 3 ? R : 0); ++rz) {
            const int fz = K + sz + rz;                 // sub-row index inside the 3 stacked strips (0 .. 3K-1)
            const int qz = D == 3 ? fz / K : 0, tz = D == 3 ? fz - qz * K : 0;
            const int wz = (D == 3 && sp.periodic[2] && (c[2] + qz - 1 < 0 || c[2] + qz - 1 >= sp.dims[2])) ? 4 : 0;
            for (int ry = -R; ry <= R; ++ry) {
                const int fy = K + sy + ry;
                const int qy = fy / K, ty = fy - qy * K;
                int ulo = K * (xin + 1) + sx - R, uhi = ulo + 2 * R;          // inclusive sub-cells
                if (EXACT) {
                    // only cells of the bird's offset table (Σβ² ≤ grid_r²)
                    const int dy = qy - 1, dz = D == 3 ? qz - 1 : 0;
                    const int rem2 = gr * gr - dy * dy - dz * dz;
                    if (rem2 < 0) continue;
                    const int hw = rem2 >= 4 ? 2 : (rem2 >= 1 ? 1 : 0);
                    ulo = max(ulo, K * (xin + 1 - hw));
                    uhi = min(uhi, K * (xin + 1 + hw) + K - 1);
                }
                const int wy = (sp.periodic[1] && (c[1] + qy - 1 < 0 || c[1] + qy - 1 >= sp.dims[1])) ? 2 : 0;
                const int wrap = wrapx | wy | wz;
                const int line = (qy + 3 * qz) * KR + (ty + K * tz);
                const uint32_t* f = sF + pbase + line * Lp1;
                const uint32_t jb = f[ulo], je = f[uhi + 1];
                for (uint32_t j = jb; j < je; ++j) {
                    const T* cj = sC + (size_t)j * (EXACT ? 2 * D : D);
                    T h[D];
                    T d2 = 0;
#pragma unroll
                    for (int d = 0; d < D; ++d) {
                        T direct = cj[d] - p[d];
                        if (wrap & (1 << d)) direct = min_image(direct, sp.extent[d]);     // utilities.jl:185-193
                        h[d] = direct;
                        const T sq = direct * direct;
                        d2 = (d == 0) ? sq : d2 + sq;
                    }
                    if (EXACT) {
                        if (j == selfj || !(d2 <= thr)) continue;                             // continuous.jl:237-238
                        ++N;
#pragma unroll
                        for (int d = 0; d < D; ++d) { cohere[d] += h[d]; match[d] += cj[D + d]; }
                    }
                    if (d2 < sep2) {                                                         // flock.jl:93
#pragma unroll
                        for (int d = 0; d < D; ++d) separate[d] -= h[d];
                    }
                }
            }
        }
    } else {
        flock_accumulate_generic<T, D, EXACT>(sp, fp, s, p, c, gr, pos, vel, cs, cap, &N, cohere, separate, match);
    }
    T np_[D], nv[D];
    flock_epilogue<T, D>(sp, p, v, cohere, separate, match, N, fp, np_, nv, err);
    int nc[D];
    if (!pos2cell0<T, D>(sp, np_, nc)) atomicOr(err, AB_EF_NUMERIC);
    const uint32_t nk = device_key<T, D>(sp, np_, nc);
#pragma unroll
    for (int d = 0; d < D; ++d) { npos[(size_t)d * cap + s] = np_[d]; nvel[(size_t)d * cap + s] = nv[d]; }
    newkey[s] = nk;
    if (ncount) ncount[ids[s]] = N;
}

// Can the tiled step (sweep + tiled re-sort) run this configuration?
template <class T, int D>
static bool plan_tiled(ab_handle* h, const ab_flock_params* p, bool exact, TiledPlan* plan) {
    if (h->desc.kind != AB_SPACE_CONTINUOUS || p->per_agent_field >= 0 || h->n < 1) return false;
    const double s = (double)(T)h->desc.spacing, r = p->visual_distance, sep = p->separation, margin = 1.0 + 1e-9;
    for (int d = 0; d < D; ++d) if (h->dims[d] < 6) return false;     // radius-2 offset table unclipped (grid_general.jl:49)
    if (!(r > 0) || !(sep >= 0) || !(p->speed >= 0)) return false;
    if (!(p->speed * margin < s)) return false;                        // re-sort: less than one reference cell per step
    if (!((r + s * sqrt((double)D) * 0.5) * margin <= 2.0 * s)) return false;   // grid_r ∈ {1, 2} for every bird
    if (exact && !(r <= s)) return false;                              // ±1 reference cell covers the radius
    int K = 0;
    if (sep * margin <= s * 0.5 && (h->ncells << D) <= h->nfine_cap) K = 2;
    else if (sep * margin <= s) K = 1;
    if (!K) return false;
    if (!ab_resort_applicable(h)) return false;
    plan->K = K;
    return true;
}
