// ab_grid.cu — GridSpace / GridSpaceSingle: index build, neighbour queries, walk!, and the
// Schelling rule.
//
//   GridSpace{D,P}        src/spaces/grid_multi.jl:4-12, 68-105   (many agents per position)
//   GridSpaceSingle{D,P}  src/spaces/grid_single.jl:11-57         (at most one agent per position)
//   nearby_ids            grid_multi.jl:120-232, grid_single.jl:92-205; offset tables
//                         grid_general.jl:47-151 under :chebyshev / :manhattan / :euclidean
//   walk!                 src/spaces/walk.jl:33-56, 112-126
//   Schelling rule        docs/src/tutorial.jl:351-376 (synchronous formulation)
//
// Device layout.  Agent columns of a grid handle stay in id order (slot = id - 1): `pos` is D x i32,
// 1-based like Julia's.  GridSpace keeps a cell list (cell_start + ids grouped by cell, ascending id
// inside a cell); GridSpaceSingle keeps the dense array grid_ids[cell] (0 = empty) — the device
// form of stored_ids::Array{Int,D} (grid_single.jl:12).  The Schelling step works on cell-major
// mirrors of the agent columns (grid_group, grid_mood): the stencil then reads 4 B per cell and
// writes 1 B per cell, all coalesced, and a relocated agent is three scattered 4-byte stores.
#include <climits>

#include "ab_internal.cuh"

// ---------------------------------------------------------------------------
// small device helpers
// ---------------------------------------------------------------------------
template <int D> struct GridDev {
    int dims[D];
    bool periodic[D];
    int metric;
};
template <int D> static GridDev<D> make_grid(const ab_handle* h) {
    GridDev<D> g;
    for (int d = 0; d < D; ++d) { g.dims[d] = h->dims[d]; g.periodic[d] = h->desc.periodic[d] != 0; }
    g.metric = h->desc.metric;
    return g;
}

// pos (1-based, id order) -> 0-based cell; false when out of bounds (BoundsError in the reference)
template <int D>
__device__ __forceinline__ bool load_cell(const GridDev<D>& g, const int32_t* __restrict__ pos, int64_t cap, int i, int* c) {
    bool ok = true;
#pragma unroll
    for (int d = 0; d < D; ++d) {
        c[d] = pos[(size_t)d * cap + i] - 1;
        if (c[d] < 0 || c[d] >= g.dims[d]) { ok = false; c[d] = 0; }
    }
    return ok;
}

// ---------------------------------------------------------------------------
// index build
// ---------------------------------------------------------------------------
template <int D>
__global__ void k_gcount(GridDev<D> g, const int32_t* __restrict__ pos, int64_t cap, int n, uint32_t* __restrict__ newkey,
                         uint32_t* __restrict__ rank, uint32_t* __restrict__ cell_count, int* __restrict__ err) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int c[D];
    if (!load_cell<D>(g, pos, cap, i, c)) atomicOr(err, AB_EF_OUTSIDE);
    const uint32_t k = lin_cell<D>(g.dims, c);
    newkey[i] = k;
    rank[i] = atomicAdd(&cell_count[k], 1u);
}
__global__ void k_gbucket(const uint32_t* __restrict__ newkey, const uint32_t* __restrict__ rank,
                          const uint32_t* __restrict__ cell_start, int n, int32_t* __restrict__ idtmp) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) idtmp[cell_start[newkey[i]] + rank[i]] = i;
}
// canonical order inside a cell: ascending id (the reference's order is insertion order, i.e. the
// same for a bulk add; results are compared as sets anyway)
__global__ void k_gplace(const uint32_t* __restrict__ newkey, const uint32_t* __restrict__ cell_start,
                         const int32_t* __restrict__ idtmp, int n, int32_t* __restrict__ sorted_ids) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t k = newkey[i];
    const uint32_t b = cell_start[k], e = cell_start[k + 1];
    uint32_t r = 0;
    for (uint32_t t = b; t < e; ++t) r += (idtmp[t] < i) ? 1u : 0u;
    sorted_ids[b + r] = i;
}
template <int D>
__global__ void k_gsingle_build(GridDev<D> g, const int32_t* __restrict__ pos, int64_t cap, int n,
                                int32_t* __restrict__ grid_ids, int* __restrict__ err) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int c[D];
    if (!load_cell<D>(g, pos, cap, i, c)) { atomicOr(err, AB_EF_OUTSIDE); return; }
    const uint32_t k = lin_cell<D>(g.dims, c);
    if (atomicCAS(&grid_ids[k], 0, i + 1) != 0) atomicOr(err, AB_EF_OCCUPIED);      // grid_single.jl:47-52
}

template <int D>
static int grid_build_t(ab_handle* h) {
    const int n = (int)h->n, TB = 256;
    GridDev<D> g = make_grid<D>(h);
    const int32_t* pos = (const int32_t*)h->fields[0].buf[h->cur];
    if (h->desc.kind == AB_SPACE_GRID_SINGLE) {
        AB_CUDA(h, cudaMemsetAsync(h->grid_ids, 0, (size_t)h->ncells * sizeof(int32_t), h->stream));
        if (n > 0) {
            AB_LAUNCH(h, "k_gsingle_build");
            k_gsingle_build<D><<<ab_blocks(n, TB), TB, 0, h->stream>>>(g, pos, h->cap, n, h->grid_ids, h->d_err);
        }
    } else {
        if (n > 0) {
            AB_LAUNCH(h, "k_gcount");
            k_gcount<D><<<ab_blocks(n, TB), TB, 0, h->stream>>>(g, pos, h->cap, n, h->newkey, h->rank, h->cell_count, h->d_err);
        }
        AB_TRY(ab_run_scan(h, h->cell_start));
        if (n > 0) {
            { AB_LAUNCH(h, "k_gbucket"); k_gbucket<<<ab_blocks(n, TB), TB, 0, h->stream>>>(h->newkey, h->rank, h->cell_start, n, h->idtmp); }
            { AB_LAUNCH(h, "k_gplace"); k_gplace<<<ab_blocks(n, TB), TB, 0, h->stream>>>(h->newkey, h->cell_start, h->idtmp, n, h->perm); }
        }
    }
    h->index_valid = true;
    h->grid_cells_valid = false;
    int rc = ab_check_device_errors(h);
    if (rc == AB_ERR_OUTSIDE) return ab_fail(h, AB_ERR_OUTSIDE, "BoundsError: position outside the grid");
    return rc;
}

int ab_grid_build(ab_handle* h) {
    if (h->grid_cols_dirty) AB_TRY(ab_grid_sync_columns(h));
    return h->D == 2 ? grid_build_t<2>(h) : grid_build_t<3>(h);
}
static int grid_ensure_index(ab_handle* h) {
    if (h->index_valid) return AB_OK;
    return ab_grid_build(h);
}

// ---------------------------------------------------------------------------
// nearby_ids — grid_multi.jl:120-232, grid_single.jl:92-205
//   position form: every id stored in the cells of offsets_within_radius(space, r) around `origin`
//                  (origin's own cell included);
//   agent form   : GridSpace filters id != agent.id (space_interaction_API.jl:284-287);
//                  GridSpaceSingle walks the table without the zero offset (grid_single.jl:201-205).
// ---------------------------------------------------------------------------
template <int D, bool SINGLE, bool FILL>
__global__ void k_gnearby(GridDev<D> g, const int32_t* __restrict__ qpos, const int64_t* __restrict__ qids, int nq,
                          int r0, int agent_form, const int32_t* __restrict__ pos, int64_t cap, int n,
                          const uint32_t* __restrict__ cell_start, const int32_t* __restrict__ sorted_ids,
                          const int32_t* __restrict__ grid_ids, int64_t* __restrict__ counts,
                          const int64_t* __restrict__ offsets, int64_t* __restrict__ out, int* __restrict__ err) {
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= nq) return;
    int c[D];
    int self = -1;
    bool ok = true;
    if (qids) {
        const int64_t id = qids[q];
        if (id < 1 || id > n) { atomicOr(err, AB_EF_OVERFLOW); ok = false; }
        else { self = (int)id - 1; ok = load_cell<D>(g, pos, cap, self, c); }
    } else {
#pragma unroll
        for (int d = 0; d < D; ++d) {
            c[d] = qpos[(size_t)q * D + d] - 1;
            if (c[d] < 0 || c[d] >= g.dims[d]) ok = false;
        }
        if (!ok) atomicOr(err, AB_EF_OUTSIDE);
    }
    if (!ok) { if (!FILL) counts[q] = 0; return; }
    if (!agent_form) self = -1;
    const bool skip_zero = SINGLE && agent_form;
    int64_t k = 0;
    const int64_t base = FILL ? offsets[q] : 0;
    auto visit = [&](uint32_t b, uint32_t e, int) {
        for (uint32_t j = b; j < e; ++j) {
            int id0;
            if (SINGLE) { id0 = grid_ids[j] - 1; if (id0 < 0) continue; }       // j = linear cell
            else { id0 = sorted_ids[j]; if (id0 == self) continue; }             // j = slot of the cell list
            if (FILL) out[base + k] = (int64_t)id0 + 1;
            ++k;
        }
    };
    if (SINGLE) for_each_run_cs<D>(g.dims, g.periodic, g.metric, c, r0, CellIdentity{}, 0, skip_zero, visit);
    else for_each_run_cs<D>(g.dims, g.periodic, g.metric, c, r0, CellStartPtr{cell_start}, 0, false, visit);
    if (!FILL) counts[q] = k;
}

int ab_grid_nearby(ab_handle* h, const void* qpos, const int64_t* qids, int64_t nq, double r, int exclude_self,
                   int64_t* counts, const int64_t* offsets, int64_t* ids_out) {
    if (nq == 0) return AB_OK;
    if (h->grid_cols_dirty) AB_TRY(ab_grid_sync_columns(h));
    AB_TRY(grid_ensure_index(h));
    // offsets_within_radius: r0 = floor(Int, r + 1) - 1 (grid_general.jl:98-110)
    double fr = floor(r + 1.0) - 1.0;
    if (!(fr < 1.0e9)) fr = 1.0e9;
    if (!(fr > -1.0e9)) fr = -1.0e9;
    const int r0 = (int)fr;
    const bool fill = ids_out != nullptr;
    const size_t qbytes = qpos ? (size_t)nq * h->D * sizeof(int32_t) : (size_t)nq * sizeof(int64_t);
    const size_t o1 = (qbytes + 255) / 256 * 256;
    const size_t cbytes = (size_t)(nq + 1) * sizeof(int64_t);
    const size_t o2 = o1 + (cbytes + 255) / 256 * 256;
    int64_t total = 0;
    if (fill) {
        if (!offsets) return ab_fail(h, AB_ERR_INVALID, "null offsets");
        total = offsets[nq];
        if (total < 0) return ab_fail(h, AB_ERR_INVALID, "bad offsets");
    }
    char* base;
    AB_TRY(ab_scratch(h, o2 + (size_t)total * sizeof(int64_t) + 256, (void**)&base));
    AB_CUDA(h, cudaMemcpyAsync(base, qpos ? qpos : (const void*)qids, qbytes, cudaMemcpyHostToDevice, h->stream));
    int64_t* cdev = (int64_t*)(base + o1);
    int64_t* odev = (int64_t*)(base + o2);
    if (fill) AB_CUDA(h, cudaMemcpyAsync(cdev, offsets, cbytes, cudaMemcpyHostToDevice, h->stream));
    const int32_t* qp = qpos ? (const int32_t*)base : nullptr;
    const int64_t* qi = qids ? (const int64_t*)base : nullptr;
    const int agent_form = (qids && exclude_self) ? 1 : 0;
    const bool single = h->desc.kind == AB_SPACE_GRID_SINGLE;
    const int32_t* pos = (const int32_t*)h->fields[0].buf[h->cur];
    {
        AB_LAUNCH(h, fill ? "k_gnearby_fill" : "k_gnearby_count");
        const int TB = 128, nb = ab_blocks(nq, TB);
#define GN(DD, S, F) k_gnearby<DD, S, F><<<nb, TB, 0, h->stream>>>(make_grid<DD>(h), qp, qi, (int)nq, r0, agent_form, pos, h->cap, (int)h->n, \
        h->cell_start, h->perm, h->grid_ids, F ? nullptr : cdev, F ? cdev : nullptr, F ? odev : nullptr, h->d_err)
        if (h->D == 2) { if (single) { if (fill) GN(2, true, true); else GN(2, true, false); } else { if (fill) GN(2, false, true); else GN(2, false, false); } }
        else { if (single) { if (fill) GN(3, true, true); else GN(3, true, false); } else { if (fill) GN(3, false, true); else GN(3, false, false); } }
#undef GN
    }
    if (fill) { if (total > 0) AB_CUDA(h, cudaMemcpyAsync(ids_out, odev, (size_t)total * sizeof(int64_t), cudaMemcpyDeviceToHost, h->stream)); }
    else AB_CUDA(h, cudaMemcpyAsync(counts, cdev, (size_t)nq * sizeof(int64_t), cudaMemcpyDeviceToHost, h->stream));
    int e = ab_check_device_errors_now(h);               // hands back host data: synchronous in every mode
    h->index_valid = true;                               // a bad query does not invalidate the index
    if (e == AB_ERR_INVALID) return ab_fail(h, AB_ERR_INVALID, "agent id out of range");
    if (e == AB_ERR_OUTSIDE) return ab_fail(h, AB_ERR_OUTSIDE, "BoundsError: query position outside the grid");
    return e;
}

// nearby_positions(pos, model, r) — grid_general.jl:185-232: the cells of offsets_within_radius_no_0 around
// `pos` (wrapped with mod1 on periodic axes, skipped outside the grid otherwise; the origin itself excluded);
// empty_nearby_positions — discrete.jl:116-129: those of them that hold no agent.
template <int D, bool SINGLE, bool FILL>
__global__ void k_gpositions(GridDev<D> g, const int32_t* __restrict__ qpos, int nq, int r0, int empty_only,
                             const uint32_t* __restrict__ cell_start, const int32_t* __restrict__ grid_ids,
                             int64_t* __restrict__ counts, const int64_t* __restrict__ offsets, int32_t* __restrict__ out,
                             int* __restrict__ err) {
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= nq) return;
    int c[D];
    bool ok = true;
#pragma unroll
    for (int d = 0; d < D; ++d) {
        c[d] = qpos[(size_t)q * D + d] - 1;
        if (c[d] < 0 || c[d] >= g.dims[d]) ok = false;
    }
    if (!ok) { atomicOr(err, AB_EF_OUTSIDE); if (!FILL) counts[q] = 0; return; }
    int64_t k = 0;
    const int64_t base = FILL ? offsets[q] : 0;
    auto visit = [&](uint32_t b, uint32_t e, int) {
        for (uint32_t j = b; j < e; ++j) {                       // j = linear (column-major) cell
            if (empty_only) {
                const bool occupied = SINGLE ? grid_ids[j] != 0 : cell_start[j + 1] != cell_start[j];
                if (occupied) continue;
            }
            if (FILL) {
                uint32_t rest = j;
#pragma unroll
                for (int d = 0; d < D; ++d) {
                    const uint32_t qd = rest / (uint32_t)g.dims[d];
                    out[(size_t)(base + k) * D + d] = (int32_t)(rest - qd * (uint32_t)g.dims[d]) + 1;
                    rest = qd;
                }
            }
            ++k;
        }
    };
    for_each_run_cs<D>(g.dims, g.periodic, g.metric, c, r0, CellIdentity{}, 0, true, visit);
    if (!FILL) counts[q] = k;
}

int ab_grid_positions(ab_handle* h, const int32_t* qpos, int64_t nq, double r, int empty_only, int64_t* counts,
                      const int64_t* offsets, int32_t* pos_out) {
    if (nq == 0) return AB_OK;
    if (!qpos) return ab_fail(h, AB_ERR_INVALID, "null query positions");
    if (h->grid_cols_dirty) AB_TRY(ab_grid_sync_columns(h));
    if (empty_only) AB_TRY(grid_ensure_index(h));
    double fr = floor(r + 1.0) - 1.0;                     // offsets_within_radius: r0 = floor(Int, r + 1) - 1
    if (!(fr < 1.0e9)) fr = 1.0e9;
    if (!(fr > -1.0e9)) fr = -1.0e9;
    const int r0 = (int)fr;
    const bool fill = pos_out != nullptr;
    const size_t qbytes = (size_t)nq * h->D * sizeof(int32_t);
    const size_t o1 = (qbytes + 255) / 256 * 256;
    const size_t cbytes = (size_t)(nq + 1) * sizeof(int64_t);
    const size_t o2 = o1 + (cbytes + 255) / 256 * 256;
    int64_t total = 0;
    if (fill) {
        if (!offsets) return ab_fail(h, AB_ERR_INVALID, "null offsets");
        total = offsets[nq];
        if (total < 0) return ab_fail(h, AB_ERR_INVALID, "bad offsets");
    } else if (!counts) return ab_fail(h, AB_ERR_INVALID, "null counts");
    char* base;
    AB_TRY(ab_scratch(h, o2 + (size_t)total * h->D * sizeof(int32_t) + 256, (void**)&base));
    AB_CUDA(h, cudaMemcpyAsync(base, qpos, qbytes, cudaMemcpyHostToDevice, h->stream));
    int64_t* cdev = (int64_t*)(base + o1);
    int32_t* odev = (int32_t*)(base + o2);
    if (fill) AB_CUDA(h, cudaMemcpyAsync(cdev, offsets, cbytes, cudaMemcpyHostToDevice, h->stream));
    const bool single = h->desc.kind == AB_SPACE_GRID_SINGLE;
    {
        AB_LAUNCH(h, fill ? "k_gpositions_fill" : "k_gpositions_count");
        const int TB = 128, nb = ab_blocks(nq, TB);
#define GP(DD, S, F) k_gpositions<DD, S, F><<<nb, TB, 0, h->stream>>>(make_grid<DD>(h), (const int32_t*)base, (int)nq, r0, empty_only, \
        h->cell_start, h->grid_ids, F ? nullptr : cdev, F ? cdev : nullptr, F ? odev : nullptr, h->d_err)
        if (h->D == 2) { if (single) { if (fill) GP(2, true, true); else GP(2, true, false); } else { if (fill) GP(2, false, true); else GP(2, false, false); } }
        else { if (single) { if (fill) GP(3, true, true); else GP(3, true, false); } else { if (fill) GP(3, false, true); else GP(3, false, false); } }
#undef GP
    }
    if (fill) { if (total > 0) AB_CUDA(h, cudaMemcpyAsync(pos_out, odev, (size_t)total * h->D * sizeof(int32_t), cudaMemcpyDeviceToHost, h->stream)); }
    else AB_CUDA(h, cudaMemcpyAsync(counts, cdev, (size_t)nq * sizeof(int64_t), cudaMemcpyDeviceToHost, h->stream));
    const bool was_valid = h->index_valid;
    int e = ab_check_device_errors_now(h);
    h->index_valid = was_valid;                          // a bad query does not invalidate the index
    if (e == AB_ERR_OUTSIDE) return ab_fail(h, AB_ERR_OUTSIDE, "BoundsError: query position outside the grid");
    return e;
}

// ---------------------------------------------------------------------------
// walk!(agent, direction, model; ifempty) — walk.jl:33-56; normalize_position walk.jl:112-126
// (periodic: mod1, else clamp).  GridSpaceSingle always behaves as ifempty = true (walk.jl:46-56).
// ---------------------------------------------------------------------------
__device__ __forceinline__ int grid_norm(int p1 /*1-based*/, int n, bool periodic) {
    if (periodic) return jl_mod1(p1, n);
    return p1 < 1 ? 1 : (p1 > n ? n : p1);
}

// All agents move (GridSpace, ifempty = false): no agent's move depends on another's.
template <int D>
__global__ void k_gwalk_all(GridDev<D> g, int32_t* __restrict__ pos, int64_t cap, int n, const int64_t* __restrict__ ids,
                            int nactive, const int32_t* __restrict__ dirs, int* __restrict__ err) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= nactive) return;
    int i = t;
    if (ids) {
        const int64_t id = ids[t];
        if (id < 1 || id > n) { atomicOr(err, AB_EF_OVERFLOW); return; }
        i = (int)id - 1;
    }
#pragma unroll
    for (int d = 0; d < D; ++d)
        pos[(size_t)d * cap + i] = grid_norm(pos[(size_t)d * cap + i] + dirs[(size_t)t * D + d], g.dims[d], g.periodic[d]);
}

// ifempty = true: whether a target is empty depends on the walks applied before it, so the
// reference's result is defined by its sequential order.  One thread replays that order on the
// occupancy array (exactly the reference's semantics; bulk conflict-free grid movement is a
// "next" item of SURVEY §8f, not a hot kernel).
template <int D, bool SINGLE>
__global__ void k_gwalk_seq(GridDev<D> g, int32_t* __restrict__ pos, int64_t cap, int n, const int64_t* __restrict__ ids,
                            int nactive, const int32_t* __restrict__ dirs, int32_t* __restrict__ occ /*ids or counts*/,
                            int* __restrict__ err) {
    if (blockIdx.x != 0 || threadIdx.x != 0) return;
    for (int t = 0; t < nactive; ++t) {
        int i = t;
        if (ids) {
            const int64_t id = ids[t];
            if (id < 1 || id > n) { atomicOr(err, AB_EF_OVERFLOW); continue; }
            i = (int)id - 1;
        }
        int c0[D], c1[D];
#pragma unroll
        for (int d = 0; d < D; ++d) {
            c0[d] = pos[(size_t)d * cap + i] - 1;
            c1[d] = grid_norm(c0[d] + 1 + dirs[(size_t)t * D + d], g.dims[d], g.periodic[d]) - 1;
        }
        const uint32_t k0 = lin_cell<D>(g.dims, c0), k1 = lin_cell<D>(g.dims, c1);
        if (occ[k1] != 0) continue;                       // target not empty (the agent's own cell counts)
        if (SINGLE) { occ[k0] = 0; occ[k1] = i + 1; }
        else { occ[k0] -= 1; occ[k1] += 1; }
#pragma unroll
        for (int d = 0; d < D; ++d) pos[(size_t)d * cap + i] = c1[d] + 1;
    }
}
__global__ void k_copy_counts(const uint32_t* __restrict__ cs, int64_t ncells, int32_t* __restrict__ occ) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < ncells) occ[i] = (int32_t)(cs[i + 1] - cs[i]);
}

// randomwalk!(agent, model::ABM{<:AbstractGridSpace}, r; ifempty, force_motion) — src/spaces/walk.jl:154-214.
// Without force_motion the step is one uniformly chosen entry of offsets_at_radius(model, r)
// (grid_general.jl:119-136), then the ordinary walk!: the directions are drawn in parallel (draw 0 of the
// agent's counter stream) and handed to the walk kernels above.  With force_motion the choice is made among
// the EMPTY targets (random_empty_pos_in_offsets, walk.jl:200-214): up to 2L random probes (draws 0..2L-1),
// then one uniform pick (draw 2L) among the empty targets in table order; the agent stays when there is none.
// Emptiness depends on the walks before it, so that form runs in the sequential replay like walk!(ifempty).
// The reference hands the chosen POSITION to walk! as if it were a direction (walk.jl:170, :193).  Both readings are
// offered: force_motion = 1 is the documented behaviour — the agent moves to the chosen empty position —
// force_motion = 2 the code as written (target = normalize_position(pos .+ choice), taken only if empty).
struct GridRw {
    const int32_t* offs;     // L x D offsets at radius, table order of the reference (first dimension fastest)
    int L;
    uint64_t seed;
    int literal;             // force_motion = 2: the chosen POSITION is handed to walk! as a direction (walk.jl:170, :193 as written)
};

template <int D>
__global__ void k_grw_dirs(int n, const int64_t* __restrict__ ids, int nactive, GridRw rw, int32_t* __restrict__ dirs) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= nactive) return;
    const int64_t id = ids ? ids[t] : (int64_t)t + 1;
    int j = (int)floor(ab_rw_u01(rw.seed, id, 0) * (double)rw.L);
    j = min(j, rw.L - 1);
#pragma unroll
    for (int d = 0; d < D; ++d) dirs[(size_t)t * D + d] = rw.offs[(size_t)j * D + d];
}

template <int D, bool SINGLE>
__global__ void k_grw_force_seq(GridDev<D> g, int32_t* __restrict__ pos, int64_t cap, int n, const int64_t* __restrict__ ids,
                                int nactive, GridRw rw, int32_t* __restrict__ occ, int* __restrict__ err) {
    if (blockIdx.x != 0 || threadIdx.x != 0) return;
    for (int t = 0; t < nactive; ++t) {
        int i = t;
        if (ids) {
            const int64_t id = ids[t];
            if (id < 1 || id > n) { atomicOr(err, AB_EF_OVERFLOW); continue; }
            i = (int)id - 1;
        }
        int c0[D], c1[D];
#pragma unroll
        for (int d = 0; d < D; ++d) c0[d] = pos[(size_t)d * cap + i] - 1;
        auto target = [&](int j) {
#pragma unroll
            for (int d = 0; d < D; ++d) c1[d] = grid_norm(c0[d] + 1 + rw.offs[(size_t)j * D + d], g.dims[d], g.periodic[d]) - 1;
            return lin_cell<D>(g.dims, c1);
        };
        bool found = false;
        uint32_t k1 = 0;
        for (int a = 0; a < 2 * rw.L && !found; ++a) {
            int j = (int)floor(ab_rw_u01(rw.seed, i + 1, (uint32_t)a) * (double)rw.L);
            j = min(j, rw.L - 1);
            k1 = target(j);
            found = occ[k1] == 0;
        }
        if (!found) {
            int cnt = 0;
            for (int j = 0; j < rw.L; ++j) cnt += occ[target(j)] == 0 ? 1 : 0;
            if (cnt == 0) continue;                                       // nothing empty around: the agent stays
            int pick = (int)floor(ab_rw_u01(rw.seed, i + 1, (uint32_t)(2 * rw.L)) * (double)cnt);
            pick = min(pick, cnt - 1);
            for (int j = 0; j < rw.L; ++j) {
                k1 = target(j);
                if (occ[k1] == 0 && pick-- == 0) break;
            }
        }
        if (rw.literal) {
            // walk!(agent, choice, model; ifempty = true) with choice = the empty POSITION found above, used as a
            // direction (the reference's code as written): target = normalize_position(agent.pos .+ choice); the agent
            // moves only if that target is empty (walk.jl:118-128, :141-149)
#pragma unroll
            for (int d = 0; d < D; ++d) c1[d] = grid_norm((c0[d] + 1) + (c1[d] + 1), g.dims[d], g.periodic[d]) - 1;
            k1 = lin_cell<D>(g.dims, c1);
            if (occ[k1] != 0) continue;
        }
        const uint32_t k0 = lin_cell<D>(g.dims, c0);
        if (SINGLE) { occ[k0] = 0; occ[k1] = i + 1; }
        else { occ[k0] -= 1; occ[k1] += 1; }
#pragma unroll
        for (int d = 0; d < D; ++d) pos[(size_t)d * cap + i] = c1[d] + 1;
    }
}

// offsets_at_radius(space, r) — grid_general.jl:119-136 over calculate_offsets (:138-151) and the
// periodicity-aware hyperrectangle (:47-89)
static int host_offsets_at_radius(ab_handle* h, double r, std::vector<int32_t>* out) {
    const int D = h->D;
    const double fr = floor(r + 1.0) - 1.0;
    if (!(fr >= 0.0) || fr > 30000.0) return ab_fail(h, AB_ERR_INVALID, "bad random-walk radius");
    const int64_t r0 = (int64_t)fr;
    bool notover = true;
    for (int d = 0; d < D; ++d) notover = notover && (h->desc.periodic[d] ? r0 < h->dims[d] / 2 : r0 < h->dims[d]);
    int64_t lo[3] = {0, 0, 0}, hi[3] = {0, 0, 0};
    for (int d = 0; d < D; ++d) {
        if (notover) { lo[d] = -r0; hi[d] = r0; }
        else if (h->desc.periodic[d]) {
            const int64_t half = h->dims[d] / 2, rm = std::min<int64_t>(r0, half);
            lo[d] = -rm;
            hi[d] = rm - ((rm == half && h->dims[d] % 2 == 0) ? 1 : 0);
        } else {
            const int64_t rm = std::min<int64_t>(r0, h->dims[d]);
            lo[d] = -rm; hi[d] = rm;
        }
    }
    out->clear();
    for (int64_t z = lo[2]; z <= hi[2]; ++z)
        for (int64_t y = lo[1]; y <= hi[1]; ++y)
            for (int64_t x = lo[0]; x <= hi[0]; ++x) {
                const int64_t b[3] = {x, y, z};
                int64_t l1 = 0, linf = 0;
                for (int d = 0; d < D; ++d) { l1 += std::llabs(b[d]); linf = std::max<int64_t>(linf, std::llabs(b[d])); }
                const bool keep = h->desc.metric == AB_METRIC_MANHATTAN ? l1 == r0 : linf == r0;
                if (!keep) continue;
                if (out->size() / D >= 60000) return ab_fail(h, AB_ERR_INVALID, "random-walk radius too large");
                for (int d = 0; d < D; ++d) out->push_back((int32_t)b[d]);
            }
    if (out->empty()) return ab_fail(h, AB_ERR_INVALID, "ArgumentError: no offsets at this radius");
    return AB_OK;
}

static int grid_walk_impl(ab_handle* h, const void* directions, const int64_t* active_ids, int64_t nactive, int ifempty,
                          const std::vector<int32_t>* rw_offs, uint64_t rw_seed, int rw_force);

int ab_grid_walk(ab_handle* h, const void* directions, const int64_t* active_ids, int64_t nactive, int ifempty) {
    if (!directions) return ab_fail(h, AB_ERR_INVALID, "null directions");
    return grid_walk_impl(h, directions, active_ids, nactive, ifempty, nullptr, 0, 0);
}

int ab_grid_randomwalk(ab_handle* h, double r, uint64_t seed, const int64_t* active_ids, int64_t nactive, int ifempty,
                       int force_motion) {
    if (h->desc.metric == AB_METRIC_EUCLIDEAN)
        return ab_fail(h, AB_ERR_INVALID, "ArgumentError: Random walks on a `GridSpace` with Euclidean metric are not defined.");
    std::vector<int32_t> offs;
    AB_TRY(host_offsets_at_radius(h, r, &offs));
    const bool single = h->desc.kind == AB_SPACE_GRID_SINGLE;
    if (single) ifempty = 1;                                              // no effect on GridSpaceSingle (walk.jl:141-143)
    if (force_motion < 0 || force_motion > 2) return ab_fail(h, AB_ERR_INVALID, "force_motion must be 0, 1 or 2");
    const int force = (force_motion && (single || ifempty)) ? force_motion : 0;     // walk.jl:165, :189
    if (!active_ids) nactive = h->n;
    return grid_walk_impl(h, nullptr, active_ids, nactive, ifempty, &offs, seed, force);
}

static int grid_walk_impl(ab_handle* h, const void* directions, const int64_t* active_ids, int64_t nactive, int ifempty,
                          const std::vector<int32_t>* rw_offs, uint64_t rw_seed, int rw_force) {
    if (nactive < 0) return ab_fail(h, AB_ERR_INVALID, "negative nactive");
    const int na = active_ids ? (int)nactive : (int)h->n;
    if (!active_ids && nactive != h->n && nactive != 0) return ab_fail(h, AB_ERR_INVALID, "directions for all agents need nactive == n");
    if (na == 0) return AB_OK;
    if (h->grid_cols_dirty) AB_TRY(ab_grid_sync_columns(h));
    const bool single = h->desc.kind == AB_SPACE_GRID_SINGLE;
    const bool seq = single || ifempty;
    if (seq) AB_TRY(grid_ensure_index(h));
    const size_t ids_bytes = active_ids ? (size_t)na * sizeof(int64_t) : 0;
    const size_t o1 = (ids_bytes + 255) / 256 * 256;
    const size_t dir_bytes = (size_t)na * h->D * sizeof(int32_t);
    const size_t o2 = o1 + (dir_bytes + 255) / 256 * 256;
    const size_t occ_bytes = (seq && !single) ? (size_t)h->ncells * sizeof(int32_t) : 0;
    const size_t o3 = o2 + (occ_bytes + 255) / 256 * 256;
    const size_t offs_bytes = rw_offs ? rw_offs->size() * sizeof(int32_t) : 0;
    char* base;
    AB_TRY(ab_scratch(h, o3 + offs_bytes + 256, (void**)&base));
    if (active_ids) AB_CUDA(h, cudaMemcpyAsync(base, active_ids, ids_bytes, cudaMemcpyHostToDevice, h->stream));
    const int64_t* ids = active_ids ? (const int64_t*)base : nullptr;
    const int32_t* dirs = (const int32_t*)(base + o1);
    GridRw rw{nullptr, 0, rw_seed, rw_force == 2 ? 1 : 0};
    if (rw_offs) {
        // the table lives in a vector of the caller: the copy is complete before this function returns
        // (ab_check_device_errors below synchronises; in asynchronous mode synchronise here)
        AB_CUDA(h, cudaMemcpyAsync(base + o3, rw_offs->data(), offs_bytes, cudaMemcpyHostToDevice, h->stream));
        if (h->async_mode) AB_CUDA(h, cudaStreamSynchronize(h->stream));
        rw.offs = (const int32_t*)(base + o3);
        rw.L = (int)(rw_offs->size() / h->D);
        if (!rw_force) {
            AB_LAUNCH(h, "k_grw_dirs");
            if (h->D == 2) k_grw_dirs<2><<<ab_blocks(na, 256), 256, 0, h->stream>>>((int)h->n, ids, na, rw, (int32_t*)(base + o1));
            else k_grw_dirs<3><<<ab_blocks(na, 256), 256, 0, h->stream>>>((int)h->n, ids, na, rw, (int32_t*)(base + o1));
        }
    } else {
        AB_CUDA(h, cudaMemcpyAsync(base + o1, directions, dir_bytes, cudaMemcpyHostToDevice, h->stream));
    }
    int32_t* pos = (int32_t*)h->fields[0].buf[h->cur];
    int32_t* occ = single ? h->grid_ids : (int32_t*)(base + o2);
    if (seq && !single) {
        AB_LAUNCH(h, "k_copy_counts");
        k_copy_counts<<<ab_blocks(h->ncells, 256), 256, 0, h->stream>>>(h->cell_start, h->ncells, occ);
    }
    if (rw_force) {
        AB_LAUNCH(h, "k_grw_force_seq");
        if (h->D == 2) {
            if (single) k_grw_force_seq<2, true><<<1, 32, 0, h->stream>>>(make_grid<2>(h), pos, h->cap, (int)h->n, ids, na, rw, occ, h->d_err);
            else k_grw_force_seq<2, false><<<1, 32, 0, h->stream>>>(make_grid<2>(h), pos, h->cap, (int)h->n, ids, na, rw, occ, h->d_err);
        } else {
            if (single) k_grw_force_seq<3, true><<<1, 32, 0, h->stream>>>(make_grid<3>(h), pos, h->cap, (int)h->n, ids, na, rw, occ, h->d_err);
            else k_grw_force_seq<3, false><<<1, 32, 0, h->stream>>>(make_grid<3>(h), pos, h->cap, (int)h->n, ids, na, rw, occ, h->d_err);
        }
    } else {
        AB_LAUNCH(h, seq ? "k_gwalk_seq" : "k_gwalk_all");
        if (h->D == 2) {
            if (!seq) k_gwalk_all<2><<<ab_blocks(na, 256), 256, 0, h->stream>>>(make_grid<2>(h), pos, h->cap, (int)h->n, ids, na, dirs, h->d_err);
            else if (single) k_gwalk_seq<2, true><<<1, 32, 0, h->stream>>>(make_grid<2>(h), pos, h->cap, (int)h->n, ids, na, dirs, occ, h->d_err);
            else k_gwalk_seq<2, false><<<1, 32, 0, h->stream>>>(make_grid<2>(h), pos, h->cap, (int)h->n, ids, na, dirs, occ, h->d_err);
        } else {
            if (!seq) k_gwalk_all<3><<<ab_blocks(na, 256), 256, 0, h->stream>>>(make_grid<3>(h), pos, h->cap, (int)h->n, ids, na, dirs, h->d_err);
            else if (single) k_gwalk_seq<3, true><<<1, 32, 0, h->stream>>>(make_grid<3>(h), pos, h->cap, (int)h->n, ids, na, dirs, occ, h->d_err);
            else k_gwalk_seq<3, false><<<1, 32, 0, h->stream>>>(make_grid<3>(h), pos, h->cap, (int)h->n, ids, na, dirs, occ, h->d_err);
        }
    }
    const bool keep_index = single;        // grid_ids was updated in place
    h->grid_cells_valid = false;
    int e = ab_check_device_errors_now(h);
    h->index_valid = keep_index && e == AB_OK;
    if (e == AB_ERR_INVALID) return ab_fail(h, AB_ERR_INVALID, "agent id out of range");
    return e;
}

// ---------------------------------------------------------------------------
// Schelling rule on GridSpaceSingle — docs/src/tutorial.jl:351-376, synchronous formulation:
//   count = number of agents in the neighbouring cells (radius 1 under the space's metric:
//   8 cells for :chebyshev, 4 for :manhattan / :euclidean; out-of-bounds cells skipped or wrapped)
//   whose group equals the agent's, taken from the snapshot at the start of the step;
//   mood = count >= min_to_be_happy; an unhappy agent moves to an empty cell.
// Relocation has no order-free definition in the reference (sequential rejection sampling on the
// model RNG, discrete.jl:96-114, 278-287).  The device rule: the unhappy agents U and the empty
// cells E of the snapshot are compacted in cell order; mover k (k' = π_U(k), a seeded pseudo-random
// permutation of 0..|U|-1) moves iff k' < |E|, to E[π_E(k')].  Every destination was empty at the
// start of the step, no two movers share one, and when |U| > |E| the |E| movers are a pseudo-random
// subset (move_agent_single! is a no-op when no empty position is left, discrete.jl:283-284).
// ---------------------------------------------------------------------------
#define SCH_TB 256

__global__ void k_fill_i32(int32_t* p, int64_t n, int32_t v) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}
__global__ void k_sch_scatter(GridDev<2> g, const int32_t* __restrict__ pos, int64_t cap, int n,
                              const int32_t* __restrict__ group, const uint8_t* __restrict__ mood,
                              int32_t* __restrict__ grid_group, uint8_t* __restrict__ grid_mood) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int c[2];
    load_cell<2>(g, pos, cap, i, c);
    const uint32_t k = lin_cell<2>(g.dims, c);
    grid_group[k] = group[i];
    grid_mood[k] = mood[i];
}
__global__ void k_sch_gather(GridDev<2> g, const int32_t* __restrict__ grid_ids, const uint8_t* __restrict__ grid_mood,
                             int64_t ncells, int64_t cap, int32_t* __restrict__ pos, uint8_t* __restrict__ mood) {
    const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= ncells) return;
    const int id = grid_ids[k];
    if (id == 0) return;
    const int y = (int)(k / g.dims[0]), x = (int)(k - (int64_t)y * g.dims[0]);
    pos[id - 1] = x + 1;
    pos[cap + id - 1] = y + 1;
    mood[id - 1] = grid_mood[k] == 1 ? 1 : 0;       // cell codes: 1 happy, 0 unhappy, 2 empty, 3 unhappy and not scheduled
}

// Happiness pass.  One thread per cell; a lane loads its own column of the three rows and takes the
// x±1 neighbours from the adjacent lanes (direct loads only at warp / row edges), so the stencil
// reads ~3 x 4 B per cell from L1 and 4 B per cell from DRAM.
template <bool RELOC>
__global__ void __launch_bounds__(SCH_TB)
k_sch_happy(GridDev<2> g, const int32_t* __restrict__ grid_group, uint8_t* __restrict__ grid_mood, int64_t ncells,
            int min_happy, int diag, uint32_t* __restrict__ blockU, uint32_t* __restrict__ blockE) {
    const int64_t k = (int64_t)blockIdx.x * SCH_TB + threadIdx.x;
    const bool live = k < ncells;
    const int W = g.dims[0], H = g.dims[1];
    const int64_t kk = live ? k : ncells - 1;
    const int y = (int)(kk / W), x = (int)(kk - (int64_t)y * W);
    const int lane = threadIdx.x & 31;
    int v[3], l[3], r[3];
    const int xl = x > 0 ? x - 1 : (g.periodic[0] ? W - 1 : -1);
    const int xr = x < W - 1 ? x + 1 : (g.periodic[0] ? 0 : -1);
#pragma unroll
    for (int j = 0; j < 3; ++j) {
        int yy = y + j - 1;
        if (yy < 0) yy = g.periodic[1] ? H - 1 : -1;
        else if (yy >= H) yy = g.periodic[1] ? 0 : -1;
        const int64_t rb = (int64_t)yy * W;
        v[j] = yy >= 0 ? grid_group[rb + x] : AB_SCH_EMPTY;
        int lv = __shfl_up_sync(0xffffffffu, v[j], 1), rv = __shfl_down_sync(0xffffffffu, v[j], 1);
        if (lane == 0 || x == 0) lv = (yy >= 0 && xl >= 0) ? grid_group[rb + xl] : AB_SCH_EMPTY;
        if (lane == 31 || x == W - 1) rv = (yy >= 0 && xr >= 0) ? grid_group[rb + xr] : AB_SCH_EMPTY;
        l[j] = lv; r[j] = rv;
    }
    const int me = v[1];
    const bool occupied = live && me != AB_SCH_EMPTY;
    int cnt = 0;
    cnt += (v[0] == me) + (v[2] == me) + (l[1] == me) + (r[1] == me);
    if (diag) cnt += (l[0] == me) + (r[0] == me) + (l[2] == me) + (r[2] == me);
    const bool happy = occupied && cnt >= min_happy;
    if (live) grid_mood[k] = happy ? 1 : 0;
    if (RELOC) {
        const int nu = __syncthreads_count(occupied && !happy);
        const int ne = __syncthreads_count(live && !occupied);
        if (threadIdx.x == 0) { blockU[blockIdx.x] = (uint32_t)nu; blockE[blockIdx.x] = (uint32_t)ne; }
    }
}

// Compaction of the unhappy agents' cells and of the empty cells, in cell order.
__global__ void __launch_bounds__(SCH_TB)
k_sch_compact(const int32_t* __restrict__ grid_group, const uint8_t* __restrict__ grid_mood, int64_t ncells,
              const uint32_t* __restrict__ baseU, const uint32_t* __restrict__ baseE, uint32_t* __restrict__ U,
              uint32_t* __restrict__ E) {
    __shared__ uint32_t s_u[SCH_TB / 32], s_e[SCH_TB / 32];
    const int64_t k = (int64_t)blockIdx.x * SCH_TB + threadIdx.x;
    const bool live = k < ncells;
    const int gv = live ? grid_group[k] : 0;
    const bool empty = live && gv == AB_SCH_EMPTY;
    const bool unhappy = live && !empty && grid_mood[k] == 0;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const unsigned mu = __ballot_sync(0xffffffffu, unhappy), me = __ballot_sync(0xffffffffu, empty);
    if (lane == 0) { s_u[wid] = __popc(mu); s_e[wid] = __popc(me); }
    __syncthreads();
    uint32_t ou = baseU[blockIdx.x], oe = baseE[blockIdx.x];
    for (int w = 0; w < wid; ++w) { ou += s_u[w]; oe += s_e[w]; }
    const unsigned below = (1u << lane) - 1u;
    if (unhappy) U[ou + __popc(mu & below)] = (uint32_t)k;
    if (empty) E[oe + __popc(me & below)] = (uint32_t)k;
}

// Fast happiness pass for row lengths that are multiples of 128 (one warp = 128 consecutive cells of
// a row, 4 per lane as one 16-byte load).  A block walks SCH4_R rows top-down keeping three rows in
// registers, so every cell's group is read ~1.1x from L2/DRAM and the x neighbours come from the
// adjacent lanes.  Mood byte written per cell: 1 happy, 0 unhappy, 2 empty — the compaction pass
// then needs this byte only.  chunkU / chunkE: unhappy / empty counts per 128-cell chunk (cell order).
#define SCH4_R 16
template <bool RELOC>
__global__ void __launch_bounds__(256)
k_sch_happy4(GridDev<2> g, const int32_t* __restrict__ grid_group, uint8_t* __restrict__ grid_mood, int min_happy,
             int diag, uint32_t* __restrict__ chunkU, uint32_t* __restrict__ chunkE) {
    const int W = g.dims[0], H = g.dims[1];
    const int lane = threadIdx.x & 31;
    const int x0 = (blockIdx.x * 256 + threadIdx.x) * 4;
    if (x0 >= W) return;                                   // whole warps leave together (W % 128 == 0)
    const int y0 = blockIdx.y * SCH4_R;
    const bool px = g.periodic[0], py = g.periodic[1];
    auto load_row = [&](int yy, int* o) {
        if (yy < 0) yy = py ? H - 1 : -1;
        else if (yy >= H) yy = py ? yy - H : -1;
        if (yy < 0) {
#pragma unroll
            for (int i = 0; i < 6; ++i) o[i] = AB_SCH_EMPTY;
            return;
        }
        const int32_t* row = grid_group + (size_t)yy * W;
        const int4 v = *reinterpret_cast<const int4*>(row + x0);
        int l = __shfl_up_sync(0xffffffffu, v.w, 1), r = __shfl_down_sync(0xffffffffu, v.x, 1);
        if (lane == 0) l = x0 > 0 ? row[x0 - 1] : (px ? row[W - 1] : AB_SCH_EMPTY);
        if (lane == 31) r = x0 + 4 < W ? row[x0 + 4] : (px ? row[0] : AB_SCH_EMPTY);
        o[0] = l; o[1] = v.x; o[2] = v.y; o[3] = v.z; o[4] = v.w; o[5] = r;
    };
    int up[6], cu[6], dn[6];
    load_row(y0 - 1, up);
    load_row(y0, cu);
#pragma unroll 4
    for (int r = 0; r < SCH4_R; ++r) {
        const int y = y0 + r;
        if (y >= H) break;
        load_row(y + 1, dn);
        uint32_t out = 0;
        int nu = 0, ne = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int me = cu[i + 1];
            int cnt = (up[i + 1] == me) + (dn[i + 1] == me) + (cu[i] == me) + (cu[i + 2] == me);
            if (diag) cnt += (up[i] == me) + (up[i + 2] == me) + (dn[i] == me) + (dn[i + 2] == me);
            const bool occupied = me != AB_SCH_EMPTY;
            const bool happy = occupied && cnt >= min_happy;
            out |= (occupied ? (happy ? 1u : 0u) : 2u) << (8 * i);
            nu += occupied && !happy;
            ne += !occupied;
        }
        *reinterpret_cast<uint32_t*>(grid_mood + (size_t)y * W + x0) = out;
        if (RELOC) {
            const int tu = __reduce_add_sync(0xffffffffu, nu), te = __reduce_add_sync(0xffffffffu, ne);
            if (lane == 0) {
                const size_t chunk = ((size_t)y * W + (size_t)(x0 - lane * 4)) >> 7;
                chunkU[chunk] = (uint32_t)tu;
                chunkE[chunk] = (uint32_t)te;
            }
        }
#pragma unroll
        for (int i = 0; i < 6; ++i) { up[i] = cu[i]; cu[i] = dn[i]; }
    }
}

// Compaction for the fast path: one warp per 128-cell chunk, reads the mood bytes only.
__global__ void __launch_bounds__(256)
k_sch_compact4(const uint8_t* __restrict__ grid_mood, int64_t nchunks, const uint32_t* __restrict__ baseU,
               const uint32_t* __restrict__ baseE, uint32_t* __restrict__ U, uint32_t* __restrict__ E) {
    const int64_t chunk = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (chunk >= nchunks) return;
    const int lane = threadIdx.x & 31;
    const uint32_t k0 = (uint32_t)(chunk * 128 + lane * 4);
    const uint32_t m = *reinterpret_cast<const uint32_t*>(grid_mood + k0);
    uint32_t fu = 0, fe = 0;                                // bit i: cell k0 + i is unhappy / empty
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const uint32_t b = (m >> (8 * i)) & 0xffu;
        fu |= (b == 0u ? 1u : 0u) << i;
        fe |= (b == 2u ? 1u : 0u) << i;
    }
    const uint32_t mine = (uint32_t)__popc(fu) | ((uint32_t)__popc(fe) << 16);
    uint32_t inc = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += t; }
    const uint32_t exc = inc - mine;
    uint32_t ou = baseU[chunk] + (exc & 0xffffu), oe = baseE[chunk] + (exc >> 16);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        if (fu & (1u << i)) U[ou++] = k0 + i;
        if (fe & (1u << i)) E[oe++] = k0 + i;
    }
}

// Scheduled subset (Schedulers.Partially / an id list): after the happiness pass, the agents that were NOT scheduled
// get their previous mood back — coded 1 (happy) or 3 (unhappy, not a mover) — and the unhappy count of the chunk
// (one CTA = one chunk of the compaction: 128 or SCH_TB cells) is recounted.
__global__ void k_sch_mask(const int32_t* __restrict__ grid_ids, const uint8_t* __restrict__ old_mood, uint8_t* __restrict__ grid_mood,
                           int64_t ncells, const uint8_t* __restrict__ active, uint32_t* __restrict__ chunkU) {
    const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    bool unhappy = false;
    if (k < ncells) {
        const int id = grid_ids[k];
        if (id != 0) {
            uint8_t m = grid_mood[k];
            if (!active[id - 1]) { m = old_mood[k] == 1 ? 1 : 3; grid_mood[k] = m; }
            unhappy = m == 0;
        }
    }
    const int nu = __syncthreads_count(unhappy);
    if (threadIdx.x == 0 && chunkU) chunkU[blockIdx.x] = (uint32_t)nu;
}

// Seeded pseudo-random permutation of [0, n): 4-round Feistel network on 2*hb bits, cycle-walked into range.
__device__ __forceinline__ uint32_t sch_mix(uint32_t x, uint64_t key) {
    uint64_t z = (uint64_t)x + key + 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return (uint32_t)((z ^ (z >> 31)) >> 16);
}
__device__ __forceinline__ uint32_t sch_perm(uint32_t i, uint32_t n, uint64_t key) {
    if (n <= 1) return 0;
    int bits = 32 - __clz(n - 1);
    if (bits < 2) bits = 2;
    const int hb = (bits + 1) >> 1;
    const uint32_t mask = (1u << hb) - 1u;
    uint32_t x = i;
    do {
        uint32_t L = x >> hb, R = x & mask;
#pragma unroll
        for (int rd = 0; rd < 4; ++rd) {
            const uint32_t t = L ^ (sch_mix(R, key + (uint64_t)rd * 0xD1B54A32D192ED03ull) & mask);
            L = R; R = t;
        }
        x = (L << hb) | R;
    } while (x >= n);
    return x;
}

__global__ void k_sch_move(GridDev<2> g, const uint32_t* __restrict__ totU, const uint32_t* __restrict__ totE,
                           const uint32_t* __restrict__ U, const uint32_t* __restrict__ E, uint64_t key,
                           int32_t* __restrict__ grid_ids, int32_t* __restrict__ grid_group, uint8_t* __restrict__ grid_mood,
                           long long* __restrict__ moved) {
    const uint32_t nU = *totU, nE = *totE;
    if (blockIdx.x == 0 && threadIdx.x == 0) *moved = (long long)(nU < nE ? nU : nE);
    for (uint32_t k = blockIdx.x * blockDim.x + threadIdx.x; k < nU; k += gridDim.x * blockDim.x) {
        const uint32_t kp = nU > nE ? sch_perm(k, nU, key) : k;
        if (kp >= nE) continue;                                      // no empty position left for this one
        const uint32_t src = U[k], dst = E[sch_perm(kp, nE, key ^ 0xA5A5A5A55A5A5A5Aull)];
        grid_ids[dst] = grid_ids[src];
        grid_group[dst] = grid_group[src];
        grid_mood[dst] = 0;
        grid_ids[src] = 0;
        grid_group[src] = AB_SCH_EMPTY;
    }
}

// ---- host side of the Schelling step -------------------------------------------------------------
static int sch_field(ab_handle* h, const char* name, int dtype) {
    const int f = ab_field_id(h, name);
    if (f < 0 || h->fields[f].dtype != dtype || h->fields[f].ncomp != 1) return -1;
    return f;
}

static int sch_ensure_cells(ab_handle* h, int fgroup, int fmood) {
    if (!h->grid_group) {
        AB_CUDA(h, cudaMalloc((void**)&h->grid_group, (size_t)h->ncells * sizeof(int32_t)));
        AB_CUDA(h, cudaMalloc((void**)&h->grid_mood, (size_t)h->ncells));
        AB_CUDA(h, cudaMallocHost((void**)&h->h_moved, sizeof(int64_t)));
    }
    if (h->grid_cells_valid) return AB_OK;
    AB_TRY(grid_ensure_index(h));
    {
        AB_LAUNCH(h, "k_fill_i32");
        k_fill_i32<<<ab_blocks(h->ncells, 256), 256, 0, h->stream>>>(h->grid_group, h->ncells, AB_SCH_EMPTY);
    }
    AB_CUDA(h, cudaMemsetAsync(h->grid_mood, 0, (size_t)h->ncells, h->stream));
    if (h->n > 0) {
        AB_LAUNCH(h, "k_sch_scatter");
        k_sch_scatter<<<ab_blocks(h->n, 256), 256, 0, h->stream>>>(make_grid<2>(h), (const int32_t*)h->fields[0].buf[h->cur], h->cap,
                                                                  (int)h->n, (const int32_t*)h->fields[fgroup].buf[h->cur],
                                                                  (const uint8_t*)h->fields[fmood].buf[h->cur], h->grid_group, h->grid_mood);
    }
    h->grid_cells_valid = true;
    return AB_OK;
}

int ab_grid_sync_columns(ab_handle* h) {
    if (!h->grid_cols_dirty) return AB_OK;
    const int fmood = sch_field(h, "mood", AB_U8);
    if (fmood < 0) return ab_fail(h, AB_ERR_STATE, "mood column disappeared");
    {
        AB_LAUNCH(h, "k_sch_gather");
        k_sch_gather<<<ab_blocks(h->ncells, 256), 256, 0, h->stream>>>(make_grid<2>(h), h->grid_ids, h->grid_mood, h->ncells, h->cap,
                                                                      (int32_t*)h->fields[0].buf[h->cur], (uint8_t*)h->fields[fmood].buf[h->cur]);
    }
    h->grid_cols_dirty = false;
    return AB_OK;
}

void ab_grid_free(ab_handle* h) {
    if (h->grid_group) cudaFree(h->grid_group);
    if (h->grid_mood) cudaFree(h->grid_mood);
    if (h->h_moved) cudaFreeHost(h->h_moved);
    h->grid_group = nullptr; h->grid_mood = nullptr; h->h_moved = nullptr;
}

extern "C" int ab_nearby_positions_count(ab_handle* h, const int32_t* qpos, int64_t nq, double r, int empty_only, int64_t* counts) {
    if (h->desc.kind == AB_SPACE_CONTINUOUS) return ab_fail(h, AB_ERR_INVALID, "nearby_positions is for GridSpace / GridSpaceSingle handles");
    AB_CUDA(h, cudaSetDevice(h->device));
    return ab_grid_positions(h, qpos, nq, r, empty_only, counts, nullptr, nullptr);
}
extern "C" int ab_nearby_positions_fill(ab_handle* h, const int32_t* qpos, int64_t nq, double r, int empty_only,
                                        const int64_t* offsets, int32_t* pos_out) {
    if (h->desc.kind == AB_SPACE_CONTINUOUS) return ab_fail(h, AB_ERR_INVALID, "nearby_positions is for GridSpace / GridSpaceSingle handles");
    if (!pos_out && nq > 0 && offsets && offsets[nq] > 0) return ab_fail(h, AB_ERR_INVALID, "null output");
    AB_CUDA(h, cudaSetDevice(h->device));
    if (nq > 0 && offsets && offsets[nq] == 0) return AB_OK;
    return ab_grid_positions(h, qpos, nq, r, empty_only, nullptr, offsets, pos_out);
}

extern "C" int ab_step_schelling(ab_handle* h, const ab_schelling_params* p, int64_t nsteps, int64_t* n_moved);
// The Schelling step for a scheduled subset (schedulers.jl:68-120): only the scheduled agents run schelling_step!
// (tutorial.jl:351-376) — the others keep their mood and their position, and stay neighbours of everyone.
extern "C" int ab_step_schelling_active(ab_handle* h, const ab_schelling_params* p, int64_t nsteps, const int64_t* active_ids,
                                        int64_t nactive, int64_t* n_moved) {
    if (!active_ids) return ab_step_schelling(h, p, nsteps, n_moved);
    if (h->desc.kind != AB_SPACE_GRID_SINGLE) return ab_fail(h, AB_ERR_INVALID, "the Schelling rule needs a 2-D GridSpaceSingle handle");
    if (n_moved) *n_moved = 0;
    if (h->n == 0 || nsteps == 0) return AB_OK;
    AB_TRY(ab_set_step_mask(h, active_ids, nactive));
    h->step_mask_cur = h->step_mask;
    const int rc = ab_step_schelling(h, p, nsteps, n_moved);
    h->step_mask_cur = nullptr;
    return rc;
}

extern "C" int ab_step_schelling(ab_handle* h, const ab_schelling_params* p, int64_t nsteps, int64_t* n_moved) {
    if (!p) return ab_fail(h, AB_ERR_INVALID, "null params");
    if (h->desc.kind != AB_SPACE_GRID_SINGLE || h->D != 2) return ab_fail(h, AB_ERR_INVALID, "the Schelling rule needs a 2-D GridSpaceSingle handle");
    if (nsteps < 0) return ab_fail(h, AB_ERR_INVALID, "negative nsteps");
    for (int d = 0; d < 2; ++d)
        if (h->desc.periodic[d] && h->dims[d] < 3) return ab_fail(h, AB_ERR_INVALID, "periodic dimensions need at least 3 cells");
    const int fgroup = sch_field(h, "group", AB_I32), fmood = sch_field(h, "mood", AB_U8);
    if (fgroup < 0 || fmood < 0) return ab_fail(h, AB_ERR_STATE, "register columns group (Int32) and mood (UInt8) first");
    AB_CUDA(h, cudaSetDevice(h->device));
    if (n_moved) *n_moved = 0;
    if (nsteps == 0) return AB_OK;
    AB_TRY(sch_ensure_cells(h, fgroup, fmood));
    const int64_t ncells = h->ncells;
    const bool fast = h->dims[0] % 128 == 0;               // whole-warp rows: k_sch_happy4 / k_sch_compact4
    const int nblocks = fast ? (int)(ncells / 128) : ab_blocks(ncells, SCH_TB);
    const size_t nbpad = ((size_t)nblocks + SCAN_TILE - 1) / SCAN_TILE * SCAN_TILE + SCAN_TILE;
    // scratch: blockU, blockE (padded), baseU, baseE (nblocks + 1), U (<= n), E (<= ncells - n), moved
    uint32_t *blockU = nullptr, *blockE = nullptr, *baseU = nullptr, *baseE = nullptr, *U = nullptr, *E = nullptr;
    long long* d_moved = nullptr;
    const int diag = h->desc.metric == AB_METRIC_CHEBYSHEV ? 1 : 0;
    if (p->relocate) {
        char* base;
        const size_t lists = ((size_t)ncells + 64) * sizeof(uint32_t);
        AB_TRY(ab_scratch2(h, 256 + 4 * nbpad * sizeof(uint32_t) + lists + 256, (void**)&base));
        d_moved = (long long*)base;
        blockU = (uint32_t*)(base + 256); blockE = blockU + nbpad; baseU = blockE + nbpad; baseE = baseU + nbpad;
        U = baseE + nbpad;                       // U and E share one ncells-sized array: U from the front,
        E = U + h->n;                            // E behind the (at most n) unhappy entries
        AB_CUDA(h, cudaMemsetAsync(base, 0, 256 + 2 * nbpad * sizeof(uint32_t), h->stream));
    }
    GridDev<2> g = make_grid<2>(h);
    uint8_t* old_mood = nullptr;
    if (h->step_mask_cur) AB_TRY(ab_scratch(h, (size_t)ncells, (void**)&old_mood));
    AB_CUDA(h, cudaEventRecord(h->ev0, h->stream));
    for (int64_t st = 0; st < nsteps; ++st) {
        if (old_mood) AB_CUDA(h, cudaMemcpyAsync(old_mood, h->grid_mood, (size_t)ncells, cudaMemcpyDeviceToDevice, h->stream));
        {
            AB_LAUNCH(h, fast ? "k_sch_happy4" : "k_sch_happy");
            const dim3 g4((h->dims[0] + 1023) / 1024, (h->dims[1] + SCH4_R - 1) / SCH4_R);
            if (fast && p->relocate) k_sch_happy4<true><<<g4, 256, 0, h->stream>>>(g, h->grid_group, h->grid_mood, p->min_to_be_happy, diag, blockU, blockE);
            else if (fast) k_sch_happy4<false><<<g4, 256, 0, h->stream>>>(g, h->grid_group, h->grid_mood, p->min_to_be_happy, diag, nullptr, nullptr);
            else if (p->relocate) k_sch_happy<true><<<nblocks, SCH_TB, 0, h->stream>>>(g, h->grid_group, h->grid_mood, ncells, p->min_to_be_happy, diag, blockU, blockE);
            else k_sch_happy<false><<<nblocks, SCH_TB, 0, h->stream>>>(g, h->grid_group, h->grid_mood, ncells, p->min_to_be_happy, diag, nullptr, nullptr);
        }
        if (old_mood) {
            AB_LAUNCH(h, "k_sch_mask");
            k_sch_mask<<<nblocks, fast ? 128 : SCH_TB, 0, h->stream>>>(h->grid_ids, old_mood, h->grid_mood, ncells, h->step_mask_cur,
                                                                     p->relocate ? blockU : nullptr);
        }
        if (p->relocate) {
            AB_TRY(ab_scan_u32(h, blockU, baseU, (uint32_t)nblocks));
            AB_TRY(ab_scan_u32(h, blockE, baseE, (uint32_t)nblocks));
            {
                AB_LAUNCH(h, fast ? "k_sch_compact4" : "k_sch_compact");
                if (fast) k_sch_compact4<<<ab_blocks((int64_t)nblocks * 32, 256), 256, 0, h->stream>>>(h->grid_mood, nblocks, baseU, baseE, U, E);
                else k_sch_compact<<<nblocks, SCH_TB, 0, h->stream>>>(h->grid_group, h->grid_mood, ncells, baseU, baseE, U, E);
            }
            {
                AB_LAUNCH(h, "k_sch_move");
                const uint64_t key = p->seed * 0x9E3779B97F4A7C15ull + (uint64_t)(h->sch_tick++) * 0xD6E8FEB86659FD93ull;
                k_sch_move<<<148 * 8, 256, 0, h->stream>>>(g, baseU + nblocks, baseE + nblocks, U, E, key, h->grid_ids, h->grid_group,
                                                          h->grid_mood, d_moved);
            }
        }
    }
    AB_CUDA(h, cudaEventRecord(h->ev1, h->stream));
    if (p->relocate) AB_CUDA(h, cudaMemcpyAsync(h->h_moved, d_moved, sizeof(long long), cudaMemcpyDeviceToHost, h->stream));
    h->grid_cols_dirty = true;
    AB_TRY(ab_check_device_errors_now(h));               // synchronous call: *h_moved and the events are read below
    h->index_valid = true;
    float ms = 0;
    AB_CUDA(h, cudaEventElapsedTime(&ms, h->ev0, h->ev1));
    h->last_step_ms = ms;
    h->last_moved = p->relocate ? *h->h_moved : 0;
    if (n_moved) *n_moved = h->last_moved;
    return AB_OK;
}
