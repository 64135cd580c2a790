// ab_internal.cuh — shared definitions of the B200-native Agents.jl hot path.
//
// Data layout in HBM (one population per handle, capacity `cap` agents):
//   * every agent column ("field") is stored SoA *per component* and in
//     CELL-SORTED order: comp c of field f lives at buf[cur] + c*cap.  Slot s
//     holds the agent id[cur][s] (0-based).  The sorted order is persistent:
//     a step writes the next state directly into the other buffer in the next
//     step's sorted order, so there is no per-step gather/scatter by id.
//   * device cells: every reference cell (pos2cell, continuous.jl:112-113) is
//     split into K^D "fine" cells (K = 1 or 2 per dimension, split at the
//     reference cell's centre), and the device key of an agent is the
//     column-major linear index of its fine cell over the fine grid
//     (dims[d]*K) — x fastest, like Julia's stored_ids[c...]
//     (src/spaces/grid_multi.jl:73).  A run of fine cells of one fine row is a
//     contiguous slot range, so both the reference's cell walks (K runs per
//     reference cell row) and the short-range sweeps of the fused rules
//     (fine rows fy-1..fy+1) read contiguous memory without restaging.
//   * cell_start[nfine+1]: first slot of each fine cell.  Within a fine cell,
//     slots are ordered by id, so every floating-point sum over a cell's
//     agents has one canonical order (bitwise reproducible runs).
//   * key[cur][s]: fine-cell key of slot s.
// Compiled with --fmad=false: predicates that decide neighbour sets and cells
// must round exactly like the reference (Julia never contracts a*b+c here);
// fma() is used explicitly only where results are tolerance-checked.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <utility>
#include <vector>

#include "../../include/agents_b200.h"

#define AB_VERSION 100

#define AB_SCH_EMPTY INT32_MIN
enum { AB_EF_OUTSIDE = 1, AB_EF_NUMERIC = 2, AB_EF_OCCUPIED = 4, AB_EF_OVERFLOW = 8, AB_EF_COMM = 16, AB_EF_ARG = 32,
       // what went wrong in the exchange (reported with AB_ERR_COMM)
       AB_EF_COMM_CAP = 1 << 8, AB_EF_COMM_HDR = 1 << 9, AB_EF_COMM_WINDOW = 1 << 10, AB_EF_COMM_TIMEOUT = 1 << 11 };

// Counter-based random stream of the bulk randomwalk! calls (the reference draws from the model's RNG,
// src/spaces/walk.jl:154-214, 389-408; its stream is not reproduced — SURVEY §8c(iv) — so the device rule is
// defined on a counter: draw c of agent `id` under `seed` is u = (splitmix64(seed ^ ((id << 16) + c)) >> 11) * 2^-53.
__host__ __device__ __forceinline__ uint64_t ab_splitmix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    uint64_t z = x;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
__host__ __device__ __forceinline__ double ab_rw_u01(uint64_t seed, int64_t id, uint32_t c) {
    return (double)(ab_splitmix64(seed ^ (((uint64_t)id << 16) + c)) >> 11) * (1.0 / 9007199254740992.0);
}

struct ab_field {
    std::string name;
    int dtype = AB_F64;
    int ncomp = 1;
    size_t esize = 8;
    void* buf[2] = {nullptr, nullptr};
};

struct ab_kernel_stat {
    const char* name;
    double ms;
    int64_t launches;
};

struct ab_pending_event {
    const char* name;
    cudaEvent_t a, b;
};

struct ab_comm;   // multi-GPU state (ab_comm.cu)

struct ab_handle {
    ab_space_desc desc{};
    int device = 0;
    cudaStream_t stream = nullptr;
    int D = 2;
    int dtype = AB_F64;
    size_t tsize = 8;
    int64_t n = 0, cap = 0;
    int64_t ncells = 1;                // reference cells (prod dims)
    int dims[3] = {1, 1, 1};
    int K = 1;                         // refinement: each reference cell holds K^D device (fine) cells
    int kb = 0;                        // log2(K)
    int ksh = 0;                       // log2(K^D)
    int64_t nfine = 1;                 // ncells << ksh : entries of cell_start - 1
    int64_t nfine_cap = 1;             // allocated (padded) length of cell_count

    std::vector<ab_field> fields;      // 0 = pos, 1 = vel (continuous); 0 = pos (grids)
    int cur = 0;
    void* tmp_pos = nullptr;           // D*cap of T: new state in OLD slot order
    void* tmp_vel = nullptr;
    int32_t* id[2] = {nullptr, nullptr};
    uint32_t* key[2] = {nullptr, nullptr};
    uint32_t* newkey = nullptr;        // per old slot
    uint32_t* rank = nullptr;          // arrival rank inside the new cell (atomic order)
    uint32_t* cell_count = nullptr;    // ncells (zero between uses)
    uint32_t* cell_start = nullptr;    // nfine + 1 (current)
    uint32_t* cell_start_alt = nullptr; // written by the tiled re-sort, then swapped in
    int32_t* idtmp = nullptr;          // ids bucketed by cell, arrival order
    int32_t* perm = nullptr;           // new slot -> old slot (for extra fields)
    int32_t* slot_of_id = nullptr;
    bool slot_map_valid = false;
    bool index_valid = false;
    bool agg_valid = false;            // cell_agg matches the current sorted state
    uint64_t agg_epoch = 1, M_epoch = 0; // cell_M was built from cell_agg version M_epoch
    int32_t* ncount = nullptr;         // neighbour count per slot of the last flock step

    // per (fine row of a) reference cell: count, Σ(pos - centre), Σ vel of the sorted state
    // (SoA: [part][1+2D][ncells] of T) — input of the aggregated approximate-search sweep (ab_flock.cu)
    void* cell_agg = nullptr;
    size_t cell_agg_bytes = 0;
    void* cell_agg1 = nullptr;         // the 4 fine-row partials of cell_agg collapsed to one (k_agg_collapse)
    size_t cell_agg1_bytes = 0;
    // sums of those aggregates over the reference's offset tables of radius 1 and 2 (k_cell_M)
    void* cell_M = nullptr;
    size_t cell_M_bytes = 0;
    cudaStream_t stream2 = nullptr;    // side stream of the overlapped sharded step (halo install + boundary sweep)
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    bool pdl = false;                  // programmatic dependent launch of the step kernels (AB_PDL=1)
    bool overlap_halo = false;         // sharded flock step: interior sweep while the halo travels — measured slower, AB_OVERLAP=1 enables
    bool fine4 = false;                // quarter-cell refinement (K = 4) on the 2-D flock path: correct, measured slower (AB_FINE4=1 enables)
    int rs_target = 100;               // average agents per re-sort tile (tuning knob)
    // re-sort tiles (ab_resort_plan): the fused step kernels count the new keys per fine cell AND per tile, a small
    // scan of the tile totals gives every tile its output range, and the tile's own warp turns its cell counts into
    // the new cell_start (there is no scan over the fine cells on this path)
    int rs_tw = 0, rs_tpr = 0, rs_ntiles = 0, rs_nrows = 0;
    uint32_t rs_magic = 0;             // ceil(2^32 / rs_tw): fx / tw == umulhi(fx, magic) for fx < 65536
    uint32_t* tile_count = nullptr;    // [2 parities][tile_cap]: rs_ntiles entries each (zero between uses)
    uint32_t* tile_base = nullptr;     // rs_ntiles + 1 (scan path only)
    size_t tile_cap = 0;
    // scan-free path (few fine rows: 2-D spaces): the step kernels also count per fine ROW; a re-sort tile sums the
    // rows before its own and the tiles before it in its row — no scan launch at all.  The counters alternate by
    // parity: step t accumulates into [t & 1], re-sort t reads them and zeroes the other half behind the previous one.
    uint32_t* row_count = nullptr;     // [2][row_cap]
    size_t row_cap = 0;
    bool rs_scanfree = false;
    int rs_par = 0;

    // single-pass scan state
    unsigned long long* scan_status = nullptr;
    size_t scan_status_len = 0;
    uint32_t* scan_ticket = nullptr;
    uint32_t scan_epoch = 0;

    int* d_err = nullptr;              // device error flag word
    int* h_err = nullptr;              // pinned mirror
    uint8_t* step_mask = nullptr;          // ab_step_flock_active: 1 per scheduled agent, by id (device, n bytes)
    size_t step_mask_bytes = 0;
    const uint8_t* step_mask_cur = nullptr; // set for the duration of a masked step call
    const int32_t* place_prio = nullptr;   // ab_layout_set: in-cell order of a restored checkpoint (rank per id)

    void* scratch = nullptr;           // grow-only device scratch for uploads / queries
    size_t scratch_bytes = 0;
    void* scratch2 = nullptr;
    size_t scratch2_bytes = 0;

    // pairs query memo
    double pairs_r = 0;
    int pairs_search = 0;
    int64_t pairs_n = -1;
    unsigned long long* pairs_offsets = nullptr;   // per slot exclusive offsets (device)
    size_t pairs_offsets_len = 0;

    // grid spaces
    int32_t* grid_ids = nullptr;       // GridSpaceSingle: id per cell (0 = empty), 1-based
    int32_t* grid_group = nullptr;     // Schelling: group of the cell's occupant (AB_SCH_EMPTY when empty)
    uint8_t* grid_mood = nullptr;      // Schelling: mood of the cell's occupant
    bool grid_cells_valid = false;     // grid_group / grid_mood mirror the agent columns
    bool grid_cols_dirty = false;      // agent columns (pos, mood) are stale w.r.t. the cell-major arrays
    int64_t last_moved = 0;
    int64_t* h_moved = nullptr;        // pinned
    uint64_t sch_tick = 0;             // steps taken so far (decorrelates the per-step permutations)

    ab_comm* comm = nullptr;
    bool cellm_old = false;            // AB_CELLM_OLD=1: the one-row-per-warp k_cell_M in 2-D too (A/B experiments)
    bool small_scan = false;           // AB_SMALL_SCAN=1: single-CTA scan of the tile totals (A/B experiments)
    bool no_split_M = false;           // AB_SPLIT_M=0: cell sums in one launch after the install (A/B experiments)
    bool no_side_unpack = false;       // AB_SIDE_UNPACK=0: halo install after the re-sort on the same stream (A/B experiments)
    bool halo_stale = false;           // a halo message timed out: the window is half-updated until ab_comm_exchange has run
    bool side_unpack_now = false;      // the halo install being queued runs beside the re-sort (one-thread flag wait first)
    bool unpack_tb = false;            // ... or the scanned tile totals do (scan path)
    bool unpack_rows = false;          // this step's per-row counters give the owned count (set by a scan-free re-sort)
    bool halo_packed = false;          // the re-sort of this step already delivered the boundary messages
    // slab window along the last axis (multi-GPU, ab_comm.cu); W == 0: the handle holds the whole space
    int win_gdl = 0, win_goff = 0, win_nl = 0, win_W = 0;

    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    double last_step_ms = 0;
    std::vector<int64_t> near_lo, near_hi;   // result of the last ab_pairs_nearest
    bool async_mode = false;           // ab_set_async: calls return once their work is queued; ab_sync collects errors
    bool step_pending = false;         // ev0/ev1 of an asynchronous step not read yet
    int64_t launches = 0;
    bool profiling = false;
    std::vector<ab_kernel_stat> kstats;
    std::vector<ab_pending_event> pending;   // recorded, not yet resolved launch timings
    std::string err;
};

// ---------------------------------------------------------------------------
// host helpers
// ---------------------------------------------------------------------------
int ab_fail(ab_handle* h, int code, const std::string& msg);
int ab_cuda_fail(ab_handle* h, cudaError_t e, const char* what);
#define AB_CUDA(h, call)                                              \
    do {                                                              \
        cudaError_t _e = (call);                                      \
        if (_e != cudaSuccess) return ab_cuda_fail((h), _e, #call);   \
    } while (0)
#define AB_TRY(expr)                \
    do {                            \
        int _rc = (expr);           \
        if (_rc != AB_OK) return _rc; \
    } while (0)

int ab_scratch(ab_handle* h, size_t bytes, void** out);     // scratch  (grow-only)
int ab_scratch2(ab_handle* h, size_t bytes, void** out);    // scratch2 (grow-only)
int ab_check_device_errors(ab_handle* h);                   // reads + clears d_err
int ab_check_device_errors_now(ab_handle* h);               // same, also in asynchronous mode (calls that hand back host data)
int ab_ensure_index(ab_handle* h);
int ab_ensure_slot_map(ab_handle* h);
int ab_set_refine(ab_handle* h, int K);
int ab_reindex_from(ab_handle* h, const void* src_pos, const void* src_vel, bool count_done);
int ab_permute_extra_fields(ab_handle* h, bool owned_only);   // applies h->perm (new slot -> old slot) to user columns, cur -> cur^1
bool ab_resort_applicable(const ab_handle* h);
int ab_resort_tiled(ab_handle* h, const void* src_pos, const void* src_vel, bool want_agg, cudaEvent_t after_scan = nullptr);
int ab_resort_plan(ab_handle* h);              // tile geometry of the next tiled re-sort (before the step kernel that counts the keys)
int ab_ensure_agg(ab_handle* h);
int ab_agg_range(ab_handle* h, int64_t cell0, int64_t ncell);
int ab_run_scan(ab_handle* h, uint32_t* dst);
// tile of the single-pass scan (k_scan, ab_core.cu): CTA size x items per thread
#ifndef SCAN_THREADS
#define SCAN_THREADS 256
#endif
#ifndef SCAN_ITEMS
#define SCAN_ITEMS 16
#endif
#define SCAN_TILE (SCAN_THREADS * SCAN_ITEMS)
int ab_scan_u32(ab_handle* h, uint32_t* count, uint32_t* start, uint32_t n);   // count padded to a multiple of SCAN_TILE             // cell_count -> exclusive offsets in dst (zeroes cell_count)
size_t ab_dtype_size(int dtype);
// cross-file internals (C++ linkage)
void ab_comm_free(ab_handle* h);
// halo refresh + migration after a re-sort (slab window): pack my boundary layers, move the messages
// (NCCL between processes / copies between in-process peers), install the received layers as my halo
int ab_comm_pack(ab_handle* h);
int ab_comm_transfer(ab_handle* h);
int ab_comm_unpack(ab_handle* h, bool want_agg);
bool ab_comm_local(const ab_handle* h);              // linked to in-process peers (ab_comm_link_local)
bool ab_comm_direct_on(const ab_handle* h);          // boundary messages travel by peer-memory stores + flags
int ab_comm_refresh_counts(ab_handle* h);            // owned / window agent counts back to the host
int ab_set_step_mask(ab_handle* h, const int64_t* active_ids, int64_t nactive);   // by-id mask of a scheduled subset -> h->step_mask
int ab_grid_build(ab_handle* h);
int ab_grid_sync_columns(ab_handle* h);   // cell-major Schelling state -> agent columns (pos, mood)
void ab_grid_free(ab_handle* h);
int ab_grid_walk(ab_handle* h, const void* directions, const int64_t* active_ids, int64_t nactive, int ifempty);
int ab_grid_randomwalk(ab_handle* h, double r, uint64_t seed, const int64_t* active_ids, int64_t nactive, int ifempty, int force_motion);
int ab_grid_nearby(ab_handle* h, const void* qpos, const int64_t* qids, int64_t nq, double r, int exclude_self,
                   int64_t* counts, const int64_t* offsets, int64_t* ids_out);

// ---- boundary messages of a slab window (ab_comm.cu) ----------------------------------------------
// [hdr 16 B][cs slice (ncs + 1) u32, padded to 16 B][pos D x C][vel D x C][id C][key C]
// ... followed by the user columns (every registered field beyond pos / vel: the r, k, mass of the pair-force rule,
// per-bird parameters, ...), each [ncomp x C] of its element size
#define AB_MSG_MAXF 8
struct MsgLayout {
    size_t off_cs, off_pos, off_vel, off_id, off_key, bytes;
    uint32_t ncs;
    int64_t C;
    int nx;                              // user columns in the message
    size_t off_x[AB_MSG_MAXF];
    int x_ncomp[AB_MSG_MAXF], x_esize[AB_MSG_MAXF];
};
// Where the tiled re-sort of a slab window delivers its boundary layers while it places them (ab_comm_pack_begin):
// the W bottom owned layers go to the lower neighbour's message, the W top ones to the upper neighbour's — into
// the neighbour's memory (direct exchange) or into this rank's send buffers.  The last boundary tile to finish
// publishes the exchange number in the neighbours' flags.  enabled == 0: a separate k_halo_pack launch packs.
struct HaloOut {
    int enabled;
    char* buf[2];              // nullptr: no neighbour on that side
    uint32_t* flag[2];         // direct exchange only
    uint32_t* done_ctr;        // direct exchange only
    uint32_t epoch, ncs, kown;
    MsgLayout m;
};
int ab_comm_pack_begin(ab_handle* h, HaloOut* out);

// launch bookkeeping: every kernel launch of the library goes through this
struct ab_launch_scope {
    ab_handle* h;
    const char* name;
    cudaEvent_t a = nullptr, b = nullptr;
    ab_launch_scope(ab_handle* h_, const char* name_);
    ~ab_launch_scope();
};
#define AB_CAT2(a, b) a##b
#define AB_CAT(a, b) AB_CAT2(a, b)
#define AB_LAUNCH(h, name) ab_launch_scope AB_CAT(_scope_, __LINE__)((h), (name))

// A fused step that ends early (a failing call between the sweep and the re-sort, or a device error flag) leaves the
// per-tile / per-row histograms of that step half-consumed and rs_par possibly out of phase with the data: the next
// ab_resort_plan must clear both halves, and the index is rebuilt from the positions the handle holds.
// The per-cell counts (cell_count: zero between uses, the sweep adds the new keys, the re-sort / scan consumes them) are
// cleared as well — the index build counts into them.  Returns rc.
int ab_step_failed(ab_handle* h, int rc);
// A fused step that failed ON THE DEVICE has also run its re-sort on incomplete keys (a bird whose new position is NaN
// or outside the space has no cell): some agents were not placed, so the columns and the slot -> id table of the handle
// no longer describe the population.  Like a reference model whose step! threw half way, the state is undefined; what
// the handle guarantees is that it can be repaired: the slot order is reset to id order, so the caller's next
// ab_upload of pos, vel and the user columns lands every row where its id says, and the index is rebuilt from them.
int ab_step_state_lost(ab_handle* h, int rc);

static inline int ab_blocks(int64_t n, int threads) { return (int)((n + threads - 1) / threads); }

// Programmatic dependent launch for the kernels of the fused step (they form one dependency chain on one stream:
// sweep -> scan -> re-sort -> aggregates -> sweep).  Every such kernel begins with ab_dep_wait(): nothing of it
// touches memory before the kernel in front of it has completed and flushed, so the only thing that overlaps is
// the launch itself and the residency of its first CTAs — the gap between two dependent kernels.
__device__ __forceinline__ void ab_dep_wait() {
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}
// Tile geometry of the re-sort as the fused step kernels see it: they histogram the new keys per fine cell
// (cell_count) and per tile (tile_count).  Tile = tw consecutive fine cells of one fine row; only the first
// `nrows` fine rows are re-sorted (all of them, or the owned ones of a slab window).
struct TileGeo {
    uint32_t* cell_count;
    uint32_t* tile_count;
    uint32_t* row_count;      // nullptr: the tile totals are scanned (many fine rows)
    uint32_t tpr, magic, nrows;
};
static inline TileGeo make_tilegeo(const ab_handle* h) {
    TileGeo g;
    g.cell_count = h->cell_count;
    g.tile_count = h->tile_count + (size_t)h->rs_par * h->tile_cap;
    g.row_count = h->rs_scanfree ? h->row_count + (size_t)h->rs_par * h->row_cap : nullptr;
    g.tpr = (uint32_t)h->rs_tpr; g.magic = h->rs_magic; g.nrows = (uint32_t)h->rs_nrows;
    return g;
}
#ifdef __CUDACC__
// one agent's new key: fine row `frow`, fine x `fx` (< 65536).  The tile counter is bumped once per group of lanes
// that share the tile (lanes of a warp are neighbours in space: one or two groups).
__device__ __forceinline__ void count_tile(const TileGeo& g, uint32_t frow, uint32_t fx) {
    const uint32_t tile = frow * g.tpr + __umulhi(fx, g.magic);
    const unsigned act = __activemask();
    const unsigned me = 1u << (threadIdx.x & 31);
    const unsigned m = __match_any_sync(act, tile);
    if ((m & (0u - m)) == me) atomicAdd(&g.tile_count[tile], (uint32_t)__popc(m));
    if (g.row_count) {
        const unsigned mr = __match_any_sync(act, frow);
        if ((mr & (0u - mr)) == me) atomicAdd(&g.row_count[frow], (uint32_t)__popc(mr));
    }
}
__device__ __forceinline__ void count_new_key(const TileGeo& g, uint32_t key, uint32_t frow, uint32_t fx) {
    if (frow >= g.nrows) return;                              // not re-sorted here (halo rows of a slab window)
    atomicAdd(&g.cell_count[key], 1u);
    count_tile(g, frow, fx);
}
#endif

template <class... KArgs, class... Args>
static inline cudaError_t ab_launch_chain(bool pdl, void (*kern)(KArgs...), dim3 grid, dim3 block, cudaStream_t st, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = 0;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}

// ---------------------------------------------------------------------------
// device-side space description and reference arithmetic
// ---------------------------------------------------------------------------
template <class T, int D> struct SpaceDev {
    T extent[D];
    T spacing;
    T inv_spacing;   // rounded 1/spacing — only for guarded shortcuts, never for a cell decision on its own
    double extent_d[D];
    double spacing_d;
    int dims[D];
    bool periodic[D];
    int metric;
    int K;        // fine cells per reference cell and dimension (1 or 2)
    int kb;       // log2(K): fine coordinate = (reference coordinate << kb) | upper-half bit
    int ksh;      // log2(K^D)
    // Slab window along the LAST axis (multi-GPU).  dims[D-1] is the number of layers this handle
    // STORES: nl owned layers, then W upper-halo layers, then W lower-halo layers, so local layer
    // l = (global layer - goff) mod (nl + 2W) and the ordinary periodic index arithmetic on dims[]
    // finds every neighbour.  Positions, extents and cell centres stay GLOBAL (bit-exact pos2cell).
    // division of a device key by the fine row lengths (dims[d] << kb, d < D-1) without a divide instruction
    // (Granlund-Montgomery round-up multiplier): q = (t + ((n - t) >> s1)) >> s2 with t = umulhi(m, n)
    uint32_t div_m[D];
    uint8_t div_s1[D], div_s2[D];
    int gdl;      // global number of layers (== dims[D-1] when W == 0)
    int goff;     // global layer of local layer 0
    int nl;       // owned layers (== dims[D-1] when W == 0)
    int W;        // halo layers on each side; 0 = not windowed
    // Periodicity of the last axis of the SPACE.  periodic[D-1] steers the cell-index arithmetic and is always true on
    // a slab window (the window wraps onto its own halo layers; beyond a non-periodic edge those layers are empty);
    // positions are normalised (walk.jl:77-91) with this flag.
    bool gper_last;
};

template <class T, int D> static inline SpaceDev<T, D> make_space(const ab_handle* h) {
    SpaceDev<T, D> s;
    for (int d = 0; d < D; ++d) {
        s.extent[d] = (T)h->desc.extent[d];
        s.extent_d[d] = (double)s.extent[d];
        s.dims[d] = h->dims[d];
        s.periodic[d] = h->desc.periodic[d] != 0;
    }
    s.spacing = (T)h->desc.spacing;
    s.inv_spacing = T(1) / s.spacing;
    s.spacing_d = (double)s.spacing;
    s.metric = h->desc.kind == AB_SPACE_CONTINUOUS ? AB_METRIC_EUCLIDEAN : h->desc.metric;
    s.K = h->K;
    s.kb = h->kb;
    s.ksh = h->ksh;
    for (int d = 0; d < D; ++d) {
        const uint32_t dv = (uint32_t)h->dims[d] << h->kb;
        int l = 0;
        while ((1ull << l) < dv) ++l;                                   // ceil(log2 dv)
        s.div_m[d] = (uint32_t)(((1ull << 32) * ((1ull << l) - dv)) / dv + 1);
        s.div_s1[d] = (uint8_t)(l < 1 ? l : 1);
        s.div_s2[d] = (uint8_t)(l > 1 ? l - 1 : 0);
    }
    s.W = h->win_W;
    s.gper_last = s.periodic[D - 1];
    if (h->win_W) s.periodic[D - 1] = true;
    s.gdl = h->win_W ? h->win_gdl : h->dims[D - 1];
    s.goff = h->win_W ? h->win_goff : 0;
    s.nl = h->win_W ? h->win_nl : h->dims[D - 1];
    return s;
}

#ifdef __CUDACC__

__device__ __forceinline__ double ab_sqrt(double x) { return sqrt(x); }
__device__ __forceinline__ float ab_sqrt(float x) { return sqrtf(x); }
__device__ __forceinline__ double ab_rsqrt(double x) { return rsqrt(x); }
__device__ __forceinline__ float ab_rsqrt(float x) { return 1.0f / sqrtf(x); }
__device__ __forceinline__ double ab_fmod(double x, double y) { return fmod(x, y); }
__device__ __forceinline__ float ab_fmod(float x, float y) { return fmodf(x, y); }
__device__ __forceinline__ double ab_next(double x, double dir) { return nextafter(x, dir); }
__device__ __forceinline__ float ab_next(float x, float dir) { return nextafterf(x, dir); }

// Julia mod(x::T, y::T) for floats — base/float.jl; used by normalize_position
// (src/spaces/walk.jl:77-79).  Can return exactly y.
template <class T> __device__ __noinline__ T jl_mod_slow(T x, T y) {
    T r;
    if (x >= T(0) && x < y) r = x;
    else r = ab_fmod(x, y);
    if (r == T(0)) return copysign(r, y);
    if ((r > T(0)) != (y > T(0))) return r + y;
    return r;
}
template <class T> __device__ __forceinline__ T jl_mod(T x, T y) {
    // the common Euler step stays inside (0, y): mod(x, y) == x exactly; everything else (wraps, zeros, negative
    // moduli, NaN) takes the out-of-line general form
    if (x > T(0) && x < y) return x;
    return jl_mod_slow<T>(x, y);
}
template <class T> __device__ __forceinline__ T jl_sign(T x) { return x > T(0) ? T(1) : (x < T(0) ? T(-1) : x); }
__device__ __forceinline__ int jl_mod1(int x, int y) {
    int m = x % y;
    if (m < 0) m += y;
    return m == 0 ? y : m;
}

// pos2cell — src/spaces/continuous.jl:112-113: max(1, ceil(Int, p / spacing)),
// returned 0-based.  IEEE division (never p * (1/spacing)).  Sets *bad when the
// reference would throw (NaN/Inf → InexactError) or index out of the grid.
template <class T, int D>
__device__ __forceinline__ bool pos2cell0(const SpaceDev<T, D>& sp, const T* p, int* c) {
    bool ok = true;
#pragma unroll
    for (int d = 0; d < D; ++d) {
        T q = p[d] / sp.spacing;
        T cq = ceil(q);
        if (!(cq <= T(2147483000.0)) || !(cq >= T(-2147483000.0))) { ok = false; cq = T(1); }   // NaN / Inf / huge
        int ci = (int)cq;
        ci = ci < 1 ? 1 : ci;
        const int lim = d == D - 1 ? sp.gdl : sp.dims[d];       // pos2cell is global; dims[D-1] may be a slab window
        if (ci > lim) { ok = false; ci = lim; }
        c[d] = ci - 1;
    }
    return ok;
}
template <int D> __device__ __forceinline__ uint32_t lin_cell(const int* dims, const int* c) {
    uint32_t lin = (uint32_t)c[D - 1];
#pragma unroll
    for (int d = D - 2; d >= 0; --d) lin = lin * (uint32_t)dims[d] + (uint32_t)c[d];
    return lin;
}
template <int D> __device__ __forceinline__ void unlin_cell(const int* dims, uint32_t lin, int* c) {
#pragma unroll
    for (int d = 0; d < D - 1; ++d) {
        uint32_t q = lin / (uint32_t)dims[d];
        c[d] = (int)(lin - q * (uint32_t)dims[d]);
        lin = q;
    }
    c[D - 1] = (int)lin;
}

// Slab window (multi-GPU): local layer <-> global layer of the last axis.
template <class T, int D> __device__ __forceinline__ int local_to_global_last(const SpaceDev<T, D>& sp, int l) {
    if (sp.W == 0) return l;
    int g = (l >= sp.nl + sp.W ? l - (sp.nl + 2 * sp.W) : l) + sp.goff;    // lower-halo layers sit below layer 0
    if (g < 0) g += sp.gdl;
    else if (g >= sp.gdl) g -= sp.gdl;
    return g;
}
template <class T, int D> __device__ __forceinline__ int global_to_local_last(const SpaceDev<T, D>& sp, int g, bool* ok) {
    if (sp.W == 0) return g;
    int l = g - sp.goff;
    if (l < 0) l += sp.gdl;
    if (l < sp.nl + sp.W) return l;                    // owned or upper halo
    l -= sp.gdl;
    if (l >= -sp.W) return l + sp.nl + 2 * sp.W;       // lower halo
    *ok = false;                                       // outside this rank's window
    return 0;
}

// Device cell key: column-major linear index of the agent's FINE cell over the
// fine grid (dims << kb).  Fine coordinate = reference coordinate * K + (position
// above the reference cell's centre — cell_center, continuous.jl:116-117), so
// fine cells nest exactly inside the bit-exact reference cells.
// `c` are GLOBAL reference-cell coordinates (pos2cell0); *inwin (optional) reports whether the cell lies in
// this handle's slab window.
template <class T, int D>
__device__ __forceinline__ uint32_t device_key(const SpaceDev<T, D>& sp, const T* p, const int* c, bool* inwin = nullptr,
                                               uint32_t* frow = nullptr) {
    uint32_t f[D];
    bool ok = true;
#pragma unroll
    for (int d = 0; d < D; ++d) {
        const int cl = d == D - 1 ? global_to_local_last<T, D>(sp, c[d], &ok) : c[d];
        f[d] = (uint32_t)cl << sp.kb;
        if (sp.K >= 2) {
            const T centre = sp.spacing * ((T)(c[d] + 1) - T(0.5));
            const bool up = p[d] > centre;
            if (sp.K == 2) f[d] |= up ? 1u : 0u;
            else {
                // K = 4: halves split at the centre, quarters at centre -+ spacing/4 (any monotone split works:
                // the sweeps only need fine cells no narrower than ~spacing/4, with a 1e-9 margin in the plan)
                const T quarter = up ? centre + sp.spacing * T(0.25) : centre - sp.spacing * T(0.25);
                f[d] |= (up ? 2u : 0u) | (p[d] > quarter ? 1u : 0u);
            }
        }
    }
    uint32_t lin = f[D - 1];
#pragma unroll
    for (int d = D - 2; d >= 0; --d) {
        if (d == 0 && frow) *frow = lin;                     // fine row = key / fd0, fine x = f[0] = key - frow * fd0
        lin = lin * ((uint32_t)sp.dims[d] << sp.kb) + f[d];
    }
    if (inwin) *inwin = ok;
    return lin;
}
// fine coordinates of a device key
template <class T, int D>
__device__ __forceinline__ void key_to_fine(const SpaceDev<T, D>& sp, uint32_t key, int* f) {
#pragma unroll
    for (int d = 0; d < D - 1; ++d) {
        const uint32_t fd = (uint32_t)sp.dims[d] << sp.kb;
        const uint32_t q = key / fd;
        f[d] = (int)(key - q * fd);
        key = q;
    }
    f[D - 1] = (int)key;
}
// the same with multiply-high instead of the divide (hot kernels)
template <class T, int D>
__device__ __forceinline__ void key_to_fine_fast(const SpaceDev<T, D>& sp, uint32_t key, int* f) {
#pragma unroll
    for (int d = 0; d < D - 1; ++d) {
        const uint32_t fd = (uint32_t)sp.dims[d] << sp.kb;
        const uint32_t t = __umulhi(sp.div_m[d], key);
        const uint32_t q = (t + ((key - t) >> sp.div_s1[d])) >> sp.div_s2[d];
        f[d] = (int)(key - q * fd);
        key = q;
    }
    f[D - 1] = (int)key;
}
// reference cell (0-based coordinates) of a device key
template <class T, int D>
__device__ __forceinline__ void key_to_cell(const SpaceDev<T, D>& sp, uint32_t key, int* c) {
    key_to_fine<T, D>(sp, key, c);
#pragma unroll
    for (int d = 0; d < D; ++d) c[d] >>= sp.kb;
}

// grid_r = ceil(Int, (r + δ)/spacing), δ = distance_from_cell_center —
// src/spaces/continuous.jl:116-121, :221-225.  `cell .- 0.5` is Float64 in
// Julia whatever T is, so this is double arithmetic on promoted values.
template <class T, int D>
__device__ __forceinline__ int grid_radius(const SpaceDev<T, D>& sp, const T* p, const int* c0, double r) {
    double s = 0;
#pragma unroll
    for (int d = 0; d < D; ++d) {
        double centre = sp.spacing_d * ((double)(c0[d] + 1) - 0.5);
        double direct = fabs((double)p[d] - centre);
        double m = sp.periodic[d] ? fmin(direct, sp.extent_d[d] - direct) : direct;
        double sq = m * m;
        s = (d == 0) ? sq : s + sq;
    }
    double delta = sqrt(s);
    double g = ceil((r + delta) / sp.spacing_d);
    if (!(g < 1.0e9)) g = 1.0e9;      // r = Inf etc.: clipped by the hyper-rectangle rule anyway
    if (!(g > -1.0e9)) g = -1.0e9;
    return (int)g;
}

// Largest s (of type T) with (double)sqrt_T(s) <= r.  The reference's exact
// filter is `euclidean_distance(...) <= r` = sqrt(Σ m²) <= r
// (src/spaces/continuous.jl:237-238); sqrt is monotone, so comparing the sum of
// squares with this threshold is bit-identical and needs no sqrt per candidate.
template <class T> __device__ __forceinline__ T exact_threshold(double r) {
    if (!(r >= 0.0)) return T(-1);
    if (isinf(r)) return (T)r;
    T s = (T)(r * r);
    const T inf = (T)INFINITY;
    while ((double)ab_sqrt(s) > r) s = ab_next(s, -inf);
    for (;;) {
        T nx = ab_next(s, inf);
        if (nx == s || !((double)ab_sqrt(nx) <= r)) break;
        s = nx;
    }
    return s;
}

// Per-dimension offset range of calculate_hyperrectangle for integer radius r0
// (src/spaces/grid_general.jl:47-89).
template <int D>
__device__ __forceinline__ void hyper_ranges(const int* dims, const bool* periodic, int r0, int* lo, int* hi) {
    bool all_notover = true;
#pragma unroll
    for (int d = 0; d < D; ++d) all_notover = all_notover && (periodic[d] ? (r0 < dims[d] / 2) : (r0 < dims[d]));
#pragma unroll
    for (int d = 0; d < D; ++d) {
        if (all_notover) { lo[d] = -r0; hi[d] = r0; }
        else if (periodic[d]) {
            int half = dims[d] / 2;
            int rm = r0 < half ? r0 : half;
            lo[d] = -rm;
            hi[d] = rm - ((rm == half && (dims[d] % 2) == 0) ? 1 : 0);
        } else {
            int rm = r0 < dims[d] ? r0 : dims[d];
            lo[d] = -rm; hi[d] = rm;
        }
    }
}

__device__ __forceinline__ int isqrt_floor(long long v) {
    if (v <= 0) return 0;
    long long w = (long long)sqrt((double)v);
    while (w * w > v) --w;
    while ((w + 1) * (w + 1) <= v) ++w;
    return (int)w;
}

// Fine rows of one reference cell row (cy, cz): their first fine-cell index.
struct FineRows {
    uint32_t base[4];
    int n;
};
template <int D>
__device__ __forceinline__ FineRows fine_rows(const int* dims, int kb, int cy, int cz) {
    FineRows fr;
    const int K = 1 << kb;
    const uint32_t fd0 = (uint32_t)dims[0] << kb;
    if (D == 2) {
        fr.n = K;
        for (int t = 0; t < K; ++t) fr.base[t] = (((uint32_t)cy << kb) + t) * fd0;
    } else {
        const uint32_t fd1 = (uint32_t)dims[1] << kb;
        fr.n = K * K;
        for (int tz = 0; tz < K; ++tz)
            for (int ty = 0; ty < K; ++ty)
                fr.base[tz * K + ty] = ((((uint32_t)cz << kb) + tz) * fd1 + (((uint32_t)cy << kb) + ty)) * fd0;
    }
    return fr;
}

// Emits the slot runs of the reference cells [a, b] (inclusive, unwrapped x cell
// coordinates, at most dims0 of them) of one reference cell row — one run per
// fine row — in ascending-offset order, wrapping (periodic) or clipping
// (non-periodic) at the x boundary.
struct CellStartPtr {            // cell_start lookup
    const uint32_t* __restrict__ p;
    __device__ __forceinline__ uint32_t operator()(size_t i) const { return p[i]; }
};
struct CellIdentity {            // "slot" = linear cell: the walk then yields runs of cells (dense grids)
    __device__ __forceinline__ uint32_t operator()(size_t i) const { return (uint32_t)i; }
};
template <class CS, class F>
__device__ __forceinline__ void emit_xrange(int dims0, bool periodic0, const CS& cell_start_,
                                            int kb, const FineRows& fr, int a, int b, int wrap, F&& f) {
    auto R = [&](int xa, int xe, int wr) {      // reference cells [xa, xe) of this row
        for (int t = 0; t < fr.n; ++t)
            f(cell_start_((size_t)fr.base[t] + ((uint32_t)xa << kb)), cell_start_((size_t)fr.base[t] + ((uint32_t)xe << kb)), wr);
    };
    if (a > b) return;
    if (!periodic0) {
        if (a < 0) a = 0;
        if (b >= dims0) b = dims0 - 1;
        if (a <= b) R(a, b + 1, wrap);
        return;
    }
    if (a < 0) {
        int e = b < -1 ? b : -1;
        R(a + dims0, e + dims0 + 1, wrap | 1);
        a = 0;
    }
    if (a <= b) {
        int e = b < dims0 - 1 ? b : dims0 - 1;
        if (a <= e) R(a, e + 1, wrap);
        if (b >= dims0) {
            int s2 = a > dims0 ? a : dims0;
            R(s2 - dims0, b - dims0 + 1, wrap | 1);
        }
    }
}

// Walks the cells of offsets_within_radius(space, r0) around cell c0 in the
// reference's order (Iterators.product: first dimension fastest —
// src/spaces/grid_general.jl:138-151) applying the wrap / skip rules of
// GridSpaceIdIterator (src/spaces/grid_multi.jl:203-232).  Cells with
// consecutive x are consecutive in the sorted layout, so the walk yields RUNS:
// f(slot_begin, slot_end, wrapmask) where bit d of wrapmask says the run was
// reached across the periodic boundary of dimension d.  skip_zero drops the
// zero offset (offsets_within_radius_no_0, grid_general.jl:165-179).
template <int D, class CS, class F>
__device__ __forceinline__ void for_each_run_cs(const int* dims, const bool* periodic, int metric, const int* c0,
                                                int r0, const CS& cell_start, int kb, bool skip_zero, F&& f) {
    if (r0 < 0) {   // empty product → calculate_offsets pushes the zero offset (grid_general.jl:149)
        if (skip_zero) return;
        const FineRows fr = fine_rows<D>(dims, kb, c0[1], D == 3 ? c0[2] : 0);
        for (int t = 0; t < fr.n; ++t)
            f(cell_start((size_t)fr.base[t] + ((uint32_t)c0[0] << kb)),
              cell_start((size_t)fr.base[t] + ((uint32_t)(c0[0] + 1) << kb)), 0);
        return;
    }
    int lo[3] = {0, 0, 0}, hi[3] = {0, 0, 0};
    hyper_ranges<D>(dims, periodic, r0, lo, hi);
    const long long r2 = (long long)r0 * r0;
    for (int bz = lo[2]; bz <= hi[2]; ++bz) {      // lo[2] == hi[2] == 0 when D == 2
        int wrapz = 0, cz = 0;
        if (D == 3) {
            cz = c0[2] + bz;
            if (cz < 0 || cz >= dims[2]) {
                if (!periodic[2]) continue;
                cz = cz < 0 ? cz + dims[2] : cz - dims[2];
                wrapz = 4;
            }
        }
        for (int by = lo[1]; by <= hi[1]; ++by) {
            int w;   // x half-width the metric allows on this row
            if (metric == AB_METRIC_EUCLIDEAN) {
                long long rem = r2 - (long long)by * by - (long long)bz * bz;
                if (rem < 0) continue;
                w = isqrt_floor(rem);
            } else if (metric == AB_METRIC_MANHATTAN) {
                w = r0 - (by < 0 ? -by : by) - (bz < 0 ? -bz : bz);
                if (w < 0) continue;
            } else w = 0x3fffffff;
            int xl = lo[0] > -w ? lo[0] : -w;
            int xh = hi[0] < w ? hi[0] : w;
            if (xl > xh) continue;
            int wrap = wrapz;
            int cy = c0[1] + by;
            if (cy < 0 || cy >= dims[1]) {
                if (!periodic[1]) continue;
                cy = cy < 0 ? cy + dims[1] : cy - dims[1];
                wrap |= 2;
            }
            const FineRows fr = fine_rows<D>(dims, kb, cy, cz);
            if (skip_zero && by == 0 && bz == 0 && xl <= 0 && xh >= 0) {
                emit_xrange(dims[0], periodic[0], cell_start, kb, fr, c0[0] + xl, c0[0] - 1, wrap, f);
                emit_xrange(dims[0], periodic[0], cell_start, kb, fr, c0[0] + 1, c0[0] + xh, wrap, f);
            } else {
                emit_xrange(dims[0], periodic[0], cell_start, kb, fr, c0[0] + xl, c0[0] + xh, wrap, f);
            }
        }
    }
}

template <int D, class F>
__device__ __forceinline__ void for_each_run(const int* dims, const bool* periodic, int metric, const int* c0,
                                             int r0, const uint32_t* __restrict__ cell_start, int kb,
                                             bool skip_zero, F&& f) {
    for_each_run_cs<D>(dims, periodic, metric, c0, r0, CellStartPtr{cell_start}, kb, skip_zero, f);
}

// Is cell `ct` visited by the walk above when started from cell `cf` with
// radius r0?  (Used to reproduce the asymmetric pair set of all_pairs!,
// src/spaces/continuous.jl:438-452, without a hash set.)
template <int D>
__device__ __forceinline__ bool cell_reachable(const int* dims, const bool* periodic, int metric, const int* cf,
                                               const int* ct, int r0) {
    int lo[3] = {0, 0, 0}, hi[3] = {0, 0, 0};
    if (r0 < 0) {
        bool same = true;
#pragma unroll
        for (int d = 0; d < D; ++d) same = same && (cf[d] == ct[d]);
        return same;
    }
    hyper_ranges<D>(dims, periodic, r0, lo, hi);
    long long s2 = 0, s1 = 0;
#pragma unroll
    for (int d = 0; d < D; ++d) {
        int diff = ct[d] - cf[d];
        int b = diff;
        bool found = (b >= lo[d] && b <= hi[d]);
        if (!found && periodic[d]) {
            b = diff + dims[d];
            found = (b >= lo[d] && b <= hi[d]);
            if (!found) { b = diff - dims[d]; found = (b >= lo[d] && b <= hi[d]); }
        }
        if (!found) return false;
        s2 += (long long)b * b;
        s1 += b < 0 ? -b : b;
    }
    if (metric == AB_METRIC_EUCLIDEAN) return s2 <= (long long)r0 * r0;
    if (metric == AB_METRIC_MANHATTAN) return s1 <= r0;
    return true;
}

// get_direction — src/spaces/utilities.jl:185-216 (one dimension)
template <class T> __device__ __forceinline__ T min_image(T direct, T L) {
    T inverse = direct - jl_sign(direct) * L;
    return (fabs(direct) <= fabs(inverse)) ? direct : inverse;
}

// normalize_position — src/spaces/walk.jl:77-91 (one dimension)
template <class T> __device__ __forceinline__ T normalize_coord(T x, T L, bool periodic) {
    if (periodic) return jl_mod(x, L);
    T hi = ab_next(L, -(T)INFINITY);            // prevfloat(L)
    return x > hi ? hi : (x < T(0) ? T(0) : x); // clamp
}

#endif  // __CUDACC__
