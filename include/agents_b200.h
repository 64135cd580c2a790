/* agents_b200.h — C ABI of the B200-native Agents.jl hot path.
 *
 * One shared library (libagents_b200.so), plain C types only: this is what a
 * Julia `ccall` (or ctypes / cgo / JNI) binds.  Agents.jl v7.0.3 has no FFI of
 * its own; every entry point below replaces the body of a Julia function of
 * the reference and cites it (paths relative to the Agents.jl tree).  The
 * Julia-side glue a maintainer would add is shown in INTEGRATION.md.
 *
 * Conventions
 *  - every call returns 0 (AB_OK) or a negative ab_status; ab_last_error()
 *    returns the message of the last failing call on that handle.  Nothing
 *    throws or aborts across the ABI.
 *  - agent ids are 1-based and dense (1..n), exactly the Vector / StructVector
 *    container rule of the reference (src/core/model_accessing_API.jl:29-32).
 *  - cell coordinates returned to the caller are 1-based like Julia's.
 *  - host arrays are row-per-agent: n x ncomp values of the column's dtype —
 *    the memory layout of a Julia Vector{SVector{D,T}} / Vector{NTuple{D,T}}.
 *  - a handle owns one CUDA stream; calls are synchronous on return unless
 *    documented otherwise.  A handle is not thread-safe; handles are
 *    independent.  There is no CPU fallback: without a CUDA device
 *    ab_create() fails with AB_ERR_CUDA.
 */
#ifndef AGENTS_B200_H
#define AGENTS_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct ab_handle ab_handle;

typedef enum {
    AB_OK = 0,
    AB_ERR_INVALID = -1,      /* bad argument / unsupported configuration        */
    AB_ERR_CUDA = -2,         /* CUDA runtime error (message has the CUDA text)   */
    AB_ERR_OUTSIDE = -3,      /* "position is outside space extent!" continuous.jl:160 */
    AB_ERR_SPACING = -4,      /* extent not divisible by spacing, continuous.jl:99-101 */
    AB_ERR_OCCUPIED = -5,     /* GridSpaceSingle cell already occupied, grid_single.jl:49 */
    AB_ERR_NUMERIC = -6,      /* NaN/Inf position reached a cell computation (InexactError in Julia) */
    AB_ERR_STATE = -7,        /* call not valid in the current state (e.g. no agents) */
    AB_ERR_COMM = -8          /* NCCL / multi-GPU exchange failure */
} ab_status;

enum { AB_SPACE_CONTINUOUS = 0, AB_SPACE_GRID_MULTI = 1, AB_SPACE_GRID_SINGLE = 2 };
enum { AB_F64 = 0, AB_F32 = 1, AB_I32 = 2, AB_I64 = 3, AB_U8 = 4 };
enum { AB_METRIC_CHEBYSHEV = 0, AB_METRIC_MANHATTAN = 1, AB_METRIC_EUCLIDEAN = 2 };
enum { AB_SEARCH_APPROXIMATE = 0, AB_SEARCH_EXACT = 1 };   /* continuous.jl:212-219 */

/* Mirrors ContinuousSpace{D,P,T} (src/spaces/continuous.jl:4-10, ctor :93-105)
 * and GridSpace / GridSpaceSingle (src/spaces/grid_multi.jl:4-12,
 * grid_single.jl:11-18). */
typedef struct {
    int32_t kind;          /* AB_SPACE_*                                         */
    int32_t D;             /* 2 or 3 on the device                               */
    double extent[3];      /* continuous: extent; ignored for grids              */
    double spacing;        /* continuous: cell spacing; must divide every extent */
    uint8_t periodic[3];   /* per-dimension periodicity (P)                      */
    uint8_t _pad[5];
    int32_t metric;        /* grids: AB_METRIC_*; continuous is always euclidean (continuous.jl:102) */
    int32_t dtype;         /* continuous: AB_F64 or AB_F32 (eltype T)            */
    int64_t grid_dims[3];  /* grids: size; continuous: filled in by ab_create (extent/spacing) */
} ab_space_desc;

/* Per-agent fields of `Bird` (examples/flock.jl:21-28).  Uniform values are
 * passed here; when `per_agent_field` >= 0 it names a registered 6-component
 * column (speed, cohere_factor, separation, separate_factor, match_factor,
 * visual_distance) that overrides them agent by agent. */
typedef struct {
    double speed, cohere_factor, separation, separate_factor, match_factor, visual_distance;
    int32_t search;            /* AB_SEARCH_*: the `search` keyword of nearby_agents   */
    int32_t per_agent_field;   /* -1, or field id of a 6-component column            */
    int32_t kernel;            /* 0 auto, 1 force generic kernel, 2 force tiled kernel */
    int32_t _pad;
} ab_flock_params;

/* `Particle` fields r, k, mass (examples/celllistmap.jl:39-43) are registered
 * 1-component columns; dt / cutoff as in celllistmap.jl:73-78. */
typedef struct {
    double dt, cutoff;
    int32_t field_r, field_k, field_mass;
    int32_t _pad;
} ab_pairforce_params;

/* Schelling rule (docs/src/tutorial.jl:351-376). */
typedef struct {
    int32_t min_to_be_happy;
    int32_t relocate;          /* 0: happiness pass only; 1: unhappy agents move to empty cells */
    uint64_t seed;
} ab_schelling_params;

/* ---- lifecycle ------------------------------------------------------------
 * ab_create replaces ContinuousSpace(...) / GridSpace(...) / GridSpaceSingle(...)
 * + construct_agent_container (src/core/agent_containers.jl:53-61): it checks
 * the spacing rule (continuous.jl:99-101), fills desc->grid_dims and allocates
 * the device-resident SoA container.  Built-in columns: continuous "pos"
 * (field 0, D comps of T) and "vel" (field 1); grids "pos" (field 0, D x i32). */
int ab_create(ab_space_desc* desc, int device, ab_handle** out);
int ab_destroy(ab_handle* h);
const char* ab_last_error(const ab_handle* h);   /* h may be NULL: last ab_create error */
int ab_sync(ab_handle* h);
/* Asynchronous mode (no reference counterpart: Agents.jl is single-task; this is what an
 * ensemblerun!-style driver, src/simulations/ensemblerun.jl, uses to keep several models in flight):
 * ab_upload / ab_download / ab_step_flock / ab_build_index return once their work is queued on
 * the handle's stream — host buffers must be pinned (ab_alloc_pinned) and stay valid — and
 * ab_sync waits and returns the first device-side failure since the previous ab_sync. */
int ab_set_async(ab_handle* h, int on);
int ab_version(void);
int64_t ab_nagents(const ab_handle* h);

/* pinned host memory for the Julia-side mirror columns (optional, faster copies) */
void* ab_alloc_pinned(int64_t bytes);
void ab_free_pinned(void* p);

/* ---- columns ---------------------------------------------------------------
 * One device column per isbits field of the @agent struct
 * (src/core/agents.jl:199-231; SoA precedent src/core/agent_containers.jl:8-35). */
int ab_register_field(ab_handle* h, const char* name, int dtype, int ncomp);  /* returns field id >= 0 */
int ab_field_id(const ab_handle* h, const char* name);
/* add_agent! at setup (space_interaction_API.jl:174-179 → push!): (re)sets the
 * population to n agents with ids 1..n.  Columns must then be uploaded. */
int ab_resize(ab_handle* h, int64_t n);
int ab_upload(ab_handle* h, int field, const void* host, int64_t n);    /* id order in */
int ab_download(ab_handle* h, int field, void* host, int64_t n);        /* id order out */

/* ---- index -----------------------------------------------------------------
 * Replaces the persistent array-of-vectors index and its incremental upkeep
 * (continuous.jl:133-153, grid_multi.jl:91-105, grid_single.jl:47-57) by a
 * per-step cell list: key = column-major cell (pos2cell, continuous.jl:112-113),
 * counting sort, cell start offsets.  Called implicitly by queries/steps when
 * positions changed.  Fails with AB_ERR_OUTSIDE / AB_ERR_NUMERIC / AB_ERR_OCCUPIED
 * exactly where the reference would throw. */
int ab_build_index(ab_handle* h);
/* Device-grid refinement: K = 1 or 2 sub-cells per reference cell and dimension
 * (default 1; the fused steps pick it themselves).  A pure layout choice —
 * neighbour sets and cells are defined on the reference cells either way. */
int ab_set_refinement(ab_handle* h, int K);
/* Checkpoint support: AgentsIO.save_checkpoint / load_checkpoint (src/submodules/io/jld2_integration.jl:85-135,
 * :196-223) serialise the agents AND the space's stored_ids (the order of the agents inside every cell), so a
 * restored model continues exactly.  The columns travel through ab_download / ab_upload; these two calls carry the
 * rest of the device state a bit-for-bit continuation needs: the slot order (1-based ids in device order; on grid
 * handles 1..n), the device-grid refinement and the step counter that decorrelates the Schelling relocations.
 * ab_layout_set is called after the columns of the restored model have been uploaded; it rebuilds the index with
 * the agents of every cell in the saved order. */
int ab_layout_get(ab_handle* h, int64_t* slot_ids, int64_t n, int32_t* refinement, int64_t* tick);
/* 1 while every index build of this handle ordered every cell canonically (by id / by the restored order): the premise
 * of bitwise reproducible runs, N-GPU == 1-GPU equality and bit-for-bit resume.  0 once a cell held more than 4096
 * agents at an index BUILD (it keeps the arrival order of the atomics; results stay correct within rounding).  The
 * reference's `stored_ids` vectors (grid_multi.jl:68-105) are ordered by insertion history, which the device does not
 * keep. */
int ab_index_canonical(ab_handle* h);
int ab_layout_set(ab_handle* h, const int64_t* slot_ids, int64_t n, int32_t refinement, int64_t tick);

/* pos2cell for every agent, id order, n x D int32, 1-based (continuous.jl:112-113) */
int ab_cells(ab_handle* h, int32_t* out_cells, int64_t n);

/* ---- neighbour queries -------------------------------------------------------
 * nearby_ids(pos, model, r; search) continuous.jl:212-239 and
 * nearby_ids(agent, model, r) space_interaction_API.jl:284-287; for grids
 * grid_multi.jl:120-128, grid_single.jl:92-205.  Exactly one of qpos / qids is
 * non-NULL: qpos = nq x D positions (T for continuous, int32 1-based for
 * grids); qids = nq agent ids (agent form; exclude_self applies to it).
 * Two-phase so the caller owns the result: counts, then fill at caller-made
 * exclusive offsets (offsets[nq] = total).  ids come back 1-based int64. */
int ab_nearby_count(ab_handle* h, const void* qpos, const int64_t* qids, int64_t nq, double r,
                    int search, int exclude_self, int64_t* counts);
int ab_nearby_fill(ab_handle* h, const void* qpos, const int64_t* qids, int64_t nq, double r,
                   int search, int exclude_self, const int64_t* offsets, int64_t* ids_out);

/* interacting_pairs(model, r, :all; search) continuous.jl:418-452: unique
 * (lo, hi) id pairs; count then fill (fill re-runs the last counted query). */
/* nearby_positions(pos, model, r) grid_general.jl:185-232 and empty_nearby_positions discrete.jl:116-129 on
 * GridSpace / GridSpaceSingle handles: for nq query positions (nq x D int32, 1-based) the positions of the
 * offset table of radius r around each (origin excluded; wrapped on periodic axes, skipped outside the grid
 * otherwise), optionally only the empty ones.  Count, then fill (CSR offsets from the counts; nq x D rows out). */
int ab_nearby_positions_count(ab_handle* h, const int32_t* qpos, int64_t nq, double r, int empty_only, int64_t* counts);
int ab_nearby_positions_fill(ab_handle* h, const int32_t* qpos, int64_t nq, double r, int empty_only,
                             const int64_t* offsets, int32_t* pos_out);

int ab_pairs_count(ab_handle* h, double r, int search, int64_t* npairs);
int ab_pairs_fill(ab_handle* h, int64_t* lo, int64_t* hi, int64_t npairs);

/* nearest_neighbor(agent, model, r) continuous.jl:298-307 for all agents:
 * out[i] = id of the nearest agent within r, or 0 (`nothing`). */
int ab_nearest_neighbor(ab_handle* h, double r, int64_t* out, int64_t n);

/* interacting_pairs(model, r, :nearest; scheduler) continuous.jl:418-436 -> true_pairs! :454-497:
 * every scheduled agent proposes (itself, its nearest neighbour within r); an id keeps only its
 * closest pair.  `order` = the scheduler's id sequence (NULL: ByID, 1..n) — the result depends on
 * it exactly as in the reference (docstring, continuous.jl:400-405).  Count, then fill (list order =
 * the reference's `pairs` vector). */
int ab_pairs_nearest(ab_handle* h, double r, const int64_t* order, int64_t norder, int64_t* npairs);
int ab_pairs_nearest_fill(ab_handle* h, int64_t* lo, int64_t* hi, int64_t npairs);

/* elastic_collision!(a, b, f) continuous.jl:333-366 applied to the pairs (a[k], b[k]) in list order
 * (2-D only, like the reference).  mass_field = a scalar f64/f32 column id (the `f` symbol, may hold
 * Inf: immovable agents) or -1 for equal masses.  happened[k] (optional) = the call's return value,
 * *ncollided = how many returned true.  An agent with infinite mass and nonzero velocity fails with
 * AB_ERR_INVALID (the reference's @assert, :350 / :355). */
int ab_elastic_collisions(ab_handle* h, const int64_t* a, const int64_t* b, int64_t npairs, int mass_field,
                          uint8_t* happened, int64_t* ncollided);

/* get_spatial_index / get_spatial_property, continuous.jl:536-573, for all agents (qpos == NULL, rows in
 * id order) or for nq query positions (nq x D of the space's element type).  `property` is a column-major
 * array of prop_dims[0..pd) (pd <= D: the front dimensions of pos index it) with ncomp components stored as
 * ncomp consecutive arrays.  out_index (optional): nq x pd, 1-based; out_value (optional): nq x ncomp.
 * An index outside the array (the reference's BoundsError, pos == extent) fails with AB_ERR_OUTSIDE. */
int ab_spatial_lookup(ab_handle* h, const int64_t* prop_dims, int pd, const double* property, int ncomp,
                      const void* qpos, int64_t nq, int64_t* out_index, double* out_value);
/* The force-field form of the `update_vel!` hook (continuous.jl:86-91; called by move_agent!(agent, model, dt)
 * :186-194 before the move): vel += scale * field[get_spatial_index(pos)] for the scheduled agents
 * (active_ids == NULL: all).  `field` has D components over prop_dims. */
int ab_update_vel_field(ab_handle* h, const int64_t* prop_dims, int pd, const double* field, double scale,
                        const int64_t* active_ids, int64_t nactive);

/* ---- movement ---------------------------------------------------------------
 * move_agent!(agent, model, dt) continuous.jl:186-194 → walk! walk.jl:58-66 →
 * normalize_position walk.jl:77-91, for `active_ids` (NULL = all agents).  The ids of a list must be
 * unique (one thread per entry; the reference would move a repeated agent twice — call twice for that). */
int ab_move(ab_handle* h, double dt, const int64_t* active_ids, int64_t nactive);
/* walk!(agent, direction, model): per-agent displacement, nactive x D of T
 * (continuous) or int32 (grids, walk.jl:33-56 with ifempty).  With active_ids == NULL the rows are
 * in id order and nactive must be the agent count (or 0): AB_ERR_INVALID otherwise. */
int ab_walk(ab_handle* h, const void* directions, const int64_t* active_ids, int64_t nactive, int ifempty);
/* randomwalk!(agent, model, r) for the agents of `active_ids` (all when NULL), src/spaces/walk.jl:154-214 (grids)
 * and :389-408 (ContinuousSpace, uniform_randomwalk!).  The reference draws from the model's RNG; its stream is not
 * reproduced: draw c of agent id is u = (splitmix64(seed ^ ((id << 16) + c)) >> 11) * 2^-53, so a call is a pure
 * function of (state, seed) and the caller passes a fresh seed per tick.
 *   ContinuousSpace: direction = v / |v| * r with v = D standard normals (Box-Muller on draws 0..3); r < 0 takes
 *     the agent's own speed (the reference's default r); vel = direction; then walk!.  r == 0 (or a resting agent
 *     with the default r) fails with AB_ERR_INVALID like the reference's ArgumentError.  ifempty / force_motion ignored.
 *   GridSpace / GridSpaceSingle: one entry of offsets_at_radius(model, r) (grid_general.jl:119-136; floor(r) steps under
 *     :chebyshev / :manhattan, :euclidean fails with AB_ERR_INVALID), index floor(u0 * L); with force_motion (and
 *     ifempty, or a GridSpaceSingle) the choice is made among the empty targets (walk.jl:200-214) in the sequential
 *     order of `active_ids`.  force_motion = 1: the agent moves to the chosen empty position (the docstring's
 *     behaviour); force_motion = 2: the reference's code as written — the chosen POSITION is handed to walk! as a
 *     direction (walk.jl:170-172, :193-195), i.e. target = normalize_position(pos .+ choice), taken only if empty. */
int ab_randomwalk(ab_handle* h, double r, uint64_t seed, const int64_t* active_ids, int64_t nactive, int ifempty,
                  int force_motion);
/* randomwalk!(agent, model[, r]; polar, azimuthal) on a ContinuousSpace{2} / {3} with angle distributions
 * (src/spaces/walk.jl:236-330): the host draws one polar angle (and in 3-D one azimuthal angle) per walking agent from
 * whatever distribution it likes; the device turns the agent's velocity by them (rotate, walk.jl:332-357), scales it to
 * length r (r < 0: keeps its length), stores it as the new velocity and walks.  A resting agent has no direction to
 * rotate: AB_ERR_INVALID. */
int ab_randomwalk_angles(ab_handle* h, double r, const double* polar, const double* azimuthal, const int64_t* active_ids,
                         int64_t nactive);

/* ---- fused rules (synchronous update) ---------------
 * step!(model, n) with a B200Step model_step! (step_standard.jl:24-31). */
int ab_step_flock(ab_handle* h, const ab_flock_params* p, int64_t nsteps);           /* examples/flock.jl:79-108 */
/* The same step for a scheduled subset (Schedulers.Partially or any id list from schedule(model),
 * src/submodules/schedulers.jl:8, 68-120): only the agents of `active_ids` update; the others keep their state and
 * remain neighbours of the rest (synchronous update: everybody reads the state at the start of the tick).  The list
 * applies to each of the nsteps ticks; a scheduler that draws a new subset per tick calls with nsteps = 1.
 * NULL = every agent (ab_step_flock).  Not available on a sharded handle. */
int ab_step_flock_active(ab_handle* h, const ab_flock_params* p, int64_t nsteps, const int64_t* active_ids, int64_t nactive);
int ab_step_pairforce(ab_handle* h, const ab_pairforce_params* p, int64_t nsteps);   /* examples/celllistmap.jl:143-191 */
/* Scheduled subset for the pair-force rule: model_step! computes the forces of every particle, agent_step! (the
 * integrator, celllistmap.jl:176-191) runs for the agents of `active_ids`; the others keep position and velocity. */
int ab_step_pairforce_active(ab_handle* h, const ab_pairforce_params* p, int64_t nsteps, const int64_t* active_ids, int64_t nactive);
/* Schelling on GridSpaceSingle.  Columns "group" (i32) and "mood" (u8) must be
 * registered.  n_moved (optional) receives the movers of the last step. */
int ab_step_schelling(ab_handle* h, const ab_schelling_params* p, int64_t nsteps, int64_t* n_moved);
/* Scheduled subset for the Schelling rule (schedulers.jl:68-120): only the agents of `active_ids` run schelling_step!
 * (docs/src/tutorial.jl:351-376); the others keep their mood and position and stay neighbours of everyone. */
int ab_step_schelling_active(ab_handle* h, const ab_schelling_params* p, int64_t nsteps, const int64_t* active_ids,
                             int64_t nactive, int64_t* n_moved);

/* neighbour count N seen by every bird in the last flock step (id order);
 * diagnostic used by the parity tests (neighbour-set flips). */
int ab_flock_last_counts(ab_handle* h, int32_t* counts, int64_t n);

/* ---- data collection ---------------------------------------------------------
 * run!(model, n; adata = [(field, aggregator[, filter])]) -> collect_agent_data!
 * (src/simulations/collect.jl:154-183, 466-502): each triple is reduced on the
 * device; one double (+ the number of agents that passed the filter) comes back
 * per triple instead of a column download per collected tick.  Plain symbols in
 * adata (whole columns) use ab_download.  mean over no agent returns NaN. */
enum { AB_RED_SUM = 0, AB_RED_MEAN = 1, AB_RED_MIN = 2, AB_RED_MAX = 3, AB_RED_COUNT = 4 };
enum { AB_CMP_NONE = 0, AB_CMP_EQ = 1, AB_CMP_NE = 2, AB_CMP_LT = 3, AB_CMP_LE = 4, AB_CMP_GT = 5, AB_CMP_GE = 6 };
typedef struct {
    int32_t field, comp;               /* column and component to aggregate            */
    int32_t op;                        /* AB_RED_*                                     */
    int32_t filter_field, filter_comp; /* filter: column component <cmp> filter_value  */
    int32_t filter_cmp;                /* AB_CMP_* (AB_CMP_NONE: every agent)          */
    double filter_value;
} ab_reduce_spec;
int ab_reduce(ab_handle* h, const ab_reduce_spec* specs, int nspec, double* out, int64_t* counts_or_null);
/* Schedulers.ByProperty(field)(model) — src/submodules/schedulers.jl:130-151: the 1-based agent ids in `sortperm`
 * order of component `comp` of a column (ascending, stable: ties keep id order; NaN last; `descending` reverses the
 * result).  Sorted on the device (bitonic network over (key, id) pairs); n ids come back; n <= 2^30.
 * Ties: always by ascending id.  The reference keeps ties in the order of its `sched.ids` — the container's key order on
 * the first call (id order for the Vector containers this ABI mirrors, arbitrary for Dict-backed models) and, on later
 * calls, the PREVIOUS permutation (`sortperm!(...; initialized = true)`, schedulers.jl:144-149) — so agents with equal
 * property values can come out in a different relative order there from the second call on. */
int ab_ids_by_property(ab_handle* h, int field, int comp, int descending, int64_t* out_ids, int64_t n);

/* The chunked reduction of ab_reduce for ONE (column, aggregator, filter) triple, for callers that must not
 * allocate device memory.  It has no counterpart in the reference (Agents.jl aggregates on the host heap,
 * collect.jl:466-502) and exists for hosts that cannot afford a cudaMalloc: a Julia session whose CUDA.jl pool (or a
 * Python process whose torch allocator) already owns almost the whole GPU, so an allocation made behind the pool's
 * back may fail while the pool still has room; an ensemble driver whose other models would all stall on the
 * device-wide synchronisation a cudaMalloc / cudaFree implies; a process inside a fixed memory budget (MIG slice,
 * MPS memory limit).  Such a caller cannot create a handle either (ab_create and a handle's scratch allocate), so
 * this call takes no handle:
 *   d_col        n values of `dtype` (AB_F64 .. AB_U8) in DEVICE memory of `device`, from the caller's own pool;
 *   d_filter_col n values of `filter_dtype`, or NULL with filter_cmp = AB_CMP_NONE;
 *   d_out        one double in device memory: the aggregate (mean over no value: NaN);
 *   d_count      one int64 in device memory or NULL: how many values passed the filter;
 *   stream       a cudaStream_t of that device (NULL: the default stream).  The work is queued there and the call
 *                returns without waiting; the results are valid in stream order.
 * The per-chunk partial results sit in a fixed 256 KiB array inside the library image — every GPU that loads the
 * library holds it until the process exits — which bounds n at 2^26 = 67 108 864 values per call; a longer column is
 * refused with AB_ERR_INVALID (split it, or use ab_reduce).  The call itself performs no cudaMalloc, cudaFree or any
 * other allocation.  All calls on one GPU share that array: calls that may overlap on the device (different streams
 * without an event between them) race and must be ordered by the caller; nothing in it needs clearing between
 * calls.  The first call on each device looks up the array's device address — make that call before any stream
 * capture begins.  Failure text: ab_last_error(NULL).  Same value as ab_reduce up to the order of the sum (ab_reduce
 * walks the handle's slot order, this walks the caller's array order). */
int ab_reduce_static(int device, void* stream, const void* d_col, int dtype, int64_t n, int op,
                     const void* d_filter_col, int filter_dtype, int filter_cmp, double filter_value,
                     double* d_out, int64_t* d_count);

/* ---- multi-GPU slabs (one process per GPU) -----------------------------------
 * The reference is single-task (step_standard.jl:9-23); this is its scale-out
 * (SURVEY 8e).  Slabs of whole cell layers along the LAST axis (periodic); each
 * rank stores its owned layers plus 3 halo layers per side and exchanges the
 * boundary layers (halo refresh + migration in one message per neighbour) over
 * NCCL after every fused step.  Agents carry GLOBAL ids, so an N-GPU run is
 * bitwise identical to the 1-GPU run.
 *   id = ab_comm_unique_id() on rank 0, broadcast by the host (torch.distributed /
 *   MPI / Julia Distributed); on every rank: ab_create(global space) ->
 *   ab_comm_init -> ab_comm_set_capacity -> ab_resize(capacity) ->
 *   ab_upload_local(pos), ab_upload_local(vel) -> ab_step_flock ... ->
 *   ab_download_local.  Queries addressed by agent id are not available on a
 *   sharded handle (AB_ERR_STATE). */
int ab_comm_unique_id(void* out128);                       /* 128 bytes (ncclUniqueId) */
/* nccl_unique_id may be NULL: window layout only, no communicator (tests) */
int ab_comm_init(ab_handle* h, int rank, int nranks, const void* nccl_unique_id);
/* Slabs of ONE process (e.g. a Julia session driving several GPUs, or several slabs on one GPU):
 * handles created with ab_comm_init(h, r, n, NULL), passed in rank order.  Neighbours exchange by
 * device-to-device copies ordered with CUDA events; step them with ab_step_flock_group. */
int ab_comm_link_local(ab_handle** handles, int n);
int ab_step_flock_group(ab_handle** handles, int n, const ab_flock_params* p, int64_t nsteps);
/* capacity (agents) of one boundary message; every rank must pass the same value */
int ab_comm_set_capacity(ab_handle* h, int64_t halo_agents);
/* Direct halo exchange over peer memory (NVLink), replacing the ncclSend/ncclRecv group of the step:
 * after ab_comm_set_capacity every rank exports the 64-byte CUDA IPC handle of its receive block, the
 * host passes the handles around (torch.distributed / MPI / any transport) and every rank imports its
 * lower and upper neighbour's.  From then on the pack kernel of a step stores the boundary layers
 * straight into the neighbours' memory — only the agents present travel — and a flag per message
 * carries the exchange number; NCCL stays the fallback for handles that did not import. */
int ab_comm_ipc_export(ab_handle* h, void* out64);
int ab_comm_ipc_import(ab_handle* h, const void* lower64, const void* upper64);
int ab_comm_direct(const ab_handle* h);                    /* 1 when the direct exchange is active */
/* Timed multi-rank runs: ab_comm_align queues a one-word all-reduce on the handle's stream, so the calls queued
 * after it start on every rank once the slowest rank has arrived ON THE DEVICE; ab_comm_wait_stats returns the
 * time k_halo_unpack spent waiting for the neighbours' flags (direct exchange) and the exchanges counted since
 * the last reset.  Neither has a counterpart in the reference (single task, step_standard.jl:9-23). */
int ab_comm_align(ab_handle* h);
/* Collective halo refresh outside a step: after ab_upload_local replaced the owned agents from the host, rebuild the
 * window's halo layers from the neighbours (index build + the step's own pack / transfer / unpack). */
int ab_comm_exchange(ab_handle* h);
int ab_comm_wait_stats(ab_handle* h, double* wait_us, int64_t* exchanges, int reset);
/* owned layers of this rank: global index of the first one (0-based) and how many */
int ab_comm_slab(const ab_handle* h, int64_t* first_layer, int64_t* n_layers);
int ab_comm_halo(const ab_handle* h);                      /* halo layers per side */
/* owned agents of this rank (migration changes it; the sum over ranks is conserved) */
int64_t ab_nlocal(const ab_handle* h);
/* rows -> slots of this rank's window (owned + halo layers), global ids 1-based.
 * field 0 (pos) defines the rows; other fields follow with the same rows. */
int ab_upload_local(ab_handle* h, int field, const void* host, const int64_t* global_ids, int64_t n);
/* owned agents in slot order with their global ids; cap = rows available in host / ids_out */
int ab_download_local(ab_handle* h, int field, void* host, int64_t* ids_out, int64_t cap, int64_t* n_out);

/* ---- instrumentation ----------------------------------------------------------
 * kernels launched by this handle so far (bench.py's gpu_launches) and device
 * time of the last step call measured with CUDA events on the handle's stream. */
int64_t ab_launch_count(const ab_handle* h);
double ab_last_step_ms(const ab_handle* h);
/* per-kernel device time of the last ab_step_* call when profiling is on:
 * names[i] (static strings), ms[i] summed over the steps; returns how many. */
/* Measured FP64 FMA throughput of `device` in TFLOP/s (independent fma chains on every SM, CUDA events, best of 3): the
 * denominator bench.py reports the FP64-heavy pair sweeps against.  No counterpart in the reference. */
int ab_fp64_peak(int device, double* tflops);
int ab_set_profiling(ab_handle* h, int on);
int ab_kernel_times(ab_handle* h, const char** names, double* ms, int64_t* launches, int cap);

#ifdef __cplusplus
}
#endif
#endif /* AGENTS_B200_H */
