// oracle/abm_oracle.cpp — CPU ORACLE (test infrastructure, NOT product code).
//
// A from-scratch CPU restatement of the Agents.jl (v7.0.3) hot path that this
// repository re-implements in CUDA: radius neighbour search in ContinuousSpace
// / GridSpace / GridSpaceSingle followed by the per-agent rules (flock, pair
// force, Schelling).  It exists only so that tests/, __graft_entry__.smoke()
// and bench.py's cpu_baseline leg can check / time the GPU path against the
// reference's *semantics*.  Nothing under agents.jl_b200/ may link or call it.
//
// Parity pinning: the reference is pure Julia and Julia is not installed in
// this image, so the reference itself cannot be executed here.  The oracle is
// pinned against every golden vector the reference's own tests hold for this
// path (tests/test_oracle_golden.py cites each test file:line).  Multi-step
// flock / pair-force / Schelling trajectories are NOT pinned by any reference
// test ("parity unpinned" for those).
//
// Data structures deliberately mirror the reference (one std::vector of ids
// per cell, swap-pop removal, offset tables in Iterators.product order) so
// that iteration order and floating-point summation order match what the
// Julia code would do.  Build with -ffp-contract=off: Julia never contracts
// a*b+c into an FMA on this path.
//
// Every function cites the reference file:line it restates
// (paths relative to the Agents.jl tree).

#include <atomic>
#include <mutex>
#include <memory>
#include <thread>
#include <algorithm>
#include <array>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <limits>
#include <string>
#include <unordered_set>
#include <vector>

namespace {

// ---------------------------------------------------------------------------
// Julia numeric primitives
// ---------------------------------------------------------------------------

// Julia Base.mod(x::T, y::T) for floats (base/float.jl): rem then sign fix.
// Can return exactly y for tiny negative x (pinned by
// test/continuous_space_tests.jl:52-58).
template <class T> T jl_mod(T x, T y) {
    T r = std::fmod(x, y);
    if (r == T(0)) return std::copysign(r, y);
    if ((r > T(0)) != (y > T(0))) return r + y;
    return r;
}
// Julia mod / mod1 on Int (floored).
inline int64_t jl_imod(int64_t x, int64_t y) {
    int64_t m = x % y;
    if (m != 0 && ((m < 0) != (y < 0))) m += y;
    return m;
}
inline int64_t jl_mod1(int64_t x, int64_t y) {
    int64_t m = jl_imod(x, y);
    return m == 0 ? y : m;
}
template <class T> T jl_sign(T x) { return x > T(0) ? T(1) : (x < T(0) ? T(-1) : x); }
template <class T> T jl_clamp(T x, T lo, T hi) { return x > hi ? hi : (x < lo ? lo : x); }
template <class T> T jl_prevfloat(T x) { return std::nextafter(x, -std::numeric_limits<T>::infinity()); }

// ---------------------------------------------------------------------------
// Offset tables — src/spaces/grid_general.jl:47-89 (calculate_hyperrectangle),
// :138-151 (calculate_offsets), :98-110 / :165-179 (cached accessors).
// metric: 0 chebyshev, 1 manhattan, 2 euclidean.  Product order: first
// dimension fastest (Iterators.product).
// ---------------------------------------------------------------------------
using Off = std::array<int64_t, 3>;

std::vector<Off> calculate_offsets(int D, const int64_t* size, const uint8_t* periodic,
                                   int metric, int64_t r) {
    // grid_general.jl:72-89 (the mixed-periodicity method; the all-true and
    // all-false methods :47-71 reduce to the same ranges).
    bool all_notover = true;
    for (int d = 0; d < D; ++d) {
        bool ok = periodic[d] ? (r < size[d] / 2) : (r < size[d]);
        all_notover = all_notover && ok;
    }
    int64_t lo[3] = {0, 0, 0}, hi[3] = {0, 0, 0};
    for (int d = 0; d < D; ++d) {
        if (all_notover) { lo[d] = -r; hi[d] = r; continue; }
        if (periodic[d]) {
            int64_t half = size[d] / 2, odd = size[d] % 2;
            int64_t rm = std::min(r, half);
            lo[d] = -rm;
            hi[d] = rm - ((rm == half && odd == 0) ? 1 : 0);
        } else {
            int64_t rm = std::min(r, size[d]);
            lo[d] = -rm; hi[d] = rm;
        }
    }
    std::vector<Off> out;
    Off b{0, 0, 0};
    // first dimension fastest
    for (int64_t z = lo[2]; z <= hi[2]; ++z)
        for (int64_t y = lo[1]; y <= hi[1]; ++y)
            for (int64_t x = lo[0]; x <= hi[0]; ++x) {
                b = {x, y, z};
                bool keep;
                if (metric == 2) keep = (x * x + y * y + z * z) <= r * r;       // :141
                else if (metric == 1) keep = (std::llabs(x) + std::llabs(y) + std::llabs(z)) <= r;  // :143
                else keep = true;                                                // :145
                if (keep) out.push_back(b);
            }
    if (out.empty()) out.push_back(Off{0, 0, 0});  // :149
    return out;
}

// offsets_within_radius(space, r): r0 = floor(Int, r+1) - 1 (grid_general.jl:98-110)
std::vector<Off> offsets_within_radius(int D, const int64_t* size, const uint8_t* periodic,
                                       int metric, double r, bool no_zero) {
    int64_t i = (int64_t)std::floor(r + 1.0);
    int64_t r0 = i - 1;
    std::vector<Off> o = calculate_offsets(D, size, periodic, metric, r0);
    if (no_zero) {  // :165-179
        o.erase(std::remove_if(o.begin(), o.end(),
                               [](const Off& b) { return b[0] == 0 && b[1] == 0 && b[2] == 0; }),
                o.end());
    }
    return o;
}

// combine_positions + bounds part of invalid_access
// (src/spaces/grid_multi.jl:203-232; grid_single.jl:128-165, :170-194).
// Returns false when the cell must be skipped (non-periodic dim out of bounds).
inline bool combine_position(int D, const int64_t* size, const uint8_t* periodic,
                             const int64_t* origin, const Off& b, int64_t* out) {
    for (int d = 0; d < D; ++d) {
        int64_t p = origin[d] + b[d];
        if (p < 1 || p > size[d]) {
            if (!periodic[d]) return false;
            p = jl_mod1(p, size[d]);
        }
        out[d] = p;
    }
    return true;
}

inline int64_t linear_index(int D, const int64_t* size, const int64_t* c /*1-based*/) {
    int64_t lin = 0, stride = 1;
    for (int d = 0; d < D; ++d) { lin += (c[d] - 1) * stride; stride *= size[d]; }
    return lin;  // column-major, like Julia Array indexing
}

// ---------------------------------------------------------------------------
// ContinuousSpace model — src/spaces/continuous.jl
// ---------------------------------------------------------------------------
struct FlockParams {       // examples/flock.jl:21-28 (per-agent `const` fields)
    double speed, cohere_factor, separation, separate_factor, match_factor, visual_distance;
};

struct ModelBase {
    virtual ~ModelBase() {}
    std::string err;
};

template <class T> struct CModel : ModelBase {
    int D = 2;
    T extent[3] = {0, 0, 0};
    T spacing = 0;
    uint8_t periodic[3] = {1, 1, 1};
    int64_t dims[3] = {1, 1, 1};
    std::vector<T> pos, vel;                       // n*D, id order (id = index+1)
    std::vector<std::vector<int64_t>> stored_ids;  // per cell, column-major
    int64_t n = 0;
    // space.offsets_within_radius: tables cached by floor(r+1) like the reference
    // (grid_general.jl:98-110) so the timed CPU baseline does not rebuild them per query
    // The workers of flock_step_sync call this concurrently on a fresh model: the table of a radius is built exactly
    // once (call_once) into a slot that never moves (the outer vector is sized up front, never resized).
    static constexpr size_t OFFS_SLOTS = 4097;
    mutable std::vector<std::vector<Off>> offs_cache = std::vector<std::vector<Off>>(OFFS_SLOTS);
    mutable std::unique_ptr<std::once_flag[]> offs_once{new std::once_flag[OFFS_SLOTS]};
    const std::vector<Off>& cached_offsets(int64_t gr) const {
        size_t i = (size_t)std::max<int64_t>(gr, 0);
        if (gr < 0 || i >= OFFS_SLOTS) { static thread_local std::vector<Off> tmp; tmp = offsets_within_radius(D, dims, periodic, 2, (double)gr, false); return tmp; }
        std::call_once(offs_once[i], [&] { offs_cache[i] = offsets_within_radius(D, dims, periodic, 2, (double)gr, false); });
        return offs_cache[i];
    }

    // continuous.jl:112-113  pos2cell = max.(1, ceil.(Int, pos ./ spacing))
    void pos2cell(const T* p, int64_t* c) const {
        for (int d = 0; d < D; ++d) {
            T q = p[d] / spacing;
            int64_t ci = (int64_t)std::ceil(q);
            c[d] = std::max<int64_t>(1, ci);
        }
    }
    bool cell_inbounds(const int64_t* c) const {
        for (int d = 0; d < D; ++d) if (c[d] < 1 || c[d] > dims[d]) return false;
        return true;
    }
    // utilities.jl:84-116 euclidean_distance (T arithmetic, dimension order)
    T euclid(const T* a, const T* b) const {
        T s = 0;
        for (int d = 0; d < D; ++d) {
            T direct = std::fabs(a[d] - b[d]);
            T m = periodic[d] ? std::min(direct, extent[d] - direct) : direct;
            T sq = m * m;
            s = (d == 0) ? sq : s + sq;
        }
        return std::sqrt(s);
    }
    // continuous.jl:116-121: centre = spacing .* (cell .- 0.5) — `cell .- 0.5`
    // is Float64 in Julia whatever T is, so δ and grid_r are Float64 math on
    // promoted values (identical to T math when T == double).
    double dist_from_cell_center(const T* p) const {
        int64_t c[3];
        pos2cell(p, c);
        double s = 0;
        for (int d = 0; d < D; ++d) {
            double centre = (double)spacing * ((double)c[d] - 0.5);
            double direct = std::fabs((double)p[d] - centre);
            double m = periodic[d] ? std::min(direct, (double)extent[d] - direct) : direct;
            double sq = m * m;
            s = (d == 0) ? sq : s + sq;
        }
        return std::sqrt(s);
    }
    // continuous.jl:221-230 grid_r = ceil(Int, (r + δ)/spacing)
    int64_t grid_r(const T* p, double r) const {
        double delta = dist_from_cell_center(p);
        return (int64_t)std::ceil((r + delta) / (double)spacing);
    }
    // utilities.jl:185-216 get_direction
    void get_direction(const T* from, const T* to, T* out) const {
        for (int d = 0; d < D; ++d) {
            T direct = to[d] - from[d];
            if (periodic[d]) {
                T inverse = direct - jl_sign(direct) * extent[d];
                out[d] = (std::fabs(direct) <= std::fabs(inverse)) ? direct : inverse;
            } else out[d] = direct;
        }
    }
    // walk.jl:77-91 normalize_position
    void normalize_position(T* p) const {
        for (int d = 0; d < D; ++d) {
            if (periodic[d]) p[d] = jl_mod(p[d], extent[d]);
            else p[d] = jl_clamp(p[d], T(0), jl_prevfloat(extent[d]));
        }
    }
    // continuous.jl:133-153
    void add_to_space(int64_t id, const int64_t* c) { stored_ids[linear_index(D, dims, c)].push_back(id); }
    bool remove_from_space(int64_t id, const int64_t* c) {
        auto& prev = stored_ids[linear_index(D, dims, c)];
        auto it = std::find(prev.begin(), prev.end(), id);   // findfirst
        if (it == prev.end()) { err = "agent is not on the space"; return false; }
        std::swap(*it, prev.back());                         // swap with last
        prev.pop_back();
        return true;
    }
    // add_agent_own_pos! (space_interaction_API.jl:174-179) with Vector-container
    // ids 1..N (model_accessing_API.jl:29-32)
    int64_t add_agent(const T* p, const T* v) {
        int64_t c[3];
        pos2cell(p, c);
        if (!cell_inbounds(c)) { err = "BoundsError: position maps outside the cell grid"; return -1; }
        int64_t id = ++n;
        for (int d = 0; d < D; ++d) { pos.push_back(p[d]); vel.push_back(v ? v[d] : T(0)); }
        add_to_space(id, c);
        return id;
    }
    // continuous.jl:157-169 move_agent!(agent, pos, model)
    int move_agent_to(int64_t id, const T* np) {
        for (int d = 0; d < D; ++d)
            if (!(T(0) <= np[d] && np[d] <= extent[d])) { err = "position is outside space extent!"; return -1; }
        T* p = &pos[(id - 1) * D];
        int64_t oc[3], nc[3];
        pos2cell(p, oc);
        pos2cell(np, nc);
        bool same = true;
        for (int d = 0; d < D; ++d) same = same && (oc[d] == nc[d]);
        if (!same) {
            if (!remove_from_space(id, oc)) return -2;
            add_to_space(id, nc);
        }
        for (int d = 0; d < D; ++d) p[d] = np[d];
        return 0;
    }
    // walk.jl:58-66
    int walk(int64_t id, const T* dir) {
        T t[3];
        const T* p = &pos[(id - 1) * D];
        for (int d = 0; d < D; ++d) t[d] = p[d] + dir[d];
        normalize_position(t);
        return move_agent_to(id, t);
    }
    // continuous.jl:186-194 move_agent!(agent, model, dt) (update_vel! = no-op)
    int move_agent_dt(int64_t id, double dt) {
        T dir[3];
        const T* v = &vel[(id - 1) * D];
        for (int d = 0; d < D; ++d) dir[d] = (T)dt * v[d];
        return walk(id, dir);
    }

    // nearby_ids(pos, model, r; search) — continuous.jl:212-239 on top of the
    // GridSpaceIdIterator walk, grid_multi.jl:121-232.  Calls f(id) in the
    // reference's iteration order.  exclude_id = 0 keeps everything
    // (space_interaction_API.jl:284-287 filters `i != id`).
    template <class F> void for_nearby(const T* p, double r, bool exact, int64_t exclude_id, F&& f) const {
        int64_t gr = grid_r(p, r);
        int64_t focal[3];
        pos2cell(p, focal);
        const std::vector<Off>& offs = cached_offsets(gr);
        int64_t c[3];
        for (const Off& b : offs) {
            if (!combine_position(D, dims, periodic, focal, b, c)) continue;
            const auto& ids = stored_ids[linear_index(D, dims, c)];
            for (int64_t id : ids) {
                if (id == exclude_id) continue;
                if (exact) {
                    // continuous.jl:237-238  euclidean_distance(pos, model[i].pos, model) ≤ r
                    T dist = euclid(p, &pos[(id - 1) * D]);
                    if (!((double)dist <= r)) continue;
                }
                f(id);
            }
        }
    }

    // ---- rules -----------------------------------------------------------
    // examples/flock.jl:79-105 — velocity part of agent_step! for bird `id`,
    // reading the current model state.  fp = the bird's own parameter fields.
    void flock_new_vel(int64_t id, const FlockParams& fp, bool exact, T* out_vel, int64_t* out_n) const {
        const T* bp = &pos[(id - 1) * D];
        const T* bv = &vel[(id - 1) * D];
        int64_t N = 0;
        T match[3] = {0, 0, 0}, separate[3] = {0, 0, 0}, cohere[3] = {0, 0, 0};
        T sep2 = (T)fp.separation * (T)fp.separation;   // bird.separation^2
        for_nearby(bp, fp.visual_distance, exact, id, [&](int64_t nid) {
            N += 1;
            T heading[3];
            get_direction(bp, &pos[(nid - 1) * D], heading);
            const T* nv = &vel[(nid - 1) * D];
            T s = 0;
            for (int d = 0; d < D; ++d) {
                cohere[d] += heading[d];
                match[d] += nv[d];
                T sq = heading[d] * heading[d];
                s = (d == 0) ? sq : s + sq;      // sum(heading .^ 2)
            }
            if (s < sep2)
                for (int d = 0; d < D; ++d) separate[d] -= heading[d];
        });
        T nv[3], s = 0;
        T denom = (T)std::max<int64_t>(N, 1);
        for (int d = 0; d < D; ++d) {
            T c = cohere[d] * (T)fp.cohere_factor;
            T sp = separate[d] * (T)fp.separate_factor;
            T m = match[d] * (T)fp.match_factor;
            nv[d] = bv[d] + ((c + sp) + m) / denom;       // flock.jl:104
            T sq = nv[d] * nv[d];
            s = (d == 0) ? sq : s + sq;
        }
        T nrm = std::sqrt(s);                               // LinearAlgebra.norm (StaticArrays)
        for (int d = 0; d < D; ++d) out_vel[d] = nv[d] / nrm;  // flock.jl:105
        if (out_n) *out_n = N;
    }
};

// ---------------------------------------------------------------------------
// Grid models (GridSpace: many ids per cell; GridSpaceSingle: one id or 0)
// src/spaces/grid_multi.jl, grid_single.jl, walk.jl:112-126
// ---------------------------------------------------------------------------
struct GModel : ModelBase {
    int D = 2;
    int single = 0;
    int metric = 0;
    uint8_t periodic[3] = {1, 1, 1};
    int64_t dims[3] = {1, 1, 1};
    int64_t ncells = 1;
    std::vector<int64_t> pos;                      // n*D, 1-based coords, id order
    std::vector<std::vector<int64_t>> stored_ids;  // GridSpace
    std::vector<int64_t> stored_single;            // GridSpaceSingle (0 = empty)
    int64_t n = 0;

    bool inb(const int64_t* c) const {
        for (int d = 0; d < D; ++d) if (c[d] < 1 || c[d] > dims[d]) return false;
        return true;
    }
    int64_t add_agent(const int64_t* p) {
        if (!inb(p)) { err = "BoundsError"; return -1; }
        int64_t lin = linear_index(D, dims, p);
        if (single) {
            // grid_single.jl:47-52
            if (stored_single[lin] != 0) { err = "Cannot add agent to occupied position"; return -2; }
        }
        int64_t id = ++n;
        for (int d = 0; d < D; ++d) pos.push_back(p[d]);
        if (single) stored_single[lin] = id; else stored_ids[lin].push_back(id);
        return id;
    }
    // generic move_agent! (space_interaction_API.jl:100-105): remove, set pos, add
    int move_agent_to(int64_t id, const int64_t* np) {
        if (!inb(np)) { err = "BoundsError"; return -1; }
        int64_t* p = &pos[(id - 1) * D];
        int64_t ol = linear_index(D, dims, p), nl = linear_index(D, dims, np);
        if (single) {
            stored_single[ol] = 0;                               // grid_single.jl:54-57
            if (stored_single[nl] != 0) { stored_single[ol] = id; err = "Cannot add agent to occupied position"; return -2; }
            stored_single[nl] = id;
        } else {
            auto& prev = stored_ids[ol];                          // grid_multi.jl:96-105
            auto it = std::find(prev.begin(), prev.end(), id);
            if (it == prev.end()) { err = "agent is not on the space"; return -3; }
            std::swap(*it, prev.back());
            prev.pop_back();
            stored_ids[nl].push_back(id);
        }
        for (int d = 0; d < D; ++d) p[d] = np[d];
        return 0;
    }
    // walk.jl:112-126 normalize_position for grids
    void normalize_position(int64_t* p) const {
        for (int d = 0; d < D; ++d)
            p[d] = periodic[d] ? jl_mod1(p[d], dims[d]) : std::min(std::max<int64_t>(p[d], 1), dims[d]);
    }
    bool is_empty(const int64_t* p) const {
        int64_t lin = linear_index(D, dims, p);
        return single ? stored_single[lin] == 0 : stored_ids[lin].empty();
    }
    // walk.jl:33-56 walk!(agent, direction, model; ifempty)
    int walk(int64_t id, const int64_t* dir, int ifempty) {
        int64_t t[3];
        const int64_t* p = &pos[(id - 1) * D];
        for (int d = 0; d < D; ++d) t[d] = p[d] + dir[d];
        normalize_position(t);
        if (single) ifempty = 1;                                  // walk.jl:46-56
        if (!ifempty || is_empty(t)) return move_agent_to(id, t);
        return 0;
    }
    // nearby_positions(pos, space, r) — grid_general.jl:185-280 (no_0 table)
    template <class F> void for_nearby_positions(const int64_t* origin, double r, F&& f) const {
        std::vector<Off> offs = offsets_within_radius(D, dims, periodic, metric, r, true);
        int64_t c[3];
        for (const Off& b : offs) {
            if (!combine_position(D, dims, periodic, origin, b, c)) continue;
            f(c);
        }
    }
    // nearby_ids(pos, model, r) — grid_multi.jl:120-232 / grid_single.jl:92-194.
    // no_zero = the GridSpaceSingle agent form (grid_single.jl:201-205).
    template <class F> void for_nearby_ids(const int64_t* origin, double r, bool no_zero, F&& f) const {
        std::vector<Off> offs = offsets_within_radius(D, dims, periodic, metric, r, no_zero);
        int64_t c[3];
        for (const Off& b : offs) {
            if (!combine_position(D, dims, periodic, origin, b, c)) continue;
            int64_t lin = linear_index(D, dims, c);
            if (single) { if (stored_single[lin] != 0) f(stored_single[lin]); }
            else for (int64_t id : stored_ids[lin]) f(id);
        }
    }
};

template <class T> CModel<T>* cm(void* h) { return static_cast<CModel<T>*>(h); }

struct Handle {
    int dtype;   // 0 = f64, 1 = f32
    ModelBase* m;
};

#define DISPATCH(h, expr_d, expr_f) (((Handle*)(h))->dtype == 0 ? (expr_d) : (expr_f))
#define M64(h) (static_cast<CModel<double>*>(((Handle*)(h))->m))
#define M32(h) (static_cast<CModel<float>*>(((Handle*)(h))->m))

template <class T> CModel<T>* make_cmodel(int D, const double* extent, double spacing, const uint8_t* periodic,
                                          std::string& err) {
    // continuous.jl:93-105
    auto* m = new CModel<T>();
    m->D = D;
    for (int d = 0; d < D; ++d) {
        if (!(extent[d] > 0)) { err = "ArgumentError: extent must be positive"; delete m; return nullptr; }
        double q = extent[d] / spacing;
        if (q != std::floor(q)) {
            err = "All dimensions in `extent` must be completely divisible by `spacing`";
            delete m; return nullptr;
        }
        m->dims[d] = (int64_t)std::floor(q);
        m->extent[d] = (T)extent[d];
        m->periodic[d] = periodic[d];
    }
    m->spacing = (T)spacing;
    int64_t nc = 1;
    for (int d = 0; d < D; ++d) nc *= m->dims[d];
    m->stored_ids.assign(nc, {});
    return m;
}

template <class T> int64_t nearby_collect(const CModel<T>* m, const T* p, double r, int exact, int64_t exclude,
                                          int64_t* out, int64_t cap) {
    int64_t k = 0;
    m->for_nearby(p, r, exact != 0, exclude, [&](int64_t id) { if (k < cap) out[k] = id; ++k; });
    return k;
}

struct PairHash {
    size_t operator()(const std::pair<int64_t, int64_t>& p) const {
        return std::hash<uint64_t>()((uint64_t)p.first * 0x9E3779B97F4A7C15ull ^ (uint64_t)p.second);
    }
};

// interacting_pairs(model, r, :all) — continuous.jl:418-452.  Discovery order
// is preserved; membership uses a hash set instead of the reference's linear
// `∉` scan (same result, not quadratic).
template <class T> int64_t pairs_all(const CModel<T>* m, double r, int exact, int64_t* out, int64_t cap) {
    std::unordered_set<std::pair<int64_t, int64_t>, PairHash> seen;
    int64_t k = 0;
    for (int64_t id = 1; id <= m->n; ++id) {
        m->for_nearby(&m->pos[(id - 1) * m->D], r, exact != 0, id, [&](int64_t nid) {
            std::pair<int64_t, int64_t> pr = id < nid ? std::make_pair(id, nid) : std::make_pair(nid, id);
            if (seen.insert(pr).second) {
                if (k < cap) { out[2 * k] = pr.first; out[2 * k + 1] = pr.second; }
                ++k;
            }
        });
    }
    return k;
}

// nearest_neighbor — continuous.jl:298-307
template <class T> int64_t nearest_neighbor(const CModel<T>* m, int64_t id, double r) {
    double d = std::numeric_limits<double>::infinity();
    int64_t j = 0;
    const T* p = &m->pos[(id - 1) * m->D];
    m->for_nearby(p, r, false, id, [&](int64_t nid) {
        double dn = (double)m->euclid(p, &m->pos[(nid - 1) * m->D]);
        if (dn < d) { d = dn; j = nid; }
    });
    return d > r ? 0 : j;
}

inline FlockParams params_for(const FlockParams* uniform, const double* per_agent, int64_t n, int64_t id) {
    if (!per_agent) return *uniform;
    FlockParams p;
    // per_agent: 6 columns of length n (SoA): speed, cohere, separation, separate, match, visual
    p.speed = per_agent[0 * n + id - 1];
    p.cohere_factor = per_agent[1 * n + id - 1];
    p.separation = per_agent[2 * n + id - 1];
    p.separate_factor = per_agent[3 * n + id - 1];
    p.match_factor = per_agent[4 * n + id - 1];
    p.visual_distance = per_agent[5 * n + id - 1];
    return p;
}

// Reference-faithful asynchronous step: step_standard.jl:9-23 with
// examples/flock.jl:79-108 applied in place in `order`.
template <class T> int flock_step_async(CModel<T>* m, const FlockParams* up, const double* pa, int exact,
                                        const int64_t* order, int64_t norder) {
    for (int64_t k = 0; k < norder; ++k) {
        int64_t id = order ? order[k] : k + 1;
        FlockParams fp = params_for(up, pa, m->n, id);
        T nv[3];
        m->flock_new_vel(id, fp, exact != 0, nv, nullptr);
        for (int d = 0; d < m->D; ++d) m->vel[(id - 1) * m->D + d] = nv[d];
        int rc = m->move_agent_dt(id, fp.speed);
        if (rc) return rc;
    }
    return 0;
}

static int g_threads = 1;     // orc_set_threads

// Synchronous formulation: all new velocities from the
// old state, then every bird moves.  Equivalent Julia: a model_step! that
// gathers new velocities into a temporary vector before a second loop applies
// move_agent!.  ncount (optional) receives the neighbour count N per bird.
template <class T> int flock_step_sync(CModel<T>* m, const FlockParams* up, const double* pa, int exact,
                                       int64_t* ncount) {
    std::vector<T> nvel(m->vel.size());
    // every bird reads the old state only: the gather parallelises over the host cores (orc_set_threads;
    // default 1).  Each bird's arithmetic is untouched, so the results do not depend on the thread count.
    auto gather = [&](int64_t lo, int64_t hi) {
        for (int64_t id = lo; id < hi; ++id) {
            FlockParams fp = params_for(up, pa, m->n, id);
            int64_t N;
            m->flock_new_vel(id, fp, exact != 0, &nvel[(id - 1) * m->D], &N);
            if (ncount) ncount[id - 1] = N;
        }
    };
    if (g_threads <= 1 || m->n < 4096) gather(1, m->n + 1);
    else {
        std::atomic<int64_t> next(1);
        auto work = [&]() {
            for (;;) {
                const int64_t lo = next.fetch_add(2048);
                if (lo > m->n) break;
                gather(lo, std::min<int64_t>(lo + 2048, m->n + 1));
            }
        };
        std::vector<std::thread> pool;
        for (int t = 1; t < g_threads; ++t) pool.emplace_back(work);
        work();
        for (auto& th : pool) th.join();
    }
    m->vel = nvel;
    for (int64_t id = 1; id <= m->n; ++id) {
        FlockParams fp = params_for(up, pa, m->n, id);
        int rc = m->move_agent_dt(id, fp.speed);
        if (rc) return rc;
    }
    return 0;
}

// Pair-force rule — examples/celllistmap.jl:143-155 (force), :176-191
// (integrator).  Pair enumeration = the reference's own
// interacting_pairs(:all, search=:exact) semantics with r = cutoff
// (CellListMap.jl itself is not in the tree).  To keep the
// summation order well defined each unique pair (i<j) is visited once, in
// all_pairs! discovery order.
template <class T> int pairforce_step(CModel<T>* m, const double* rad, const double* kk, const double* mass,
                                      double dt, double cutoff) {
    const int D = m->D;
    std::vector<T> f(m->pos.size(), T(0));   // forces zeroed every step
    std::unordered_set<std::pair<int64_t, int64_t>, PairHash> seen;
    for (int64_t id = 1; id <= m->n; ++id) {
        const T* x = &m->pos[(id - 1) * D];
        m->for_nearby(x, cutoff, true, id, [&](int64_t nid) {
            std::pair<int64_t, int64_t> pr = id < nid ? std::make_pair(id, nid) : std::make_pair(nid, id);
            if (!seen.insert(pr).second) return;
            int64_t i = pr.first, j = pr.second;
            T dr[3];
            m->get_direction(&m->pos[(i - 1) * D], &m->pos[(j - 1) * D], dr);   // y - x, minimum image
            T d2 = 0;
            for (int d = 0; d < D; ++d) { T sq = dr[d] * dr[d]; d2 = (d == 0) ? sq : d2 + sq; }
            T dist = std::sqrt(d2);
            T rs = (T)rad[i - 1] + (T)rad[j - 1];
            if (dist <= rs) {
                // fij = 2 * (k_i*k_j) * (d2 - rs^2) * (dr / d)
                T pref = (T(2) * ((T)kk[i - 1] * (T)kk[j - 1])) * (d2 - rs * rs);
                for (int d = 0; d < D; ++d) {
                    T fij = pref * (dr[d] / dist);
                    f[(i - 1) * D + d] += fij;
                    f[(j - 1) * D + d] -= fij;
                }
            }
        });
    }
    T tdt = (T)dt;
    for (int64_t id = 1; id <= m->n; ++id) {
        T x[3];
        for (int d = 0; d < D; ++d) {
            T a = f[(id - 1) * D + d] / (T)mass[id - 1];
            T v = m->vel[(id - 1) * D + d] + a * tdt;
            x[d] = (m->pos[(id - 1) * D + d] + v * tdt) + (a / T(2)) * (tdt * tdt);
            m->vel[(id - 1) * D + d] = v;
        }
        m->normalize_position(x);
        int rc = m->move_agent_to(id, x);
        if (rc) return rc;
    }
    return 0;
}

}  // namespace

// ===========================================================================
// C API (ctypes-facing)
// ===========================================================================
extern "C" {

double orc_mod_f64(double x, double y) { return jl_mod(x, y); }
float orc_mod_f32(float x, float y) { return jl_mod(x, y); }
int64_t orc_mod1(int64_t x, int64_t y) { return jl_mod1(x, y); }

// Offsets table; returns the count (writes at most cap rows of D int64).
int64_t orc_grid_offsets(int D, const int64_t* dims, const uint8_t* periodic, int metric, double r,
                         int no_zero, int64_t* out, int64_t cap) {
    int64_t sz[3] = {1, 1, 1};
    uint8_t per[3] = {0, 0, 0};
    for (int d = 0; d < D; ++d) { sz[d] = dims[d]; per[d] = periodic[d]; }
    auto o = offsets_within_radius(D, sz, per, metric, r, no_zero != 0);
    for (int64_t i = 0; i < (int64_t)o.size() && i < cap; ++i)
        for (int d = 0; d < D; ++d) out[i * D + d] = o[i][d];
    return (int64_t)o.size();
}

// ---- ContinuousSpace ------------------------------------------------------
void* orc_cs_create(int D, const double* extent, double spacing, const uint8_t* periodic, int dtype,
                    char* errbuf, int errlen) {
    std::string err;
    ModelBase* m = nullptr;
    if (D < 1 || D > 3) err = "D must be 1..3";
    else if (dtype == 0) m = make_cmodel<double>(D, extent, spacing, periodic, err);
    else if (dtype == 1) m = make_cmodel<float>(D, extent, spacing, periodic, err);
    else err = "bad dtype";
    if (!m) { if (errbuf && errlen > 0) snprintf(errbuf, errlen, "%s", err.c_str()); return nullptr; }
    return new Handle{dtype, m};
}
void orc_destroy(void* h) { if (h) { delete ((Handle*)h)->m; delete (Handle*)h; } }
const char* orc_last_error(void* h) { return ((Handle*)h)->m->err.c_str(); }

int orc_cs_dims(void* h, int64_t* dims) {
    int D = DISPATCH(h, M64(h)->D, M32(h)->D);
    for (int d = 0; d < D; ++d) dims[d] = DISPATCH(h, M64(h)->dims[d], M32(h)->dims[d]);
    return D;
}
int64_t orc_cs_nagents(void* h) { return DISPATCH(h, M64(h)->n, M32(h)->n); }

// pos/vel: n*D values of the model's dtype (row per agent).  Returns the last id or <0.
int64_t orc_cs_add_agents(void* h, const void* pos, const void* vel, int64_t n) {
    int64_t id = 0;
    if (((Handle*)h)->dtype == 0) {
        auto* m = M64(h);
        for (int64_t i = 0; i < n; ++i) {
            id = m->add_agent((const double*)pos + i * m->D, vel ? (const double*)vel + i * m->D : nullptr);
            if (id < 0) return id;
        }
    } else {
        auto* m = M32(h);
        for (int64_t i = 0; i < n; ++i) {
            id = m->add_agent((const float*)pos + i * m->D, vel ? (const float*)vel + i * m->D : nullptr);
            if (id < 0) return id;
        }
    }
    return id;
}
int orc_cs_get_state(void* h, void* pos, void* vel) {
    if (((Handle*)h)->dtype == 0) {
        auto* m = M64(h);
        if (pos) memcpy(pos, m->pos.data(), m->pos.size() * sizeof(double));
        if (vel) memcpy(vel, m->vel.data(), m->vel.size() * sizeof(double));
    } else {
        auto* m = M32(h);
        if (pos) memcpy(pos, m->pos.data(), m->pos.size() * sizeof(float));
        if (vel) memcpy(vel, m->vel.data(), m->vel.size() * sizeof(float));
    }
    return 0;
}
int orc_cs_set_vel(void* h, const void* vel) {
    if (((Handle*)h)->dtype == 0) { auto* m = M64(h); memcpy(m->vel.data(), vel, m->vel.size() * sizeof(double)); }
    else { auto* m = M32(h); memcpy(m->vel.data(), vel, m->vel.size() * sizeof(float)); }
    return 0;
}
// cells: n*D int32, 1-based (continuous.jl:112-113)
int orc_cs_pos2cell(void* h, const void* pos, int64_t n, int32_t* cells) {
    int64_t c[3];
    if (((Handle*)h)->dtype == 0) {
        auto* m = M64(h);
        for (int64_t i = 0; i < n; ++i) { m->pos2cell((const double*)pos + i * m->D, c); for (int d = 0; d < m->D; ++d) cells[i * m->D + d] = (int32_t)c[d]; }
    } else {
        auto* m = M32(h);
        for (int64_t i = 0; i < n; ++i) { m->pos2cell((const float*)pos + i * m->D, c); for (int d = 0; d < m->D; ++d) cells[i * m->D + d] = (int32_t)c[d]; }
    }
    return 0;
}
// ids currently stored in a cell (1-based cell coords), in stored order
int64_t orc_cs_cell_ids(void* h, const int64_t* cell, int64_t* out, int64_t cap) {
    const std::vector<int64_t>* v;
    if (((Handle*)h)->dtype == 0) { auto* m = M64(h); v = &m->stored_ids[linear_index(m->D, m->dims, cell)]; }
    else { auto* m = M32(h); v = &m->stored_ids[linear_index(m->D, m->dims, cell)]; }
    for (int64_t i = 0; i < (int64_t)v->size() && i < cap; ++i) out[i] = (*v)[i];
    return (int64_t)v->size();
}
int64_t orc_cs_grid_r(void* h, const void* pos, double r) {
    return DISPATCH(h, M64(h)->grid_r((const double*)pos, r), M32(h)->grid_r((const float*)pos, r));
}
// nearby_ids(pos, model, r; search) — includes any agent at pos (self too)
int64_t orc_cs_nearby_ids(void* h, const void* qpos, double r, int exact, int64_t* out, int64_t cap) {
    return DISPATCH(h, nearby_collect(M64(h), (const double*)qpos, r, exact, 0, out, cap),
                    nearby_collect(M32(h), (const float*)qpos, r, exact, 0, out, cap));
}
// nearby_ids(agent, model, r; search) — excludes the agent's own id
int64_t orc_cs_nearby_ids_agent(void* h, int64_t id, double r, int exact, int64_t* out, int64_t cap) {
    if (((Handle*)h)->dtype == 0) { auto* m = M64(h); return nearby_collect(m, &m->pos[(id - 1) * m->D], r, exact, id, out, cap); }
    auto* m = M32(h);
    return nearby_collect(m, &m->pos[(id - 1) * m->D], r, exact, id, out, cap);
}
// Batched agent-form counts for all agents (id order); used for parity at scale.
int orc_cs_nearby_counts_all(void* h, double r, int exact, int64_t* counts) {
    if (((Handle*)h)->dtype == 0) {
        auto* m = M64(h);
        for (int64_t id = 1; id <= m->n; ++id) { int64_t k = 0; m->for_nearby(&m->pos[(id - 1) * m->D], r, exact != 0, id, [&](int64_t) { ++k; }); counts[id - 1] = k; }
    } else {
        auto* m = M32(h);
        for (int64_t id = 1; id <= m->n; ++id) { int64_t k = 0; m->for_nearby(&m->pos[(id - 1) * m->D], r, exact != 0, id, [&](int64_t) { ++k; }); counts[id - 1] = k; }
    }
    return 0;
}
int64_t orc_cs_pairs_all(void* h, double r, int exact, int64_t* out, int64_t cap) {
    return DISPATCH(h, pairs_all(M64(h), r, exact, out, cap), pairs_all(M32(h), r, exact, out, cap));
}
int64_t orc_cs_nearest_neighbor(void* h, int64_t id, double r) {
    return DISPATCH(h, nearest_neighbor(M64(h), id, r), nearest_neighbor(M32(h), id, r));
}
int orc_cs_move_agent_dt(void* h, int64_t id, double dt) {
    return DISPATCH(h, M64(h)->move_agent_dt(id, dt), M32(h)->move_agent_dt(id, dt));
}
int orc_cs_move_agent_to(void* h, int64_t id, const void* pos) {
    return DISPATCH(h, M64(h)->move_agent_to(id, (const double*)pos), M32(h)->move_agent_to(id, (const float*)pos));
}
int orc_cs_walk(void* h, int64_t id, const void* dir) {
    return DISPATCH(h, M64(h)->walk(id, (const double*)dir), M32(h)->walk(id, (const float*)dir));
}
int orc_cs_normalize_position(void* h, void* pos) {
    if (((Handle*)h)->dtype == 0) M64(h)->normalize_position((double*)pos); else M32(h)->normalize_position((float*)pos);
    return 0;
}
double orc_cs_euclidean_distance(void* h, const void* a, const void* b) {
    return DISPATCH(h, (double)M64(h)->euclid((const double*)a, (const double*)b),
                    (double)M32(h)->euclid((const float*)a, (const float*)b));
}
int orc_cs_get_direction(void* h, const void* from, const void* to, void* out) {
    if (((Handle*)h)->dtype == 0) M64(h)->get_direction((const double*)from, (const double*)to, (double*)out);
    else M32(h)->get_direction((const float*)from, (const float*)to, (float*)out);
    return 0;
}

// params: 6 doubles (speed, cohere_factor, separation, separate_factor,
// match_factor, visual_distance); per_agent: NULL or 6 columns × n doubles.
int orc_flock_step_async(void* h, const double* params, const double* per_agent, int exact,
                         const int64_t* order, int64_t norder, int64_t nsteps) {
    FlockParams fp{params[0], params[1], params[2], params[3], params[4], params[5]};
    for (int64_t s = 0; s < nsteps; ++s) {
        int rc = DISPATCH(h, flock_step_async(M64(h), &fp, per_agent, exact, order, norder),
                          flock_step_async(M32(h), &fp, per_agent, exact, order, norder));
        if (rc) return rc;
    }
    return 0;
}
// threads of the synchronous step's gather phase (the asynchronous, reference-faithful step is sequential)
int orc_set_threads(int n) {
    if (n < 1) n = (int)std::thread::hardware_concurrency();
    g_threads = n < 1 ? 1 : n;
    return g_threads;
}

int orc_flock_step_sync(void* h, const double* params, const double* per_agent, int exact, int64_t nsteps,
                        int64_t* ncount) {
    FlockParams fp{params[0], params[1], params[2], params[3], params[4], params[5]};
    for (int64_t s = 0; s < nsteps; ++s) {
        int rc = DISPATCH(h, flock_step_sync(M64(h), &fp, per_agent, exact, ncount),
                          flock_step_sync(M32(h), &fp, per_agent, exact, ncount));
        if (rc) return rc;
    }
    return 0;
}
int orc_pairforce_step(void* h, const double* rad, const double* k, const double* mass, double dt,
                       double cutoff, int64_t nsteps) {
    for (int64_t s = 0; s < nsteps; ++s) {
        int rc = DISPATCH(h, pairforce_step(M64(h), rad, k, mass, dt, cutoff),
                          pairforce_step(M32(h), rad, k, mass, dt, cutoff));
        if (rc) return rc;
    }
    return 0;
}

// ---- Grid spaces ----------------------------------------------------------
void* orc_grid_create(int D, const int64_t* dims, const uint8_t* periodic, int metric, int single) {
    auto* g = new GModel();
    g->D = D; g->metric = metric; g->single = single;
    g->ncells = 1;
    for (int d = 0; d < D; ++d) { g->dims[d] = dims[d]; g->periodic[d] = periodic[d]; g->ncells *= dims[d]; }
    if (single) g->stored_single.assign(g->ncells, 0); else g->stored_ids.assign(g->ncells, {});
    return new Handle{2, g};
}
#define GM(h) (static_cast<GModel*>(((Handle*)(h))->m))
int64_t orc_grid_add_agents(void* h, const int64_t* pos, int64_t n) {
    int64_t id = 0;
    for (int64_t i = 0; i < n; ++i) { id = GM(h)->add_agent(pos + i * GM(h)->D); if (id < 0) return id; }
    return id;
}
int orc_grid_get_pos(void* h, int64_t* pos) { memcpy(pos, GM(h)->pos.data(), GM(h)->pos.size() * sizeof(int64_t)); return 0; }
int orc_grid_get_stored_single(void* h, int64_t* out) { memcpy(out, GM(h)->stored_single.data(), GM(h)->stored_single.size() * sizeof(int64_t)); return 0; }
int orc_grid_move_agent_to(void* h, int64_t id, const int64_t* pos) { return GM(h)->move_agent_to(id, pos); }
int orc_grid_walk(void* h, int64_t id, const int64_t* dir, int ifempty) { return GM(h)->walk(id, dir, ifempty); }
int64_t orc_grid_nearby_positions(void* h, const int64_t* origin, double r, int64_t* out, int64_t cap) {
    int64_t k = 0;
    int D = GM(h)->D;
    GM(h)->for_nearby_positions(origin, r, [&](const int64_t* c) { if (k < cap) for (int d = 0; d < D; ++d) out[k * D + d] = c[d]; ++k; });
    return k;
}
// by position (includes the id(s) at origin)
int64_t orc_grid_nearby_ids(void* h, const int64_t* origin, double r, int64_t* out, int64_t cap) {
    int64_t k = 0;
    GM(h)->for_nearby_ids(origin, r, false, [&](int64_t id) { if (k < cap) out[k] = id; ++k; });
    return k;
}
// agent form: GridSpace filters the id (space_interaction_API.jl:284-287);
// GridSpaceSingle uses the no_0 table (grid_single.jl:201-205)
int64_t orc_grid_nearby_ids_agent(void* h, int64_t id, double r, int64_t* out, int64_t cap) {
    GModel* g = GM(h);
    int64_t k = 0;
    const int64_t* p = &g->pos[(id - 1) * g->D];
    if (g->single) g->for_nearby_ids(p, r, true, [&](int64_t j) { if (k < cap) out[k] = j; ++k; });
    else g->for_nearby_ids(p, r, false, [&](int64_t j) { if (j == id) return; if (k < cap) out[k] = j; ++k; });
    return k;
}

// Schelling happiness pass — docs/src/tutorial.jl:351-372 on GridSpaceSingle,
// counts from the snapshot at the start of the step (synchronous form).
// group: per-agent int32; out_count / out_mood per agent (id order).
int orc_schelling_happiness(void* h, const int32_t* group, int min_to_be_happy, int32_t* out_count,
                            uint8_t* out_mood) {
    GModel* g = GM(h);
    if (!g->single) { g->err = "Schelling needs GridSpaceSingle"; return -1; }
    for (int64_t id = 1; id <= g->n; ++id) {
        int32_t cnt = 0;
        g->for_nearby_ids(&g->pos[(id - 1) * g->D], 1.0, true, [&](int64_t j) { if (group[j - 1] == group[id - 1]) ++cnt; });
        out_count[id - 1] = cnt;
        out_mood[id - 1] = cnt >= min_to_be_happy ? 1 : 0;
    }
    return 0;
}

// Reference-faithful asynchronous Schelling step (CPU baseline only): agents
// in id order, unhappy ones jump to a uniformly random empty cell by
// rejection sampling (discrete.jl:96-114, :278-287).  RNG: splitmix64 stream
// (the reference's MersenneTwister stream is not reproduced; SURVEY §8c iv).
static inline uint64_t splitmix64(uint64_t& s) {
    uint64_t z = (s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
int orc_schelling_step_async(void* h, const int32_t* group, int min_to_be_happy, uint8_t* mood, uint64_t seed,
                             int64_t nsteps) {
    GModel* g = GM(h);
    if (!g->single) return -1;
    uint64_t s = seed;
    for (int64_t st = 0; st < nsteps; ++st)
        for (int64_t id = 1; id <= g->n; ++id) {
            int32_t cnt = 0;
            g->for_nearby_ids(&g->pos[(id - 1) * g->D], 1.0, true, [&](int64_t j) { if (group[j - 1] == group[id - 1]) ++cnt; });
            if (cnt >= min_to_be_happy) { mood[id - 1] = 1; continue; }
            mood[id - 1] = 0;
            if (g->n >= g->ncells) continue;   // no empty position: no-op
            int64_t p[3];
            for (;;) {
                for (int d = 0; d < g->D; ++d) p[d] = 1 + (int64_t)(splitmix64(s) % (uint64_t)g->dims[d]);
                if (g->is_empty(p)) break;
            }
            g->move_agent_to(id, p);
        }
    return 0;
}

}  // extern "C"
