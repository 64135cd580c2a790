// ab_core.cu — lifecycle, agent columns, and the per-step cell-list build.
//
// The cell list replaces the reference's persistent Array{Vector{Int},D} index
// and its incremental upkeep (src/spaces/continuous.jl:133-153,
// src/spaces/grid_multi.jl:68-105).  Build = counting sort by column-major
// cell key:
//   k_count   key = pos2cell (continuous.jl:112-113); arrival rank by atomicAdd
//   k_scan    single-pass decoupled-look-back exclusive scan → cell_start
//   k_bucket  ids into their cell segment in arrival order
//   k_place   canonical slot = cell_start + #(ids in the cell smaller than mine);
//             moves pos/vel/id/key to the other buffer (and records perm)
// so a cell's agents are always stored in id order whatever the atomic order was.
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "ab_internal.cuh"

// ---------------------------------------------------------------------------
// error helpers / bookkeeping
// ---------------------------------------------------------------------------
static std::string g_create_error;

int ab_fail(ab_handle* h, int code, const std::string& msg) {
    if (h) h->err = msg; else g_create_error = msg;
    return code;
}
int ab_cuda_fail(ab_handle* h, cudaError_t e, const char* what) {
    std::string m = std::string("CUDA error: ") + cudaGetErrorString(e) + " in " + what;
    return ab_fail(h, AB_ERR_CUDA, m);
}
size_t ab_dtype_size(int dtype) {
    switch (dtype) {
        case AB_F64: case AB_I64: return 8;
        case AB_F32: case AB_I32: return 4;
        case AB_U8: return 1;
    }
    return 0;
}

// Per-launch timing (ab_set_profiling): events are only RECORDED around the launch — no host
// synchronisation inside the step, so the timed region keeps its shape — and resolved later by
// ab_kernel_times().
ab_launch_scope::ab_launch_scope(ab_handle* h_, const char* name_) : h(h_), name(name_) {
    h->launches++;
    if (h->profiling) {
        cudaEventCreate(&a);
        cudaEventCreate(&b);
        cudaEventRecord(a, h->stream);
    }
}
ab_launch_scope::~ab_launch_scope() {
    if (h->profiling && a) {
        cudaEventRecord(b, h->stream);
        h->pending.push_back({name, a, b});
    }
}
static void resolve_pending(ab_handle* h) {
    if (h->pending.empty()) return;
    cudaStreamSynchronize(h->stream);
    for (auto& pe : h->pending) {
        float ms = 0;
        if (cudaEventElapsedTime(&ms, pe.a, pe.b) == cudaSuccess) {
            bool found = false;
            for (auto& k : h->kstats)
                if (k.name == pe.name || strcmp(k.name, pe.name) == 0) { k.ms += ms; k.launches++; found = true; break; }
            if (!found) h->kstats.push_back({pe.name, (double)ms, 1});
        }
        cudaEventDestroy(pe.a);
        cudaEventDestroy(pe.b);
    }
    h->pending.clear();
}

static int grow(ab_handle* h, void** p, size_t* have, size_t bytes) {
    if (bytes <= *have) return AB_OK;
    if (*p) { AB_CUDA(h, cudaStreamSynchronize(h->stream)); AB_CUDA(h, cudaFree(*p)); *p = nullptr; *have = 0; }
    size_t want = bytes + bytes / 4 + 256;
    AB_CUDA(h, cudaMalloc(p, want));
    *have = want;
    return AB_OK;
}
int ab_scratch(ab_handle* h, size_t bytes, void** out) {
    AB_TRY(grow(h, &h->scratch, &h->scratch_bytes, bytes));
    *out = h->scratch;
    return AB_OK;
}
int ab_scratch2(ab_handle* h, size_t bytes, void** out) {
    AB_TRY(grow(h, &h->scratch2, &h->scratch2_bytes, bytes));
    *out = h->scratch2;
    return AB_OK;
}

int ab_check_device_errors_now(ab_handle* h) {
    const bool was_async = h->async_mode;
    h->async_mode = false;
    const int rc = ab_check_device_errors(h);
    h->async_mode = was_async;
    return rc;
}

int ab_step_failed(ab_handle* h, int rc) {
    h->rs_ntiles = 0;
    h->index_valid = false;
    h->agg_valid = false;
    if (h->cell_count) cudaMemsetAsync(h->cell_count, 0, (size_t)h->nfine_cap * sizeof(uint32_t), h->stream);   // best effort
    return rc;
}

__global__ void k_iota(int32_t* a, int32_t* b, int n);
int ab_step_state_lost(ab_handle* h, int rc) {
    ab_step_failed(h, rc);
    if (!h->win_W && h->n > 0 && h->desc.kind == AB_SPACE_CONTINUOUS && h->id[h->cur] && h->slot_of_id) {
        k_iota<<<ab_blocks(h->n, 256), 256, 0, h->stream>>>(h->id[h->cur], h->slot_of_id, (int)h->n);
        h->slot_map_valid = true;
    }
    return rc;
}

int ab_check_device_errors(ab_handle* h) {
    if (h->async_mode) return AB_OK;      // deferred to ab_sync
    AB_CUDA(h, cudaMemcpyAsync(h->h_err, h->d_err, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    AB_CUDA(h, cudaStreamSynchronize(h->stream));
    int e = *h->h_err;
    if (!e) return AB_OK;
    AB_CUDA(h, cudaMemsetAsync(h->d_err, 0, sizeof(int), h->stream));
    ab_step_failed(h, 0);                 // index invalid, re-sort histograms cleared by the next plan
    if (e & AB_EF_NUMERIC) return ab_fail(h, AB_ERR_NUMERIC, "InexactError: a NaN/Inf position reached a cell computation");
    if (e & AB_EF_OUTSIDE) return ab_fail(h, AB_ERR_OUTSIDE, "position is outside space extent!");
    if (e & AB_EF_OCCUPIED) return ab_fail(h, AB_ERR_OCCUPIED, "Cannot add agent to occupied position");
    if (e & AB_EF_ARG) return ab_fail(h, AB_ERR_INVALID, "ArgumentError: The displacement must be larger than 0 (a resting agent has no direction to take it from).");
    if (e & AB_EF_COMM) {
        std::string m = "halo exchange failed:";
        if (e & AB_EF_COMM_CAP) m += " more agents in the boundary layers than the message capacity (ab_comm_set_capacity);";
        if (e & AB_EF_COMM_WINDOW) m += " owned + halo agents exceed the handle's capacity (ab_resize);";
        if (e & AB_EF_COMM_HDR) m += " a neighbour's message has a foreign header (ranks disagree on refinement or capacity);";
        if (e & AB_EF_COMM_TIMEOUT) {
            m += " a neighbour's message did not arrive within the halo timeout (AB_HALO_TIMEOUT_S, default 120 s): the halo was not installed, call ab_comm_exchange on every rank before the next step;";
            h->halo_stale = true;
        }
        return ab_fail(h, AB_ERR_COMM, m);
    }
    if (e & AB_EF_OVERFLOW) return ab_fail(h, AB_ERR_INVALID, "device-side failure: id out of range, or an agent left its cell neighbourhood in one tiled step");
    return ab_fail(h, AB_ERR_INVALID, "device-side failure");
}

// ---------------------------------------------------------------------------
// kernels
// ---------------------------------------------------------------------------
template <class T, int D>
__global__ void k_count(SpaceDev<T, D> sp, const T* __restrict__ pos, int64_t cap, int n,
                        uint32_t* __restrict__ newkey, uint32_t* __restrict__ rank,
                        uint32_t* __restrict__ cell_count, int* __restrict__ err) {
    int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    T p[D];
    int c[D];
#pragma unroll
    for (int d = 0; d < D; ++d) p[d] = pos[(size_t)d * cap + s];
    bool ok = pos2cell0<T, D>(sp, p, c);
    if (!ok) {
        bool nan = false;
#pragma unroll
        for (int d = 0; d < D; ++d) nan = nan || !(p[d] == p[d]) || isinf(p[d]);
        atomicOr(err, nan ? AB_EF_NUMERIC : AB_EF_OUTSIDE);
    }
    bool inwin = true;
    uint32_t k = device_key<T, D>(sp, p, c, &inwin);
    if (!inwin) { atomicOr(err, AB_EF_OVERFLOW); k = 0; }     // agent uploaded to a rank whose slab window does not hold it
    newkey[s] = k;
    rank[s] = atomicAdd(&cell_count[k], 1u);
}

// Single-pass exclusive scan (decoupled look-back) of cell_count[0..n) into
// cell_start[0..n]; zeroes cell_count behind itself.  status word:
// [epoch:30 | flag:2 | value:32]; flag 1 = tile aggregate, 2 = inclusive prefix.
// The look-back window is the whole CTA (256 predecessors per round, not a
// warp's 32): all tiles of a small scan start together and publish their
// aggregates at about the same time, so a 32-wide window needs ntiles/32
// dependent L2 round trips (measured: 20 us for 313 tiles) where the
// CTA-wide window needs one or two.
__global__ void __launch_bounds__(SCAN_THREADS)
k_scan(uint32_t* __restrict__ count, uint32_t* __restrict__ start, uint32_t n, uint32_t ntiles,
       volatile unsigned long long* status, uint32_t* ticket, uint32_t epoch) {
    __shared__ uint32_t s_tile;
    __shared__ uint32_t s_warp[SCAN_THREADS / 32];
    __shared__ uint32_t s_red[SCAN_THREADS / 32];
    __shared__ int s_first[SCAN_THREADS / 32];
    __shared__ uint32_t s_excl;
    ab_dep_wait();
    if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1u);
    __syncthreads();
    const uint32_t tile = s_tile;
    const uint32_t base = tile * SCAN_TILE + threadIdx.x * SCAN_ITEMS;
    uint32_t v[SCAN_ITEMS];
#pragma unroll
    for (int q = 0; q < SCAN_ITEMS / 4; ++q) {
        const uint4 a = *reinterpret_cast<const uint4*>(count + base + 4 * q);
        v[4 * q] = a.x; v[4 * q + 1] = a.y; v[4 * q + 2] = a.z; v[4 * q + 3] = a.w;
    }
#pragma unroll
    for (int q = 0; q < SCAN_ITEMS / 4; ++q) *reinterpret_cast<uint4*>(count + base + 4 * q) = make_uint4(0, 0, 0, 0);
    uint32_t tsum = 0;
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; ++i) { uint32_t t = v[i]; v[i] = tsum; tsum += t; }
    // block exclusive scan of tsum
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    uint32_t inc = tsum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { uint32_t t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += t; }
    if (lane == 31) s_warp[wid] = inc;
    __syncthreads();
    uint32_t woff = 0, aggregate = 0;
#pragma unroll
    for (int w = 0; w < SCAN_THREADS / 32; ++w) { const uint32_t t = s_warp[w]; if (w < wid) woff += t; aggregate += t; }
    const unsigned long long ep = (unsigned long long)epoch << 34;
    if (threadIdx.x == 0) status[tile] = ep | ((tile == 0 ? 2ull : 1ull) << 32) | aggregate;
    // CTA-wide look-back: thread i inspects tile (p - i)
    uint32_t excl = 0;
    int p = (int)tile - 1;
    while (p >= 0) {                                   // p is uniform across the CTA
        const int idx = p - (int)threadIdx.x;
        unsigned long long st = ep | (2ull << 32);      // virtual tile before the first: inclusive 0
        if (idx >= 0) {
            do { st = status[idx]; } while (((st >> 34) != (unsigned long long)epoch) || (((st >> 32) & 3ull) == 0));
        }
        const bool incl = ((st >> 32) & 3ull) == 2ull;
        // nearest inclusive predecessor in this window
        const unsigned m = __ballot_sync(0xffffffffu, incl);
        if (lane == 0) s_first[wid] = m ? (wid * 32 + __ffs(m) - 1) : 0x7fffffff;
        __syncthreads();
        int first = 0x7fffffff;
#pragma unroll
        for (int w = 0; w < SCAN_THREADS / 32; ++w) first = min(first, s_first[w]);
        uint32_t val = ((int)threadIdx.x <= first) ? (uint32_t)st : 0u;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) val += __shfl_xor_sync(0xffffffffu, val, o);
        if (lane == 0) s_red[wid] = val;
        __syncthreads();
#pragma unroll
        for (int w = 0; w < SCAN_THREADS / 32; ++w) excl += s_red[w];
        __syncthreads();                               // s_first / s_red are reused by the next round
        if (first != 0x7fffffff) break;
        p -= SCAN_THREADS;
    }
    if (threadIdx.x == 0) {
        if (tile != 0) status[tile] = ep | (2ull << 32) | (unsigned long long)(excl + aggregate);
        if (tile == ntiles - 1) { start[n] = excl + aggregate; *ticket = 0; }
    }
    const uint32_t off = excl + woff + (inc - tsum);
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; ++i)
        if (base + i < n) start[base + i] = off + v[i];
}

// Small scans (the tile totals of the tiled re-sort: ~10 k entries) in ONE CTA: no ticket, no status words, no
// look-back — thread t owns the items [t * ipt, (t + 1) * ipt), ipt <= 32.  Zeroes count behind itself like k_scan.
#define SCAN_SMALL_MAX 32768u
__global__ void __launch_bounds__(1024)
k_scan_small(uint32_t* __restrict__ count, uint32_t* __restrict__ start, uint32_t n) {
    __shared__ uint32_t s_w[32];
    ab_dep_wait();
    const uint32_t ipt = (n + 1023u) >> 10;
    const uint32_t b = threadIdx.x * ipt;
    uint32_t v[32];
    uint32_t sum = 0;
#pragma unroll
    for (uint32_t q = 0; q < 32; ++q) {
        v[q] = 0;
        if (q < ipt && b + q < n) { v[q] = count[b + q]; count[b + q] = 0; }
        sum += v[q];
    }
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    uint32_t inc = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += t; }
    if (lane == 31) s_w[wid] = inc;
    __syncthreads();
    if (wid == 0) {
        uint32_t w = s_w[lane], wi = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, wi, o); if (lane >= o) wi += t; }
        s_w[lane] = wi - w;                                   // exclusive prefix of the warp totals
        if (lane == 31) start[n] = wi;
    }
    __syncthreads();
    uint32_t run = s_w[wid] + (inc - sum);
#pragma unroll
    for (uint32_t q = 0; q < 32; ++q) {
        if (q < ipt && b + q < n) { start[b + q] = run; run += v[q]; }
    }
}

__global__ void k_bucket(const uint32_t* __restrict__ newkey, const uint32_t* __restrict__ rank,
                         const uint32_t* __restrict__ cell_start, const int32_t* __restrict__ id_old, int n,
                         int32_t* __restrict__ idtmp) {
    int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    idtmp[cell_start[newkey[s]] + rank[s]] = id_old[s];
}

#define AB_CANON_MAX 4096   // cells fuller than this keep arrival order (still correct, not canonical)
template <class T, int D, bool HAS_VEL>
__global__ void k_place(const uint32_t* __restrict__ newkey, const uint32_t* __restrict__ rank,
                        const uint32_t* __restrict__ cell_start, const int32_t* __restrict__ id_old,
                        const int32_t* __restrict__ idtmp, const T* __restrict__ src_pos,
                        const T* __restrict__ src_vel, int64_t cap, int n, T* __restrict__ dst_pos,
                        T* __restrict__ dst_vel, int32_t* __restrict__ id_new, uint32_t* __restrict__ key_new,
                        int32_t* __restrict__ perm, int32_t* __restrict__ slot_of_id, const int32_t* __restrict__ prio,
                        int* __restrict__ noncanon) {
    int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    const uint32_t k = newkey[s];
    const uint32_t b = cell_start[k], e = cell_start[k + 1];
    const int32_t me = id_old[s];
    uint32_t r;
    if (prio) {                       // restored layout (ab_layout_set): in-cell order of the checkpoint
        r = 0;
        const int32_t mine = prio[me];
        for (uint32_t t = b; t < e; ++t) r += (prio[idtmp[t]] < mine) ? 1u : 0u;
    } else if (e - b <= AB_CANON_MAX) {
        r = 0;
        for (uint32_t t = b; t < e; ++t) r += (idtmp[t] < me) ? 1u : 0u;
    } else {
        r = rank[s];                  // arrival order: correct, but not a function of the state alone
        if (s == 0 || newkey[s - 1] != k) atomicOr(noncanon, 1);      // flagged once per crowded cell or so (ab_index_canonical)
    }
    const uint32_t dst = b + r;
#pragma unroll
    for (int d = 0; d < D; ++d) dst_pos[(size_t)d * cap + dst] = src_pos[(size_t)d * cap + s];
    if (HAS_VEL) {
#pragma unroll
        for (int d = 0; d < D; ++d) dst_vel[(size_t)d * cap + dst] = src_vel[(size_t)d * cap + s];
    }
    id_new[dst] = me;
    key_new[dst] = k;
    if (perm) perm[dst] = s;
    if (slot_of_id) slot_of_id[me] = (int32_t)dst;
}

template <class E>
__global__ void k_permute(const E* __restrict__ src, E* __restrict__ dst, const int32_t* __restrict__ perm,
                          int64_t cap, int n, int ncomp, const uint32_t* __restrict__ nlive) {
    int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= (nlive ? (int)*nlive : n)) return;        // slab window: the owned count lives on the device
    int o = perm[s];
    for (int c = 0; c < ncomp; ++c) dst[(size_t)c * cap + s] = src[(size_t)c * cap + o];
}

__global__ void k_iota(int32_t* a, int32_t* b, int n) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { a[i] = i; if (b) b[i] = i; }
}
__global__ void k_slot_map(const int32_t* __restrict__ id, int32_t* __restrict__ slot_of_id, int n) {
    int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s < n) slot_of_id[id[s]] = s;
}

// host row-per-agent (id order) -> device SoA at the agent's slot, and back
template <class E>
__global__ void k_upload(const E* __restrict__ host_rows, E* __restrict__ dst, const int32_t* __restrict__ slot_of_id,
                         int64_t cap, int n, int ncomp) {
    int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= (int64_t)n * ncomp) return;
    int i = (int)(t / ncomp), c = (int)(t - (int64_t)i * ncomp);
    dst[(size_t)c * cap + slot_of_id[i]] = host_rows[t];
}
template <class E>
__global__ void k_download(const E* __restrict__ src, E* __restrict__ host_rows, const int32_t* __restrict__ id,
                           int64_t cap, int n, int ncomp) {
    int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    int i = id[s];
    for (int c = 0; c < ncomp; ++c) host_rows[(size_t)i * ncomp + c] = src[(size_t)c * cap + s];
}
template <class T, int D>
__global__ void k_cells(SpaceDev<T, D> sp, const T* __restrict__ pos, const int32_t* __restrict__ id, int64_t cap,
                        int n, int32_t* __restrict__ out) {
    int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    T p[D];
    int c[D];
#pragma unroll
    for (int d = 0; d < D; ++d) p[d] = pos[(size_t)d * cap + s];
    pos2cell0<T, D>(sp, p, c);
    int i = id[s];
#pragma unroll
    for (int d = 0; d < D; ++d) out[(size_t)i * D + d] = c[d] + 1;
}

// ---------------------------------------------------------------------------
// index build
// ---------------------------------------------------------------------------
// Exclusive scan of count[0..n) into start[0..n] (start[n] = total); zeroes count.  `count` must be
// readable / writable up to the next multiple of SCAN_TILE.
int ab_scan_u32(ab_handle* h, uint32_t* count, uint32_t* start, uint32_t n) {
    if (n > 0 && n <= SCAN_SMALL_MAX && h->small_scan) {      // measured slower than three look-back tiles (18.7 vs 10.5 us): off
        AB_LAUNCH(h, "k_scan");
        AB_CUDA(h, ab_launch_chain(h->pdl, k_scan_small, dim3(1), dim3(1024), h->stream, count, start, n));
        return AB_OK;
    }
    uint32_t ntiles = (n + SCAN_TILE - 1) / SCAN_TILE;
    if (ntiles == 0) ntiles = 1;
    if (ntiles > h->scan_status_len) return ab_fail(h, AB_ERR_INVALID, "scan larger than the handle's status buffer");
    h->scan_epoch = (h->scan_epoch + 1) & 0x3fffffffu;
    if (h->scan_epoch == 0) {   // epoch wrapped: clear stale status words once
        AB_CUDA(h, cudaMemsetAsync(h->scan_status, 0, h->scan_status_len * sizeof(unsigned long long), h->stream));
        h->scan_epoch = 1;
    }
    AB_LAUNCH(h, "k_scan");
    AB_CUDA(h, ab_launch_chain(h->pdl, k_scan, dim3(ntiles), dim3(SCAN_THREADS), h->stream, count, start, n, ntiles, h->scan_status, h->scan_ticket, h->scan_epoch));
    return AB_OK;
}
int ab_run_scan(ab_handle* h, uint32_t* dst) { return ab_scan_u32(h, h->cell_count, dst, (uint32_t)h->nfine); }

int ab_permute_extra_fields(ab_handle* h, bool owned_only) {
    const int cur = h->cur, nxt = cur ^ 1, TB = 256;
    // slab window: only the owned slots were re-sorted (the halo columns arrive with the neighbours' messages); their
    // number is the new cell_start entry behind the owned keys
    const bool live = owned_only && h->win_W;
    const int n = live ? (int)h->cap : (int)h->n;
    const uint32_t* nlive = live ? h->cell_start + (h->nfine / h->dims[h->D - 1]) * h->win_nl : nullptr;
    const size_t first = h->desc.kind == AB_SPACE_CONTINUOUS ? 2 : 1;
    for (size_t f = first; f < h->fields.size(); ++f) {
        ab_field& fl = h->fields[f];
        AB_LAUNCH(h, "k_permute");
        if (fl.esize == 8)
            k_permute<uint64_t><<<ab_blocks(n, TB), TB, 0, h->stream>>>((const uint64_t*)fl.buf[cur], (uint64_t*)fl.buf[nxt], h->perm, h->cap, n, fl.ncomp, nlive);
        else if (fl.esize == 4)
            k_permute<uint32_t><<<ab_blocks(n, TB), TB, 0, h->stream>>>((const uint32_t*)fl.buf[cur], (uint32_t*)fl.buf[nxt], h->perm, h->cap, n, fl.ncomp, nlive);
        else
            k_permute<uint8_t><<<ab_blocks(n, TB), TB, 0, h->stream>>>((const uint8_t*)fl.buf[cur], (uint8_t*)fl.buf[nxt], h->perm, h->cap, n, fl.ncomp, nlive);
    }
    return AB_OK;
}

template <class T, int D>
static int reindex_t(ab_handle* h, const void* src_pos, const void* src_vel, bool count_done) {
    const int n = (int)h->n;
    const int cur = h->cur, nxt = cur ^ 1;
    const int TB = 256;
    SpaceDev<T, D> sp = make_space<T, D>(h);
    const bool has_vel = h->desc.kind == AB_SPACE_CONTINUOUS;
    if (n > 0 && !count_done) {
        AB_LAUNCH(h, "k_count");
        k_count<T, D><<<ab_blocks(n, TB), TB, 0, h->stream>>>(sp, (const T*)src_pos, h->cap, n, h->newkey, h->rank,
                                                              h->cell_count, h->d_err);
    }
    AB_TRY(ab_run_scan(h, h->cell_start));
    if (n > 0) {
        {
            AB_LAUNCH(h, "k_bucket");
            k_bucket<<<ab_blocks(n, TB), TB, 0, h->stream>>>(h->newkey, h->rank, h->cell_start, h->id[cur], n, h->idtmp);
        }
        const bool extra = h->fields.size() > (has_vel ? 2u : 1u);
        {
            AB_LAUNCH(h, "k_place");
            if (has_vel)
                k_place<T, D, true><<<ab_blocks(n, TB), TB, 0, h->stream>>>(
                    h->newkey, h->rank, h->cell_start, h->id[cur], h->idtmp, (const T*)src_pos, (const T*)src_vel,
                    h->cap, n, (T*)h->fields[0].buf[nxt], (T*)h->fields[1].buf[nxt], h->id[nxt], h->key[nxt],
                    extra ? h->perm : nullptr, nullptr, h->place_prio, h->d_err + 2);
            else
                k_place<T, D, false><<<ab_blocks(n, TB), TB, 0, h->stream>>>(
                    h->newkey, h->rank, h->cell_start, h->id[cur], h->idtmp, (const T*)src_pos, nullptr, h->cap, n,
                    (T*)h->fields[0].buf[nxt], nullptr, h->id[nxt], h->key[nxt], extra ? h->perm : nullptr, nullptr, h->place_prio, h->d_err + 2);
        }
        AB_TRY(ab_permute_extra_fields(h, false));
    }
    h->cur = nxt;
    h->slot_map_valid = false;
    h->index_valid = true;
    h->agg_valid = false;
    h->pairs_n = -1;
    return AB_OK;
}

// Moves the population into the other buffer in the sorted order given by
// src_pos (src_* are indexed by the CURRENT slot order).  count_done: a fused
// step kernel already produced newkey/rank/cell_count.
int ab_reindex_from(ab_handle* h, const void* src_pos, const void* src_vel, bool count_done) {
    if (h->dtype == AB_F64) {
        return h->D == 2 ? reindex_t<double, 2>(h, src_pos, src_vel, count_done)
                         : reindex_t<double, 3>(h, src_pos, src_vel, count_done);
    }
    return h->D == 2 ? reindex_t<float, 2>(h, src_pos, src_vel, count_done)
                     : reindex_t<float, 3>(h, src_pos, src_vel, count_done);
}

int ab_ensure_index(ab_handle* h) {
    if (h->index_valid) return AB_OK;
    if (h->desc.kind != AB_SPACE_CONTINUOUS) return ab_fail(h, AB_ERR_INVALID, "cell-list index is for ContinuousSpace handles");
    AB_TRY(ab_reindex_from(h, h->fields[0].buf[h->cur], h->fields[1].buf[h->cur], false));
    return ab_check_device_errors(h);
}

// Device-grid refinement (1 or 2 sub-cells per reference cell and dimension).
// Pure layout choice: every query walks reference cells, so results do not
// depend on it; the fused sweeps use the sub-cells to cut candidate counts.
int ab_set_refine(ab_handle* h, int K) {
    if (K != 1 && K != 2 && !(K == 4 && h->D == 2)) return ab_fail(h, AB_ERR_INVALID, "refinement must be 1, 2 or (2-D only) 4");
    if (h->desc.kind != AB_SPACE_CONTINUOUS && K != 1) return ab_fail(h, AB_ERR_INVALID, "grid spaces are not refined");
    const int kb = K == 4 ? 2 : (K == 2 ? 1 : 0);
    const int ksh = kb * h->D;
    if ((h->ncells << ksh) > h->nfine_cap) return ab_fail(h, AB_ERR_INVALID, "grid too large to refine");
    if (K == h->K) return AB_OK;
    h->K = K;
    h->kb = kb;
    h->ksh = ksh;
    h->nfine = h->ncells << ksh;
    h->index_valid = false;
    h->agg_valid = false;
    h->pairs_n = -1;
    return AB_OK;
}

int ab_ensure_slot_map(ab_handle* h) {
    if (h->win_W) return ab_fail(h, AB_ERR_STATE, "id-addressed calls are not available on a sharded (slab-window) handle; use ab_upload_local / ab_download_local");
    if (h->slot_map_valid) return AB_OK;
    if (h->n > 0) {
        AB_LAUNCH(h, "k_slot_map");
        k_slot_map<<<ab_blocks(h->n, 256), 256, 0, h->stream>>>(h->id[h->cur], h->slot_of_id, (int)h->n);
    }
    h->slot_map_valid = true;
    return AB_OK;
}

// ---------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------
extern "C" {

int ab_version(void) { return AB_VERSION; }

int ab_set_refinement(ab_handle* h, int K) {
    AB_CUDA(h, cudaSetDevice(h->device));
    AB_CUDA(h, cudaStreamSynchronize(h->stream));
    return ab_set_refine(h, K);
}

const char* ab_last_error(const ab_handle* h) { return h ? h->err.c_str() : g_create_error.c_str(); }

static void free_population(ab_handle* h) {
    for (auto& f : h->fields)
        for (int b = 0; b < 2; ++b) { if (f.buf[b]) cudaFree(f.buf[b]); f.buf[b] = nullptr; }
    void** ptrs[] = {&h->tmp_pos, &h->tmp_vel, (void**)&h->id[0], (void**)&h->id[1], (void**)&h->key[0],
                     (void**)&h->key[1], (void**)&h->newkey, (void**)&h->rank, (void**)&h->idtmp, (void**)&h->perm,
                     (void**)&h->slot_of_id, (void**)&h->ncount};
    for (void** p : ptrs) { if (*p) cudaFree(*p); *p = nullptr; }
    h->cap = 0;
    h->n = 0;
}

int ab_create(ab_space_desc* desc, int device, ab_handle** out) {
    if (!desc || !out) return ab_fail(nullptr, AB_ERR_INVALID, "null argument");
    *out = nullptr;
    int ndev = 0;
    cudaError_t ce = cudaGetDeviceCount(&ndev);
    if (ce != cudaSuccess || ndev == 0)
        return ab_fail(nullptr, AB_ERR_CUDA, std::string("no CUDA device available (there is no CPU fallback): ") +
                                                 (ce != cudaSuccess ? cudaGetErrorString(ce) : "device count is 0"));
    if (device < 0 || device >= ndev) return ab_fail(nullptr, AB_ERR_INVALID, "bad device ordinal");
    if (desc->D < 2 || desc->D > 3) return ab_fail(nullptr, AB_ERR_INVALID, "D must be 2 or 3 on the device");
    ab_handle* h = new ab_handle();
    h->device = device;
    if (const char* e = getenv("AB_OVERLAP")) h->overlap_halo = atoi(e) != 0;
    if (const char* e = getenv("AB_PDL")) h->pdl = atoi(e) != 0;
    if (const char* e = getenv("AB_FINE4")) h->fine4 = atoi(e) != 0;     // A/B experiments
    if (const char* e = getenv("AB_SMALL_SCAN")) h->small_scan = atoi(e) != 0;
    if (const char* e = getenv("AB_CELLM_OLD")) h->cellm_old = atoi(e) != 0;
    if (const char* e = getenv("AB_SIDE_UNPACK")) h->no_side_unpack = atoi(e) == 0;
    if (const char* e = getenv("AB_SPLIT_M")) h->no_split_M = atoi(e) == 0;
    if (const char* e = getenv("AB_RS_TARGET")) { int v = atoi(e); if (v >= 4 && v <= 2048) h->rs_target = v; }   // tuning experiments
    h->D = desc->D;
    int64_t ncells = 1;
    if (desc->kind == AB_SPACE_CONTINUOUS) {
        if (desc->dtype != AB_F64 && desc->dtype != AB_F32) { delete h; return ab_fail(nullptr, AB_ERR_INVALID, "dtype must be AB_F64 or AB_F32"); }
        if (!(desc->spacing > 0)) { delete h; return ab_fail(nullptr, AB_ERR_INVALID, "spacing must be positive"); }
        for (int d = 0; d < desc->D; ++d) {
            if (!(desc->extent[d] > 0)) { delete h; return ab_fail(nullptr, AB_ERR_INVALID, "ArgumentError: extent must be positive"); }
            double q = desc->extent[d] / desc->spacing;
            if (q != floor(q)) {   // src/spaces/continuous.jl:99-101
                delete h;
                return ab_fail(nullptr, AB_ERR_SPACING, "All dimensions in `extent` must be completely divisible by `spacing`");
            }
            desc->grid_dims[d] = (int64_t)floor(q);
        }
        h->dtype = desc->dtype;
    } else if (desc->kind == AB_SPACE_GRID_MULTI || desc->kind == AB_SPACE_GRID_SINGLE) {
        h->dtype = AB_I32;
        if (desc->metric < 0 || desc->metric > 2) { delete h; return ab_fail(nullptr, AB_ERR_INVALID, "Unknown metric type"); }
    } else { delete h; return ab_fail(nullptr, AB_ERR_INVALID, "unknown space kind"); }
    for (int d = 0; d < desc->D; ++d) {
        if (desc->grid_dims[d] < 1 || desc->grid_dims[d] > 2000000000LL) { delete h; return ab_fail(nullptr, AB_ERR_INVALID, "bad grid size"); }
        h->dims[d] = (int)desc->grid_dims[d];
        ncells *= desc->grid_dims[d];
    }
    if (ncells > 2000000000LL) { delete h; return ab_fail(nullptr, AB_ERR_INVALID, "more than 2e9 cells"); }
    h->ncells = ncells;
    h->nfine = ncells;
    // continuous spaces may refine every reference cell into 2^D sub-cells (ab_set_refine)
    // finest refinement a fused step may choose: K = 4 in 2-D (while the fine grid stays below 2^26 keys), K = 2 otherwise
    int max_ksh = (desc->kind == AB_SPACE_CONTINUOUS && (ncells << desc->D) < 4000000000LL) ? desc->D : 0;
    if (max_ksh && desc->D == 2 && (ncells << 4) <= (1LL << 26)) max_ksh = 4;
    h->tsize = ab_dtype_size(h->dtype);
    h->desc = *desc;

    cudaError_t e;
#define CK(call) if ((e = (call)) != cudaSuccess) { std::string m = std::string("CUDA error: ") + cudaGetErrorString(e) + " in " #call; ab_destroy(h); return ab_fail(nullptr, AB_ERR_CUDA, m); }
    CK(cudaSetDevice(device));
    CK(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
    CK(cudaEventCreate(&h->ev0));
    CK(cudaEventCreate(&h->ev1));
    size_t padded = (((size_t)ncells << max_ksh) + SCAN_TILE - 1) / SCAN_TILE * SCAN_TILE;
    h->nfine_cap = (int64_t)padded;
    CK(cudaMalloc((void**)&h->cell_count, padded * sizeof(uint32_t)));
    CK(cudaMemset(h->cell_count, 0, padded * sizeof(uint32_t)));
    CK(cudaMalloc((void**)&h->cell_start, (padded + 1) * sizeof(uint32_t)));
    CK(cudaMemset(h->cell_start, 0, (padded + 1) * sizeof(uint32_t)));
    if (desc->kind == AB_SPACE_CONTINUOUS) {
        CK(cudaMalloc((void**)&h->cell_start_alt, (padded + 1) * sizeof(uint32_t)));
        CK(cudaMemset(h->cell_start_alt, 0, (padded + 1) * sizeof(uint32_t)));
    }
    h->scan_status_len = padded / SCAN_TILE + 1;
    CK(cudaMalloc((void**)&h->scan_status, h->scan_status_len * sizeof(unsigned long long)));
    CK(cudaMemset(h->scan_status, 0, h->scan_status_len * sizeof(unsigned long long)));
    CK(cudaMalloc((void**)&h->scan_ticket, sizeof(uint32_t)));
    CK(cudaMemset(h->scan_ticket, 0, sizeof(uint32_t)));
    CK(cudaMalloc((void**)&h->d_err, 4 * sizeof(int)));      // [0] error flags, [1] "moved further than one cell" flag
    CK(cudaMemset(h->d_err, 0, 4 * sizeof(int)));
    CK(cudaMallocHost((void**)&h->h_err, 4 * sizeof(int)));
    if (desc->kind == AB_SPACE_GRID_SINGLE) {
        CK(cudaMalloc((void**)&h->grid_ids, (size_t)ncells * sizeof(int32_t)));
        CK(cudaMemset(h->grid_ids, 0, (size_t)ncells * sizeof(int32_t)));
    }
#undef CK
    ab_field pos;
    pos.name = "pos"; pos.dtype = h->dtype; pos.ncomp = h->D; pos.esize = h->tsize;
    h->fields.push_back(pos);
    if (desc->kind == AB_SPACE_CONTINUOUS) {
        ab_field vel = pos;
        vel.name = "vel";
        h->fields.push_back(vel);
    }
    *out = h;
    return AB_OK;
}

int ab_destroy(ab_handle* h) {
    if (!h) return AB_OK;
    cudaSetDevice(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    resolve_pending(h);
    ab_comm_free(h);
    ab_grid_free(h);
    free_population(h);
    void* ptrs[] = {h->cell_count, h->cell_start, h->scan_status, h->scan_ticket, h->d_err, h->scratch, h->scratch2,
                    h->cell_agg, h->cell_agg1, h->cell_M, h->pairs_offsets, h->grid_ids, h->cell_start_alt, h->step_mask,
                    h->tile_count, h->tile_base, h->row_count};
    for (void* p : ptrs) if (p) cudaFree(p);
    if (h->h_err) cudaFreeHost(h->h_err);
    if (h->ev0) cudaEventDestroy(h->ev0);
    if (h->ev1) cudaEventDestroy(h->ev1);
    if (h->ev_fork) cudaEventDestroy(h->ev_fork);
    if (h->ev_join) cudaEventDestroy(h->ev_join);
    if (h->stream2) cudaStreamDestroy(h->stream2);
    if (h->stream) cudaStreamDestroy(h->stream);
    delete h;
    return AB_OK;
}

int ab_sync(ab_handle* h) {
    AB_CUDA(h, cudaSetDevice(h->device));
    AB_CUDA(h, cudaStreamSynchronize(h->stream));
    const bool stepped = h->step_pending;
    if (h->step_pending) {
        float ms = 0;
        if (cudaEventElapsedTime(&ms, h->ev0, h->ev1) == cudaSuccess) h->last_step_ms = ms;
        h->step_pending = false;
    }
    if (h->async_mode) {                  // device-side failures of everything queued since the last ab_sync
        h->async_mode = false;
        const int rc = ab_check_device_errors(h);
        h->async_mode = true;
        return rc != AB_OK && stepped ? ab_step_state_lost(h, rc) : rc;
    }
    return AB_OK;
}
// Asynchronous mode: ab_upload / ab_download / ab_step_flock / ab_build_index return as soon as their
// copies and kernels are queued on the handle's stream (host buffers must be pinned and stay valid);
// ab_sync waits and reports what failed on the device in between.  Handles own separate streams, so
// independent models queued this way overlap their transfers and steps (ensemble runs).
int ab_set_async(ab_handle* h, int on) {
    if (h->win_W) return ab_fail(h, AB_ERR_STATE, "not available on a sharded handle");
    if (!on && h->async_mode) { const int rc = ab_sync(h); h->async_mode = false; return rc; }
    h->async_mode = on != 0;
    return AB_OK;
}
int64_t ab_nagents(const ab_handle* h) { return h ? h->n : 0; }
int64_t ab_launch_count(const ab_handle* h) { return h ? h->launches : 0; }
double ab_last_step_ms(const ab_handle* h) { return h ? h->last_step_ms : 0; }

void* ab_alloc_pinned(int64_t bytes) {
    void* p = nullptr;
    if (bytes <= 0 || cudaMallocHost(&p, (size_t)bytes) != cudaSuccess) return nullptr;
    return p;
}
void ab_free_pinned(void* p) { if (p) cudaFreeHost(p); }

int ab_set_profiling(ab_handle* h, int on) {
    resolve_pending(h);
    h->profiling = on != 0;
    h->kstats.clear();
    return AB_OK;
}
int ab_kernel_times(ab_handle* h, const char** names, double* ms, int64_t* launches, int cap) {
    resolve_pending(h);
    int k = 0;
    for (auto& s : h->kstats) {
        if (k < cap) { if (names) names[k] = s.name; if (ms) ms[k] = s.ms; if (launches) launches[k] = s.launches; }
        ++k;
    }
    return k;
}

int ab_field_id(const ab_handle* h, const char* name) {
    for (size_t i = 0; i < h->fields.size(); ++i) if (h->fields[i].name == name) return (int)i;
    return AB_ERR_INVALID;
}

static int alloc_field(ab_handle* h, ab_field& f) {
    for (int b = 0; b < 2; ++b) {
        if (f.buf[b]) continue;
        size_t bytes = (size_t)f.ncomp * h->cap * f.esize;
        AB_CUDA(h, cudaMalloc(&f.buf[b], bytes ? bytes : 16));
        AB_CUDA(h, cudaMemsetAsync(f.buf[b], 0, bytes ? bytes : 16, h->stream));
    }
    return AB_OK;
}

int ab_register_field(ab_handle* h, const char* name, int dtype, int ncomp) {
    if (!name || ncomp < 1 || ncomp > 16 || ab_dtype_size(dtype) == 0) return ab_fail(h, AB_ERR_INVALID, "bad field spec");
    if (ab_field_id(h, name) >= 0) return ab_fail(h, AB_ERR_INVALID, std::string("field already registered: ") + name);
    AB_CUDA(h, cudaSetDevice(h->device));
    ab_field f;
    f.name = name; f.dtype = dtype; f.ncomp = ncomp; f.esize = ab_dtype_size(dtype);
    if (h->cap > 0) AB_TRY(alloc_field(h, f));
    h->fields.push_back(f);
    return (int)h->fields.size() - 1;
}

int ab_resize(ab_handle* h, int64_t n) {
    if (n < 0 || n > 2000000000LL) return ab_fail(h, AB_ERR_INVALID, "bad agent count");
    AB_CUDA(h, cudaSetDevice(h->device));
    AB_CUDA(h, cudaStreamSynchronize(h->stream));
    if (n > h->cap) {
        free_population(h);
        h->cap = (n + 63) / 64 * 64;    // 64-agent multiple keeps every component array 256 B aligned
        for (auto& f : h->fields) AB_TRY(alloc_field(h, f));
        size_t tb = (size_t)h->D * h->cap * h->tsize;
        AB_CUDA(h, cudaMalloc(&h->tmp_pos, tb));
        AB_CUDA(h, cudaMalloc(&h->tmp_vel, tb));
        size_t ib = (size_t)h->cap * 4;
        void** ints[] = {(void**)&h->id[0], (void**)&h->id[1], (void**)&h->key[0], (void**)&h->key[1],
                         (void**)&h->newkey, (void**)&h->rank, (void**)&h->idtmp, (void**)&h->perm,
                         (void**)&h->slot_of_id, (void**)&h->ncount};
        for (void** p : ints) { AB_CUDA(h, cudaMalloc(p, ib)); AB_CUDA(h, cudaMemsetAsync(*p, 0, ib, h->stream)); }
    }
    h->n = n;
    if (n > 0) {
        AB_LAUNCH(h, "k_iota");
        k_iota<<<ab_blocks(n, 256), 256, 0, h->stream>>>(h->id[h->cur], h->slot_of_id, (int)n);
    }
    h->slot_map_valid = true;
    h->index_valid = false;
    h->pairs_n = -1;
    h->grid_cells_valid = false;
    h->grid_cols_dirty = false;
    if (h->grid_ids) AB_CUDA(h, cudaMemsetAsync(h->grid_ids, 0, (size_t)h->ncells * sizeof(int32_t), h->stream));
    AB_CUDA(h, cudaStreamSynchronize(h->stream));
    return AB_OK;
}

int ab_upload(ab_handle* h, int field, const void* host, int64_t n) {
    if (field < 0 || field >= (int)h->fields.size()) return ab_fail(h, AB_ERR_INVALID, "bad field id");
    if (n != h->n) return ab_fail(h, AB_ERR_INVALID, "upload length must equal the agent count (ab_resize first)");
    if (n == 0) return AB_OK;
    if (!host) return ab_fail(h, AB_ERR_INVALID, "null host pointer");
    AB_CUDA(h, cudaSetDevice(h->device));
    if (h->grid_cols_dirty) AB_TRY(ab_grid_sync_columns(h));
    h->grid_cells_valid = false;
    ab_field& f = h->fields[field];
    size_t bytes = (size_t)n * f.ncomp * f.esize;
    void* stage;
    AB_TRY(ab_scratch(h, bytes, &stage));
    AB_CUDA(h, cudaMemcpyAsync(stage, host, bytes, cudaMemcpyHostToDevice, h->stream));
    AB_TRY(ab_ensure_slot_map(h));
    int64_t total = n * f.ncomp;
    {
        AB_LAUNCH(h, "k_upload");
        if (f.esize == 8) k_upload<uint64_t><<<ab_blocks(total, 256), 256, 0, h->stream>>>((const uint64_t*)stage, (uint64_t*)f.buf[h->cur], h->slot_of_id, h->cap, (int)n, f.ncomp);
        else if (f.esize == 4) k_upload<uint32_t><<<ab_blocks(total, 256), 256, 0, h->stream>>>((const uint32_t*)stage, (uint32_t*)f.buf[h->cur], h->slot_of_id, h->cap, (int)n, f.ncomp);
        else k_upload<uint8_t><<<ab_blocks(total, 256), 256, 0, h->stream>>>((const uint8_t*)stage, (uint8_t*)f.buf[h->cur], h->slot_of_id, h->cap, (int)n, f.ncomp);
    }
    if (!h->async_mode) AB_CUDA(h, cudaStreamSynchronize(h->stream));
    if (field == 0) { h->index_valid = false; h->pairs_n = -1; }
    // the per-cell sums Σ(pos - centre), Σ vel of the approximate flock sweep were built from the old columns
    if (field <= 1 && h->desc.kind == AB_SPACE_CONTINUOUS) h->agg_valid = false;
    return AB_OK;
}

int ab_download(ab_handle* h, int field, void* host, int64_t n) {
    if (field < 0 || field >= (int)h->fields.size()) return ab_fail(h, AB_ERR_INVALID, "bad field id");
    if (h->win_W) return ab_fail(h, AB_ERR_STATE, "use ab_download_local on a sharded handle");
    if (n != h->n) return ab_fail(h, AB_ERR_INVALID, "download length must equal the agent count");
    if (n == 0) return AB_OK;
    if (!host) return ab_fail(h, AB_ERR_INVALID, "null host pointer");
    AB_CUDA(h, cudaSetDevice(h->device));
    if (h->grid_cols_dirty) AB_TRY(ab_grid_sync_columns(h));
    ab_field& f = h->fields[field];
    size_t bytes = (size_t)n * f.ncomp * f.esize;
    void* stage;
    AB_TRY(ab_scratch(h, bytes, &stage));
    {
        AB_LAUNCH(h, "k_download");
        if (f.esize == 8) k_download<uint64_t><<<ab_blocks(n, 256), 256, 0, h->stream>>>((const uint64_t*)f.buf[h->cur], (uint64_t*)stage, h->id[h->cur], h->cap, (int)n, f.ncomp);
        else if (f.esize == 4) k_download<uint32_t><<<ab_blocks(n, 256), 256, 0, h->stream>>>((const uint32_t*)f.buf[h->cur], (uint32_t*)stage, h->id[h->cur], h->cap, (int)n, f.ncomp);
        else k_download<uint8_t><<<ab_blocks(n, 256), 256, 0, h->stream>>>((const uint8_t*)f.buf[h->cur], (uint8_t*)stage, h->id[h->cur], h->cap, (int)n, f.ncomp);
    }
    AB_CUDA(h, cudaMemcpyAsync(host, stage, bytes, cudaMemcpyDeviceToHost, h->stream));
    if (!h->async_mode) AB_CUDA(h, cudaStreamSynchronize(h->stream));
    return AB_OK;
}

int ab_build_index(ab_handle* h) {
    AB_CUDA(h, cudaSetDevice(h->device));
    if (h->desc.kind != AB_SPACE_CONTINUOUS) {
        return ab_grid_build(h);
    }
    h->index_valid = false;
    return ab_ensure_index(h);
}

// Checkpoint support — AgentsIO.save_checkpoint / load_checkpoint (src/submodules/io/jld2_integration.jl:85-135,
// :196-223) keep the space's stored_ids, i.e. the order of the agents inside every cell, and a resumed run
// continues exactly.  The device equivalent of that order is the slot order: floating-point sums over a cell
// follow it, so a run restored with ab_layout_set continues bit for bit.
int ab_layout_get(ab_handle* h, int64_t* slot_ids, int64_t n, int32_t* refinement, int64_t* tick) {
    if (h->win_W) return ab_fail(h, AB_ERR_STATE, "not available on a sharded handle");
    if (n != h->n) return ab_fail(h, AB_ERR_INVALID, "length must equal the agent count");
    AB_CUDA(h, cudaSetDevice(h->device));
    if (refinement) *refinement = h->K;
    if (tick) *tick = (int64_t)h->sch_tick;
    if (!slot_ids || n == 0) return AB_OK;
    if (h->desc.kind != AB_SPACE_CONTINUOUS) {               // grid columns stay in id order
        for (int64_t i = 0; i < n; ++i) slot_ids[i] = i + 1;
        return AB_OK;
    }
    AB_TRY(ab_ensure_index(h));
    std::vector<int32_t> ids((size_t)n);
    AB_CUDA(h, cudaMemcpyAsync(ids.data(), h->id[h->cur], (size_t)n * 4, cudaMemcpyDeviceToHost, h->stream));
    AB_CUDA(h, cudaStreamSynchronize(h->stream));
    for (int64_t i = 0; i < n; ++i) slot_ids[i] = (int64_t)ids[(size_t)i] + 1;
    return AB_OK;
}

int ab_layout_set(ab_handle* h, const int64_t* slot_ids, int64_t n, int32_t refinement, int64_t tick) {
    if (h->win_W) return ab_fail(h, AB_ERR_STATE, "not available on a sharded handle");
    if (n != h->n) return ab_fail(h, AB_ERR_INVALID, "length must equal the agent count");
    if (tick < 0) return ab_fail(h, AB_ERR_INVALID, "negative tick");
    AB_CUDA(h, cudaSetDevice(h->device));
    AB_CUDA(h, cudaStreamSynchronize(h->stream));
    h->sch_tick = (uint64_t)tick;
    if (h->desc.kind != AB_SPACE_CONTINUOUS) return AB_OK;
    AB_TRY(ab_set_refine(h, refinement));
    if (!slot_ids || n == 0) return AB_OK;
    std::vector<int32_t> prio((size_t)n, -1);
    for (int64_t i = 0; i < n; ++i) {
        const int64_t id = slot_ids[i];
        if (id < 1 || id > n || prio[(size_t)id - 1] >= 0) return ab_fail(h, AB_ERR_INVALID, "slot order is not a permutation of 1..n");
        prio[(size_t)id - 1] = (int32_t)i;
    }
    int32_t* dprio = nullptr;
    AB_CUDA(h, cudaMalloc(&dprio, (size_t)n * 4));
    cudaError_t ce = cudaMemcpyAsync(dprio, prio.data(), (size_t)n * 4, cudaMemcpyHostToDevice, h->stream);
    int rc = AB_OK;
    if (ce == cudaSuccess) {
        h->place_prio = dprio;
        h->index_valid = false;
        rc = ab_reindex_from(h, h->fields[0].buf[h->cur], h->fields[1].buf[h->cur], false);
        h->place_prio = nullptr;
        if (rc == AB_OK) {
            rc = ab_check_device_errors_now(h);                // synchronises: prio and dprio may go
        } else cudaStreamSynchronize(h->stream);
    }
    cudaFree(dprio);
    AB_CUDA(h, ce);
    return rc;
}

// 1 when every index build so far ordered every cell canonically (by id, or by the restored checkpoint order) — the
// premise of bitwise reproducible runs, N-GPU == 1-GPU equality and bit-for-bit checkpoint resume.  A cell holding
// more than AB_CANON_MAX (4096) agents at an index BUILD keeps the arrival order of the atomics instead (ranking is
// O(m^2) per cell); results stay correct within rounding, the flag says the guarantee is gone.  The tiled re-sort of the
// fused steps is canonical for any occupancy.
int ab_index_canonical(ab_handle* h) {
    if (!h) return AB_ERR_INVALID;
    AB_CUDA(h, cudaSetDevice(h->device));
    int v = 0;
    AB_CUDA(h, cudaMemcpyAsync(&v, h->d_err + 2, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    AB_CUDA(h, cudaStreamSynchronize(h->stream));
    return v ? 0 : 1;
}

int ab_cells(ab_handle* h, int32_t* out_cells, int64_t n) {
    if (h->desc.kind != AB_SPACE_CONTINUOUS) return ab_fail(h, AB_ERR_INVALID, "ab_cells is for ContinuousSpace handles");
    if (h->win_W) return ab_fail(h, AB_ERR_STATE, "not available on a sharded handle");
    if (n != h->n) return ab_fail(h, AB_ERR_INVALID, "length must equal the agent count");
    if (n == 0) return AB_OK;
    AB_CUDA(h, cudaSetDevice(h->device));
    size_t bytes = (size_t)n * h->D * sizeof(int32_t);
    void* stage;
    AB_TRY(ab_scratch(h, bytes, &stage));
    {
        AB_LAUNCH(h, "k_cells");
        const int TB = 256;
#define CELLS(T, DD) k_cells<T, DD><<<ab_blocks(n, TB), TB, 0, h->stream>>>(make_space<T, DD>(h), (const T*)h->fields[0].buf[h->cur], h->id[h->cur], h->cap, (int)n, (int32_t*)stage)
        if (h->dtype == AB_F64) { if (h->D == 2) CELLS(double, 2); else CELLS(double, 3); }
        else { if (h->D == 2) CELLS(float, 2); else CELLS(float, 3); }
#undef CELLS
    }
    AB_CUDA(h, cudaMemcpyAsync(out_cells, stage, bytes, cudaMemcpyDeviceToHost, h->stream));
    AB_CUDA(h, cudaStreamSynchronize(h->stream));
    return AB_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------
// FP64 FMA peak of the device, measured: the denominator for every "FP64-pipe bound" statement about the pair sweeps
// (secondary limiter).  8 independent fma chains per thread, 1024 threads per SM resident, CUDA events
// around the launch; best of 3.  Nothing is allocated: the sink is a device global that is never actually written
// (the condition cannot hold), it only keeps the chains alive.
// ---------------------------------------------------------------------------
__device__ double g_fp64_sink;
#define FP64_PEAK_ITERS 8192
#define FP64_PEAK_CHAINS 8
__global__ void __launch_bounds__(256) k_fp64_peak(double x, double y) {
    double a[FP64_PEAK_CHAINS];
#pragma unroll
    for (int k = 0; k < FP64_PEAK_CHAINS; ++k) a[k] = (double)(threadIdx.x + k) * 1e-3;
#pragma unroll 1
    for (int it = 0; it < FP64_PEAK_ITERS / 4; ++it) {
#pragma unroll
        for (int u = 0; u < 4; ++u)
#pragma unroll
            for (int k = 0; k < FP64_PEAK_CHAINS; ++k) a[k] = fma(a[k], x, y);
    }
    double s = 0;
#pragma unroll
    for (int k = 0; k < FP64_PEAK_CHAINS; ++k) s += a[k];
    if (s == 1.2345e300) g_fp64_sink = s;
}

extern "C" int ab_fp64_peak(int device, double* tflops) {
    if (!tflops) return ab_fail(nullptr, AB_ERR_INVALID, "ab_fp64_peak: null output");
    AB_CUDA(nullptr, cudaSetDevice(device));
    cudaDeviceProp pr;
    AB_CUDA(nullptr, cudaGetDeviceProperties(&pr, device));
    const int blocks = pr.multiProcessorCount * 4 * 4;             // 4 resident CTAs of 256 threads per SM, 4 waves
    cudaEvent_t e0, e1;
    AB_CUDA(nullptr, cudaEventCreate(&e0));
    AB_CUDA(nullptr, cudaEventCreate(&e1));
    double best = 0;
    for (int rep = 0; rep < 4; ++rep) {                            // rep 0 warms up
        cudaEventRecord(e0, 0);
        k_fp64_peak<<<blocks, 256>>>(0.999999, 1e-9);
        cudaEventRecord(e1, 0);
        const cudaError_t e = cudaEventSynchronize(e1);
        if (e != cudaSuccess) { cudaEventDestroy(e0); cudaEventDestroy(e1); return ab_cuda_fail(nullptr, e, "k_fp64_peak"); }
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        const double flop = 2.0 * FP64_PEAK_ITERS * FP64_PEAK_CHAINS * 256.0 * blocks;
        if (rep > 0 && ms > 0) best = std::max(best, flop / (ms * 1e-3) / 1e12);
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    *tflops = best;
    return AB_OK;
}
