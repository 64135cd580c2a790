// ab_query.cu — batched neighbour queries on the cell list.
//   nearby_ids(pos | agent, model, r; search)   src/spaces/continuous.jl:212-239,
//                                               src/core/space_interaction_API.jl:284-287
//   interacting_pairs(model, r, :all; search)   src/spaces/continuous.jl:418-452
//   nearest_neighbor(agent, model, r)           src/spaces/continuous.jl:298-307
// Every query walks exactly the cells the reference's GridSpaceIdIterator would
// (per-query grid_r, euclidean offset mask, wrap / skip rules), so id sets are
// bit-identical to the reference's — including its quirks (the approximate set
// is asymmetric; `:exact` only filters the approximate candidates).
#include <unordered_map>
#include <unordered_set>

#include "ab_internal.cuh"

// euclidean_distance² as the reference sums it (utilities.jl:84-116), in T
template <class T, int D>
__device__ __forceinline__ T ref_dist2(const SpaceDev<T, D>& sp, const T* a, const T* b) {
    T s = 0;
#pragma unroll
    for (int d = 0; d < D; ++d) {
        T direct = fabs(a[d] - b[d]);
        T m = sp.periodic[d] ? (direct < sp.extent[d] - direct ? direct : sp.extent[d] - direct) : direct;
        T sq = m * m;
        s = (d == 0) ? sq : s + sq;
    }
    return s;
}

template <class T, int D>
__device__ __forceinline__ bool load_query(const SpaceDev<T, D>& sp, const T* __restrict__ qpos,
                                           const int64_t* __restrict__ qids, const int32_t* __restrict__ slot_of_id,
                                           const T* __restrict__ pos, int64_t cap, int n, int q, T* p, int* self_slot,
                                           int* err) {
    *self_slot = -1;
    if (qids) {
        int64_t id = qids[q];
        if (id < 1 || id > n) { atomicOr(err, AB_EF_OVERFLOW); return false; }
        int s = slot_of_id[id - 1];
        *self_slot = s;
#pragma unroll
        for (int d = 0; d < D; ++d) p[d] = pos[(size_t)d * cap + s];
    } else {
#pragma unroll
        for (int d = 0; d < D; ++d) p[d] = qpos[(size_t)q * D + d];
    }
    return true;
}

// FILL = false: counts[q]; FILL = true: ids at offsets[q]
template <class T, int D, bool FILL>
__global__ void k_nearby(SpaceDev<T, D> sp, const T* __restrict__ qpos, const int64_t* __restrict__ qids, int nq,
                         double r, int exact, int exclude_self, const T* __restrict__ pos,
                         const int32_t* __restrict__ id, const int32_t* __restrict__ slot_of_id,
                         const uint32_t* __restrict__ cell_start, int64_t cap, int n, int64_t* __restrict__ counts,
                         const int64_t* __restrict__ offsets, int64_t* __restrict__ out, int* __restrict__ err) {
    int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= nq) return;
    T p[D];
    int self_slot;
    if (!load_query<T, D>(sp, qpos, qids, slot_of_id, pos, cap, n, q, p, &self_slot, err)) {
        if (!FILL) counts[q] = 0;
        return;
    }
    if (!exclude_self) self_slot = -1;
    int c[D];
    if (!pos2cell0<T, D>(sp, p, c)) {   // query point beyond the grid (or NaN)
        atomicOr(err, AB_EF_OUTSIDE);
        if (!FILL) counts[q] = 0;
        return;
    }
    const int gr = grid_radius<T, D>(sp, p, c, r);
    int64_t k = 0;
    int64_t base = FILL ? offsets[q] : 0;
    for_each_run<D>(sp.dims, sp.periodic, AB_METRIC_EUCLIDEAN, c, gr, cell_start, sp.kb, false,
                    [&](uint32_t b, uint32_t e, int) {
        for (uint32_t j = b; j < e; ++j) {
            if ((int)j == self_slot) continue;
            if (exact) {
                T o[D];
#pragma unroll
                for (int d = 0; d < D; ++d) o[d] = pos[(size_t)d * cap + j];
                T dist = ab_sqrt(ref_dist2<T, D>(sp, p, o));
                if (!((double)dist <= r)) continue;     // continuous.jl:237-238
            }
            if (FILL) out[base + k] = (int64_t)id[j] + 1;
            ++k;
        }
    });
    if (!FILL) counts[q] = k;
}

// interacting_pairs(:all).  Thread = slot s (agent i).  For each j in N(i):
// the pair is emitted by the smaller id if it sees the larger one, otherwise by
// the larger id (all_pairs! adds (min,max) when EITHER direction finds it).
template <class T, int D, bool FILL>
__global__ void k_pairs(SpaceDev<T, D> sp, double r, int exact, const T* __restrict__ pos,
                        const int32_t* __restrict__ id, const uint32_t* __restrict__ key,
                        const uint32_t* __restrict__ cell_start, int64_t cap, int n,
                        unsigned long long* __restrict__ counts, const unsigned long long* __restrict__ offsets,
                        int64_t* __restrict__ lo, int64_t* __restrict__ hi) {
    int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    T p[D];
    int c[D];
#pragma unroll
    for (int d = 0; d < D; ++d) p[d] = pos[(size_t)d * cap + s];
    key_to_cell<T, D>(sp, key[s], c);
    const int gr = grid_radius<T, D>(sp, p, c, r);
    const int me = id[s];
    unsigned long long k = 0;
    const unsigned long long base = FILL ? offsets[s] : 0ull;
    for_each_run<D>(sp.dims, sp.periodic, AB_METRIC_EUCLIDEAN, c, gr, cell_start, sp.kb, false,
                    [&](uint32_t b, uint32_t e, int) {
        for (uint32_t j = b; j < e; ++j) {
            if ((int)j == s) continue;
            T o[D];
#pragma unroll
            for (int d = 0; d < D; ++d) o[d] = pos[(size_t)d * cap + j];
            if (exact) {
                T dist = ab_sqrt(ref_dist2<T, D>(sp, p, o));
                if (!((double)dist <= r)) continue;
            }
            const int other = id[j];
            if (other < me) {
                // does `other` see me?  then it emits the pair itself
                int cj[D];
                key_to_cell<T, D>(sp, key[j], cj);
                const int grj = grid_radius<T, D>(sp, o, cj, r);
                if (cell_reachable<D>(sp.dims, sp.periodic, AB_METRIC_EUCLIDEAN, cj, c, grj)) continue;
            }
            if (FILL) {
                lo[base + k] = (int64_t)(other < me ? other : me) + 1;
                hi[base + k] = (int64_t)(other < me ? me : other) + 1;
            }
            ++k;
        }
    });
    if (!FILL) counts[s] = k;
}

template <class T, int D>
__global__ void k_nearest(SpaceDev<T, D> sp, double r, const T* __restrict__ pos, const int32_t* __restrict__ id,
                          const uint32_t* __restrict__ key, const uint32_t* __restrict__ cell_start, int64_t cap,
                          int n, int64_t* __restrict__ out, double* __restrict__ dist) {
    int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    T p[D];
    int c[D];
#pragma unroll
    for (int d = 0; d < D; ++d) p[d] = pos[(size_t)d * cap + s];
    key_to_cell<T, D>(sp, key[s], c);
    const int gr = grid_radius<T, D>(sp, p, c, r);
    double best = INFINITY;
    int bj = -1;
    for_each_run<D>(sp.dims, sp.periodic, AB_METRIC_EUCLIDEAN, c, gr, cell_start, sp.kb, false,
                    [&](uint32_t b, uint32_t e, int) {
        for (uint32_t j = b; j < e; ++j) {
            if ((int)j == s) continue;
            T o[D];
#pragma unroll
            for (int d = 0; d < D; ++d) o[d] = pos[(size_t)d * cap + j];
            double dn = (double)ab_sqrt(ref_dist2<T, D>(sp, p, o));
            if (dn < best) { best = dn; bj = (int)j; }      // continuous.jl:301-304
        }
    });
    out[id[s]] = (best > r || bj < 0) ? 0 : (int64_t)id[bj] + 1;     // continuous.jl:306
    if (dist) dist[id[s]] = best;                                    // euclidean_distance to it (true_pairs!, :469)
}

// --- small 64-bit exclusive scan (query helper, not on the step path) --------
#define QS_T 256
__global__ void k_qscan_partial(const unsigned long long* __restrict__ in, int n, unsigned long long* __restrict__ part) {
    __shared__ unsigned long long sh[QS_T / 32];
    int i = blockIdx.x * QS_T + threadIdx.x;
    unsigned long long v = i < n ? in[i] : 0ull;
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x == 0) { unsigned long long t = 0; for (int w = 0; w < QS_T / 32; ++w) t += sh[w]; part[blockIdx.x] = t; }
}
__global__ void k_qscan_top(unsigned long long* part, int nb, unsigned long long* total) {
    // one thread: nb = n/256 (≤ a few 10^5), query-time only
    unsigned long long run = 0;
    for (int i = 0; i < nb; ++i) { unsigned long long t = part[i]; part[i] = run; run += t; }
    *total = run;
}
__global__ void k_qscan_final(const unsigned long long* __restrict__ in, int n, const unsigned long long* __restrict__ part,
                              unsigned long long* __restrict__ out) {
    __shared__ unsigned long long sh[QS_T / 32];
    int i = blockIdx.x * QS_T + threadIdx.x;
    unsigned long long v = i < n ? in[i] : 0ull, inc = v;
    int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (int o = 1; o < 32; o <<= 1) { unsigned long long t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += t; }
    if (lane == 31) sh[w] = inc;
    __syncthreads();
    unsigned long long woff = 0;
    for (int k = 0; k < w; ++k) woff += sh[k];
    if (i < n) out[i] = part[blockIdx.x] + woff + inc - v;
}

template <class T, int D>
static int nearby_t(ab_handle* h, const void* qpos_dev, const int64_t* qids_dev, int nq, double r, int search,
                    int exclude_self, int64_t* counts_dev, const int64_t* offsets_dev, int64_t* out_dev) {
    SpaceDev<T, D> sp = make_space<T, D>(h);
    const int cur = h->cur;
    AB_LAUNCH(h, out_dev ? "k_nearby_fill" : "k_nearby_count");
    if (out_dev)
        k_nearby<T, D, true><<<ab_blocks(nq, 128), 128, 0, h->stream>>>(
            sp, (const T*)qpos_dev, qids_dev, nq, r, search == AB_SEARCH_EXACT, exclude_self,
            (const T*)h->fields[0].buf[cur], h->id[cur], h->slot_of_id, h->cell_start, h->cap, (int)h->n, nullptr,
            offsets_dev, out_dev, h->d_err);
    else
        k_nearby<T, D, false><<<ab_blocks(nq, 128), 128, 0, h->stream>>>(
            sp, (const T*)qpos_dev, qids_dev, nq, r, search == AB_SEARCH_EXACT, exclude_self,
            (const T*)h->fields[0].buf[cur], h->id[cur], h->slot_of_id, h->cell_start, h->cap, (int)h->n, counts_dev,
            nullptr, nullptr, h->d_err);
    return AB_OK;
}

static int nearby_common(ab_handle* h, const void* qpos, const int64_t* qids, int64_t nq, double r, int search,
                         int exclude_self, int64_t* counts, const int64_t* offsets, int64_t* ids_out) {
    if (h->win_W) return ab_fail(h, AB_ERR_STATE, "neighbour queries are not available on a sharded (slab-window) handle");
    if ((qpos == nullptr) == (qids == nullptr)) return ab_fail(h, AB_ERR_INVALID, "exactly one of qpos / qids must be given");
    if (search != AB_SEARCH_APPROXIMATE && search != AB_SEARCH_EXACT)
        return ab_fail(h, AB_ERR_INVALID, "`search` keyword should be either `:approximate` or `:exact`");
    if (nq < 0 || nq > 2000000000LL) return ab_fail(h, AB_ERR_INVALID, "bad query count");
    AB_CUDA(h, cudaSetDevice(h->device));
    if (h->desc.kind != AB_SPACE_CONTINUOUS) {
        return ab_grid_nearby(h, qpos, qids, nq, r, exclude_self, counts, offsets, ids_out);
    }
    if (nq == 0) return AB_OK;
    AB_TRY(ab_ensure_index(h));
    if (qids) AB_TRY(ab_ensure_slot_map(h));
    const bool fill = ids_out != nullptr;
    size_t qbytes = qpos ? (size_t)nq * h->D * h->tsize : (size_t)nq * sizeof(int64_t);
    size_t o1 = (qbytes + 255) / 256 * 256;
    size_t cbytes = (size_t)(nq + 1) * sizeof(int64_t);
    size_t o2 = o1 + (cbytes + 255) / 256 * 256;
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
    const void* qp = qpos ? base : nullptr;
    const int64_t* qi = qids ? (const int64_t*)base : nullptr;
    int rc;
#define NB(T, DD) nearby_t<T, DD>(h, qp, qi, (int)nq, r, search, exclude_self, fill ? nullptr : cdev, fill ? cdev : nullptr, fill ? odev : nullptr)
    if (h->dtype == AB_F64) rc = h->D == 2 ? NB(double, 2) : NB(double, 3);
    else rc = h->D == 2 ? NB(float, 2) : NB(float, 3);
#undef NB
    AB_TRY(rc);
    if (fill) { if (total > 0) AB_CUDA(h, cudaMemcpyAsync(ids_out, odev, (size_t)total * sizeof(int64_t), cudaMemcpyDeviceToHost, h->stream)); }
    else AB_CUDA(h, cudaMemcpyAsync(counts, cdev, (size_t)nq * sizeof(int64_t), cudaMemcpyDeviceToHost, h->stream));
    int e = ab_check_device_errors_now(h);               // hands back host data: synchronous in every mode
    if (e == AB_ERR_INVALID) return ab_fail(h, AB_ERR_INVALID, "agent id out of range");
    return e;
}

template <class T, int D>
static int pairs_t(ab_handle* h, double r, int search, bool fill, unsigned long long* counts, int64_t* lo, int64_t* hi) {
    SpaceDev<T, D> sp = make_space<T, D>(h);
    const int cur = h->cur;
    const int n = (int)h->n;
    AB_LAUNCH(h, fill ? "k_pairs_fill" : "k_pairs_count");
    if (fill)
        k_pairs<T, D, true><<<ab_blocks(n, 128), 128, 0, h->stream>>>(sp, r, search == AB_SEARCH_EXACT, (const T*)h->fields[0].buf[cur], h->id[cur], h->key[cur], h->cell_start, h->cap, n, nullptr, h->pairs_offsets, lo, hi);
    else
        k_pairs<T, D, false><<<ab_blocks(n, 128), 128, 0, h->stream>>>(sp, r, search == AB_SEARCH_EXACT, (const T*)h->fields[0].buf[cur], h->id[cur], h->key[cur], h->cell_start, h->cap, n, counts, nullptr, nullptr, nullptr);
    return AB_OK;
}

extern "C" {

int ab_nearby_count(ab_handle* h, const void* qpos, const int64_t* qids, int64_t nq, double r, int search,
                    int exclude_self, int64_t* counts) {
    if (!counts && nq > 0) return ab_fail(h, AB_ERR_INVALID, "null counts");
    return nearby_common(h, qpos, qids, nq, r, search, exclude_self, counts, nullptr, nullptr);
}
int ab_nearby_fill(ab_handle* h, const void* qpos, const int64_t* qids, int64_t nq, double r, int search,
                   int exclude_self, const int64_t* offsets, int64_t* ids_out) {
    if (!ids_out && nq > 0 && offsets && offsets[nq] > 0) return ab_fail(h, AB_ERR_INVALID, "null ids_out");
    if (!ids_out) { static int64_t dummy; ids_out = &dummy; }
    return nearby_common(h, qpos, qids, nq, r, search, exclude_self, nullptr, offsets, ids_out);
}

int ab_pairs_count(ab_handle* h, double r, int search, int64_t* npairs) {
    if (h->win_W) return ab_fail(h, AB_ERR_STATE, "neighbour queries are not available on a sharded (slab-window) handle");
    if (h->desc.kind != AB_SPACE_CONTINUOUS) return ab_fail(h, AB_ERR_INVALID, "interacting_pairs needs a ContinuousSpace handle");
    if (search != AB_SEARCH_APPROXIMATE && search != AB_SEARCH_EXACT) return ab_fail(h, AB_ERR_INVALID, "bad search mode");
    AB_CUDA(h, cudaSetDevice(h->device));
    *npairs = 0;
    if (h->n == 0) { h->pairs_n = 0; h->pairs_r = r; h->pairs_search = search; return AB_OK; }
    AB_TRY(ab_ensure_index(h));
    const int n = (int)h->n;
    const int nb = ab_blocks(n, QS_T);
    if (h->pairs_offsets_len < (size_t)n + 1) {
        if (h->pairs_offsets) AB_CUDA(h, cudaFree(h->pairs_offsets));
        h->pairs_offsets = nullptr;
        AB_CUDA(h, cudaMalloc((void**)&h->pairs_offsets, ((size_t)n + 1) * sizeof(unsigned long long)));
        h->pairs_offsets_len = (size_t)n + 1;
    }
    char* base;
    size_t cb = ((size_t)n * 8 + 255) / 256 * 256;
    AB_TRY(ab_scratch(h, cb + (size_t)(nb + 1) * 8 + 256, (void**)&base));
    unsigned long long* counts = (unsigned long long*)base;
    unsigned long long* part = (unsigned long long*)(base + cb);
    int rc;
    if (h->dtype == AB_F64) rc = h->D == 2 ? pairs_t<double, 2>(h, r, search, false, counts, nullptr, nullptr) : pairs_t<double, 3>(h, r, search, false, counts, nullptr, nullptr);
    else rc = h->D == 2 ? pairs_t<float, 2>(h, r, search, false, counts, nullptr, nullptr) : pairs_t<float, 3>(h, r, search, false, counts, nullptr, nullptr);
    AB_TRY(rc);
    {
        AB_LAUNCH(h, "k_qscan");
        k_qscan_partial<<<nb, QS_T, 0, h->stream>>>(counts, n, part);
        k_qscan_top<<<1, 1, 0, h->stream>>>(part, nb, part + nb);
        k_qscan_final<<<nb, QS_T, 0, h->stream>>>(counts, n, part, h->pairs_offsets);
    }
    unsigned long long total = 0;
    AB_CUDA(h, cudaMemcpyAsync(&total, part + nb, 8, cudaMemcpyDeviceToHost, h->stream));
    AB_CUDA(h, cudaStreamSynchronize(h->stream));
    h->pairs_n = (int64_t)total;
    h->pairs_r = r;
    h->pairs_search = search;
    *npairs = (int64_t)total;
    return AB_OK;
}

int ab_pairs_fill(ab_handle* h, int64_t* lo, int64_t* hi, int64_t npairs) {
    if (h->pairs_n < 0) return ab_fail(h, AB_ERR_STATE, "ab_pairs_count must be called first (and the positions unchanged since)");
    if (npairs != h->pairs_n) return ab_fail(h, AB_ERR_INVALID, "npairs does not match the last ab_pairs_count");
    if (npairs == 0) return AB_OK;
    if (!lo || !hi) return ab_fail(h, AB_ERR_INVALID, "null output");
    AB_CUDA(h, cudaSetDevice(h->device));
    int64_t* dev;
    AB_TRY(ab_scratch2(h, (size_t)npairs * 16 + 256, (void**)&dev));
    int64_t* dlo = dev;
    int64_t* dhi = dev + npairs;
    int rc;
    const double r = h->pairs_r;
    const int search = h->pairs_search;
    if (h->dtype == AB_F64) rc = h->D == 2 ? pairs_t<double, 2>(h, r, search, true, nullptr, dlo, dhi) : pairs_t<double, 3>(h, r, search, true, nullptr, dlo, dhi);
    else rc = h->D == 2 ? pairs_t<float, 2>(h, r, search, true, nullptr, dlo, dhi) : pairs_t<float, 3>(h, r, search, true, nullptr, dlo, dhi);
    AB_TRY(rc);
    AB_CUDA(h, cudaMemcpyAsync(lo, dlo, (size_t)npairs * 8, cudaMemcpyDeviceToHost, h->stream));
    AB_CUDA(h, cudaMemcpyAsync(hi, dhi, (size_t)npairs * 8, cudaMemcpyDeviceToHost, h->stream));
    AB_CUDA(h, cudaStreamSynchronize(h->stream));
    return AB_OK;
}

static int nearest_common(ab_handle* h, double r, int64_t* out, double* dist, int64_t n) {
    if (h->win_W) return ab_fail(h, AB_ERR_STATE, "neighbour queries are not available on a sharded (slab-window) handle");
    if (h->desc.kind != AB_SPACE_CONTINUOUS) return ab_fail(h, AB_ERR_INVALID, "nearest_neighbor needs a ContinuousSpace handle");
    if (n != h->n) return ab_fail(h, AB_ERR_INVALID, "length must equal the agent count");
    if (n == 0) return AB_OK;
    AB_CUDA(h, cudaSetDevice(h->device));
    AB_TRY(ab_ensure_index(h));
    int64_t* dev;
    AB_TRY(ab_scratch(h, (size_t)n * 16 + 256, (void**)&dev));
    double* ddev = dist ? (double*)(dev + n) : nullptr;
    {
        AB_LAUNCH(h, "k_nearest");
        const int cur = h->cur;
#define NN(T, DD) k_nearest<T, DD><<<ab_blocks(n, 128), 128, 0, h->stream>>>(make_space<T, DD>(h), r, (const T*)h->fields[0].buf[cur], h->id[cur], h->key[cur], h->cell_start, h->cap, (int)n, dev, ddev)
        if (h->dtype == AB_F64) { if (h->D == 2) NN(double, 2); else NN(double, 3); }
        else { if (h->D == 2) NN(float, 2); else NN(float, 3); }
#undef NN
    }
    AB_CUDA(h, cudaMemcpyAsync(out, dev, (size_t)n * 8, cudaMemcpyDeviceToHost, h->stream));
    if (dist) AB_CUDA(h, cudaMemcpyAsync(dist, ddev, (size_t)n * 8, cudaMemcpyDeviceToHost, h->stream));
    AB_CUDA(h, cudaStreamSynchronize(h->stream));
    return AB_OK;
}

int ab_nearest_neighbor(ab_handle* h, double r, int64_t* out, int64_t n) { return nearest_common(h, r, out, nullptr, n); }

// interacting_pairs(model, r, :nearest; scheduler) — true_pairs!, continuous.jl:454-497.  The nearest
// neighbours (and their distances) of ALL agents come from one k_nearest launch; the reference's pair
// bookkeeping is sequential in scheduler order and is restated here on the host with hash maps in place
// of its linear `findfirst` scans (same results: every id heads at most one pair, :471-479).
int ab_pairs_nearest(ab_handle* h, double r, const int64_t* order, int64_t norder, int64_t* npairs) {
    if (!npairs) return ab_fail(h, AB_ERR_INVALID, "null npairs");
    const int64_t n = h->n;
    h->near_lo.clear();
    h->near_hi.clear();
    *npairs = 0;
    if (n == 0) return AB_OK;
    std::vector<int64_t> nn((size_t)n);
    std::vector<double> dd((size_t)n);
    AB_TRY(nearest_common(h, r, nn.data(), dd.data(), n));
    std::vector<int64_t> lo, hi;
    std::vector<double> dist;
    std::unordered_map<int64_t, size_t> head;           // first element -> index of its (only) pair
    std::unordered_set<uint64_t> present;
    auto key = [](int64_t a, int64_t b) { return ((uint64_t)a << 32) | (uint64_t)b; };
    const int64_t cnt = order ? norder : n;
    for (int64_t t = 0; t < cnt; ++t) {
        const int64_t id = order ? order[t] : t + 1;
        if (id < 1 || id > n) return ab_fail(h, AB_ERR_INVALID, "scheduled id out of range");
        const int64_t j = nn[(size_t)id - 1];
        if (!j) continue;                                // nearest_neighbor === nothing (:459)
        const int64_t a = id < j ? id : j, b = id < j ? j : id;
        if (present.count(key(a, b))) continue;          // new_pair ∉ pairs (:462)
        const double d = dd[(size_t)id - 1];
        auto it = head.find(a);
        if (it == head.end()) {
            head[a] = lo.size();
            lo.push_back(a); hi.push_back(b); dist.push_back(d);
            present.insert(key(a, b));
        } else if (dist[it->second] > d) {               // replace: not the true neighbour (:474-478)
            const size_t i = it->second;
            present.erase(key(lo[i], hi[i]));
            hi[i] = b; dist[i] = d;
            present.insert(key(a, b));
        }
    }
    // ids that occur in more than one pair keep only their closest pair (:481-496)
    std::unordered_map<int64_t, int> occ;
    std::unordered_map<int64_t, std::vector<size_t>> tails;   // last element -> indices, ascending
    for (size_t i = 0; i < lo.size(); ++i) { occ[lo[i]]++; occ[hi[i]]++; tails[hi[i]].push_back(i); }
    std::vector<size_t> rm;
    for (auto& kv : occ) {
        if (kv.second < 2) continue;
        auto f = head.find(kv.first);
        auto& tl = tails[kv.first];
        if (f != head.end()) {
            if (tl.empty()) continue;
            const size_t fidx = f->second, lidx = tl[0];
            rm.push_back(dist[fidx] <= dist[lidx] ? lidx : fidx);
        } else {
            size_t keep = 0;
            for (size_t q = 1; q < tl.size(); ++q) if (dist[tl[q]] < dist[tl[keep]]) keep = q;   // findmin: first minimum
            for (size_t q = 0; q < tl.size(); ++q) if (q != keep) rm.push_back(tl[q]);
        }
    }
    std::vector<char> dead(lo.size(), 0);
    for (size_t i : rm) dead[i] = 1;
    for (size_t i = 0; i < lo.size(); ++i)
        if (!dead[i]) { h->near_lo.push_back(lo[i]); h->near_hi.push_back(hi[i]); }
    *npairs = (int64_t)h->near_lo.size();
    return AB_OK;
}

int ab_pairs_nearest_fill(ab_handle* h, int64_t* lo, int64_t* hi, int64_t npairs) {
    if (npairs != (int64_t)h->near_lo.size()) return ab_fail(h, AB_ERR_INVALID, "npairs does not match the last ab_pairs_nearest");
    if (npairs && (!lo || !hi)) return ab_fail(h, AB_ERR_INVALID, "null output");
    for (int64_t i = 0; i < npairs; ++i) { lo[i] = h->near_lo[(size_t)i]; hi[i] = h->near_hi[(size_t)i]; }
    return AB_OK;
}

}  // extern "C"
