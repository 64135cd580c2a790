// ab_resort.cu — tile-based, deterministic re-sort of the population after a
// fused step, and the per-reference-cell aggregates of the sorted state.
//
// The reference keeps its index current by moving one id between two cell
// vectors per moved agent (src/spaces/continuous.jl:133-153, :161-166).  Here a
// step leaves the new state in OLD slot order (tmp_pos / tmp_vel / newkey); this
// kernel brings it into the next step's cell-sorted order in ONE pass:
//
//   tile  = a run of `tw` fine cells of one FINE row = a contiguous range of
//           device keys = a contiguous range of output slots.
//   before  the fused step kernel has counted the new keys per fine cell
//           (red.add into cell_count) and k_scan has turned the counts into the
//           NEW cell_start, so every tile knows its output range up front and
//           tiles are independent of each other (a first version chained the
//           tiles by decoupled look-back: ~55 % of its time was spent waiting).
//   CTA   = one tile.  It scans the OLD slots of the 3^D
//           neighbouring fine cells of its tile (agents move less than one
//           fine cell per step on this path), keeps those whose new key
//           falls into the tile, buckets them per device cell in shared
//           memory, orders every cell's agents by id (canonical
//           order: bitwise reproducible runs whatever the atomic order was),
//           and writes pos / vel / id / key coalesced and (optionally) count, Σ(pos - centre), Σ vel of the
//           tile's part of every reference cell it touches (one partial per
//           fine row of a reference cell, combined in fixed order by k_cell_M).
//
// Nothing here is specific to the flock rule: any fused rule that bounds the
// per-step displacement by one reference cell can use it.
#include "ab_internal.cuh"

#define RS_THREADS 256
#define RS_MAXSEG 27
#define RS_FC_MAX 1024          // device cells per tile (shared histogram)
#define RS_CAP 2048             // agents per tile handled in shared memory; larger tiles use global scratch

struct ResortArgs {
    int tw, tpr, ntiles, nrows;
    int has_vel, want_agg;
};

template <class T, int D>
__global__ void __launch_bounds__(RS_THREADS)
k_resort(SpaceDev<T, D> sp, ResortArgs a, const uint32_t* __restrict__ cs_old, const uint32_t* __restrict__ newkey,
         const int32_t* __restrict__ id_old, const T* __restrict__ src_pos, const T* __restrict__ src_vel,
         int64_t cap, int n, T* __restrict__ dst_pos, T* __restrict__ dst_vel, int32_t* __restrict__ id_new,
         uint32_t* __restrict__ key_new, const uint32_t* __restrict__ cs_new, int32_t* __restrict__ perm,
         T* __restrict__ agg, int64_t agg_stride, int32_t* __restrict__ g_ent_id, int32_t* __restrict__ g_ent_s,
         int* __restrict__ err) {
    __shared__ uint32_t s_seg_beg[RS_MAXSEG], s_seg_end[RS_MAXSEG];
    __shared__ int s_nseg;
    __shared__ uint32_t s_cnt[RS_FC_MAX];        // per-cell cursor
    __shared__ uint32_t s_start[RS_FC_MAX + 1];  // new cell_start of the tile's cells, relative to the tile's base
    __shared__ uint32_t s_found;
    __shared__ int32_t s_ent_id[RS_CAP], s_ent_s[RS_CAP];
    __shared__ uint16_t s_ent_lc[RS_CAP];
    __shared__ int32_t s_sorted_s[RS_CAP];

    const int tid = threadIdx.x;
    const int tile = blockIdx.x;
    const int frow = tile / a.tpr, seg = tile - frow * a.tpr;      // fine row, segment of it
    const int kb = sp.kb;
    const int fd0 = sp.dims[0] << kb, fd1 = sp.dims[1] << kb, fd2 = D == 3 ? (sp.dims[2] << kb) : 1;
    const int x0 = seg * a.tw;
    const int x1 = min(fd0, x0 + a.tw);          // exclusive, fine cells
    const uint32_t kbeg = (uint32_t)frow * (uint32_t)fd0 + (uint32_t)x0;
    const uint32_t kend = (uint32_t)frow * (uint32_t)fd0 + (uint32_t)x1;
    const int fc = (int)(kend - kbeg);
    const int fy0 = D == 2 ? frow : frow % fd1;
    const int fz0 = D == 2 ? 0 : frow / fd1;

    // ---- source segments: old slots of the neighbouring fine cells ----------------------
    if (tid == 0) {
        int ns = 0;
        for (int dz = (D == 3 ? -1 : 0); dz <= (D == 3 ? 1 : 0); ++dz) {
            int z = fz0 + dz;
            if (D == 3 && (z < 0 || z >= fd2)) {
                if (!sp.periodic[2]) continue;
                z = z < 0 ? z + fd2 : z - fd2;
            }
            for (int dy = -1; dy <= 1; ++dy) {
                int y = fy0 + dy;
                if (y < 0 || y >= fd1) {
                    if (!sp.periodic[1]) continue;
                    y = y < 0 ? y + fd1 : y - fd1;
                }
                const uint32_t rb = (D == 3 ? (uint32_t)z * (uint32_t)fd1 + (uint32_t)y : (uint32_t)y) * (uint32_t)fd0;
                const int lo = max(x0 - 1, 0), hi = min(x1, fd0 - 1);      // inclusive fine cells
                s_seg_beg[ns] = cs_old[(size_t)rb + lo];
                s_seg_end[ns] = cs_old[(size_t)rb + hi + 1];
                ++ns;
                if (sp.periodic[0]) {
                    if (x0 == 0 && hi < fd0 - 1) {        // left neighbour wraps to the last cell of the row
                        s_seg_beg[ns] = cs_old[(size_t)rb + fd0 - 1];
                        s_seg_end[ns] = cs_old[(size_t)rb + fd0];
                        ++ns;
                    }
                    if (x1 == fd0 && lo > 0) {            // right neighbour wraps to the first cell
                        s_seg_beg[ns] = cs_old[(size_t)rb];
                        s_seg_end[ns] = cs_old[(size_t)rb + 1];
                        ++ns;
                    }
                }
            }
        }
        s_nseg = ns;
    }
    const uint32_t base = cs_new[kbeg];
    for (int i = tid; i <= fc; i += RS_THREADS) {
        const uint32_t st = cs_new[(size_t)kbeg + i] - base;
        s_start[i] = st;
        if (i < fc) s_cnt[i] = st;                       // cursor = start
    }
    if (tid == 0) s_found = 0;
    __syncthreads();
    const int nseg = s_nseg;
    const uint32_t total = s_start[fc];
    const bool small = total <= RS_CAP;
    int32_t* ent_id = small ? s_ent_id : g_ent_id + base;
    int32_t* ent_s = small ? s_ent_s : g_ent_s + base;
    uint32_t found = 0;
    if (total != 0) {
    // ---- entries into their cell segment (arrival order) ----------------------------------
    for (int sg = 0; sg < nseg; ++sg) {
        const uint32_t e = s_seg_end[sg];
        for (uint32_t s = s_seg_beg[sg] + tid; s < e; s += RS_THREADS) {
            const uint32_t k = newkey[s];
            if (k >= kbeg && k < kend) {
                const uint32_t lc = k - kbeg;
                const uint32_t p = atomicAdd(&s_cnt[lc], 1u);
                if (p < s_start[lc + 1]) {               // (always, unless the counts are inconsistent)
                    ent_id[p] = id_old[s];
                    ent_s[p] = (int32_t)s;
                    if (small) s_ent_lc[p] = (uint16_t)lc;
                }
                ++found;
            }
        }
    }
    // every agent counted into this tile must have been found in the 3^D neighbourhood
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) found += __shfl_xor_sync(0xffffffffu, found, o);
    if ((tid & 31) == 0 && found) atomicAdd(&s_found, found);
    __syncthreads();
    if (tid == 0 && s_found != total) atomicOr(err, AB_EF_OVERFLOW);     // an agent left its 3^D neighbourhood
    if (s_found != total) return;

    // ---- canonical order inside every cell (by id), then coalesced output ------------------
    if (small) {
        for (uint32_t p = tid; p < total; p += RS_THREADS) {
            const uint32_t lc = s_ent_lc[p];
            const uint32_t b = s_start[lc], e = s_start[lc + 1];
            const int32_t me = s_ent_id[p];
            uint32_t r = 0;
            for (uint32_t t = b; t < e; ++t) r += (s_ent_id[t] < me) ? 1u : 0u;
            s_sorted_s[b + r] = (int32_t)p;        // sorted position -> entry
        }
        __syncthreads();
        for (uint32_t t = tid; t < total; t += RS_THREADS) {
            const int p = s_sorted_s[t];
            const int s = s_ent_s[p];
            const size_t o = (size_t)base + t;
#pragma unroll
            for (int d = 0; d < D; ++d) dst_pos[(size_t)d * cap + o] = src_pos[(size_t)d * cap + s];
            if (a.has_vel) {
#pragma unroll
                for (int d = 0; d < D; ++d) dst_vel[(size_t)d * cap + o] = src_vel[(size_t)d * cap + s];
            }
            id_new[o] = s_ent_id[p];
            key_new[o] = kbeg + s_ent_lc[p];
            if (perm) perm[o] = s;
        }
    } else {
        // crowded tile: entries live in global scratch at [base, base + total); same steps through L2
        for (uint32_t p = tid; p < total; p += RS_THREADS) {
            const uint32_t k = newkey[ent_s[p]];
            const uint32_t lc = k - kbeg;
            const uint32_t b = s_start[lc], e = s_start[lc + 1];
            const int32_t me = ent_id[p];
            uint32_t r = p - b;                                  // arrival order for very crowded cells
            if (e - b <= 4096u) { r = 0; for (uint32_t t = b; t < e; ++t) r += (ent_id[t] < me) ? 1u : 0u; }
            const int s = ent_s[p];
            const size_t o = (size_t)base + b + r;
#pragma unroll
            for (int d = 0; d < D; ++d) dst_pos[(size_t)d * cap + o] = src_pos[(size_t)d * cap + s];
            if (a.has_vel) {
#pragma unroll
                for (int d = 0; d < D; ++d) dst_vel[(size_t)d * cap + o] = src_vel[(size_t)d * cap + s];
            }
            id_new[o] = me;
            key_new[o] = k;
            if (perm) perm[o] = s;
        }
    }

    }   // total != 0

    // ---- per reference cell: count, Σ(pos - centre), Σ vel of this fine row's part, canonical order ---
    if (a.want_agg) {
        __syncthreads();     // the tile's output slots are complete (global writes of this CTA are visible to it)
        constexpr int NC = 1 + 2 * D;
        const int K = sp.K;
        const int ncell = (x1 - x0) >> kb;                    // x0 and tw are multiples of K
        const int grp = tid >> 3, gl = tid & 7;               // 8 lanes per reference cell
        const int cy = fy0 >> kb, cz = fz0 >> kb;
        const int part = (fy0 & (K - 1)) + K * (fz0 & (K - 1));
        for (int c = grp; c < ncell; c += RS_THREADS / 8) {
            const uint32_t b = base + s_start[(size_t)c << kb], e = base + s_start[(size_t)(c + 1) << kb];
            const int cx = (x0 >> kb) + c;
            const uint32_t refcell = (D == 3 ? ((uint32_t)cz * sp.dims[1] + cy) : (uint32_t)cy) * (uint32_t)sp.dims[0] + (uint32_t)cx;
            int cc[3] = {cx, cy, cz};
            cc[D - 1] = local_to_global_last<T, D>(sp, cc[D - 1]);     // centres are global (slab window)
            T acc[2 * D];
#pragma unroll
            for (int d = 0; d < 2 * D; ++d) acc[d] = T(0);
            for (uint32_t t = b + gl; t < e; t += 8) {
#pragma unroll
                for (int d = 0; d < D; ++d) {
                    const T centre = sp.spacing * ((T)(cc[d] + 1) - T(0.5));
                    acc[d] += dst_pos[(size_t)d * cap + t] - centre;
                    acc[D + d] += dst_vel[(size_t)d * cap + t];
                }
            }
#pragma unroll
            for (int d = 0; d < 2 * D; ++d) {
                acc[d] += __shfl_xor_sync(0xffffffffu, acc[d], 4);
                acc[d] += __shfl_xor_sync(0xffffffffu, acc[d], 2);
                acc[d] += __shfl_xor_sync(0xffffffffu, acc[d], 1);
            }
            if (gl == 0) {
                T* dst = agg + (size_t)part * NC * agg_stride + refcell;
                dst[0] = (T)(e - b);
#pragma unroll
                for (int d = 0; d < 2 * D; ++d) dst[(size_t)(1 + d) * agg_stride] = acc[d];
            }
        }
    }
}

// Aggregates of an already sorted state (after a generic index build): one
// 8-lane group per (reference cell, fine row of it), same summation order as k_resort.
template <class T, int D>
__global__ void __launch_bounds__(256)
k_cell_agg(SpaceDev<T, D> sp, const uint32_t* __restrict__ cs, const T* __restrict__ pos, const T* __restrict__ vel,
           int64_t cap, uint32_t cell0, uint32_t ncells, T* __restrict__ agg, int64_t agg_stride) {
    constexpr int NC = 1 + 2 * D;
    const int K = sp.K, kb = sp.kb;
    const int nfr = D == 2 ? K : K * K;
    const uint32_t g = (blockIdx.x * blockDim.x + threadIdx.x) >> 3;
    const int gl = threadIdx.x & 7;
    const bool live = g < ncells * (uint32_t)nfr;
    const uint32_t c = cell0 + (live ? g / (uint32_t)nfr : 0);          // cells [cell0, cell0 + ncells)
    const int part = live ? (int)(g - (c - cell0) * (uint32_t)nfr) : 0;
    int cc[D];
    unlin_cell<D>(sp.dims, c, cc);
    const FineRows fr = fine_rows<D>(sp.dims, kb, cc[1], D == 3 ? cc[2] : 0);
    const uint32_t b = cs[(size_t)fr.base[part] + ((uint32_t)cc[0] << kb)];
    const uint32_t e = live ? cs[(size_t)fr.base[part] + ((uint32_t)(cc[0] + 1) << kb)] : b;
    cc[D - 1] = local_to_global_last<T, D>(sp, cc[D - 1]);               // centres are global (slab window)
    T acc[2 * D];
#pragma unroll
    for (int d = 0; d < 2 * D; ++d) acc[d] = T(0);
    for (uint32_t t = b + gl; t < e; t += 8) {
#pragma unroll
        for (int d = 0; d < D; ++d) {
            const T centre = sp.spacing * ((T)(cc[d] + 1) - T(0.5));
            acc[d] += pos[(size_t)d * cap + t] - centre;
            acc[D + d] += vel[(size_t)d * cap + t];
        }
    }
#pragma unroll
    for (int d = 0; d < 2 * D; ++d) {
        acc[d] += __shfl_xor_sync(0xffffffffu, acc[d], 4);
        acc[d] += __shfl_xor_sync(0xffffffffu, acc[d], 2);
        acc[d] += __shfl_xor_sync(0xffffffffu, acc[d], 1);
    }
    if (live && gl == 0) {
        T* dst = agg + (size_t)part * NC * agg_stride + c;
        dst[0] = (T)(e - b);
#pragma unroll
        for (int d = 0; d < 2 * D; ++d) dst[(size_t)(1 + d) * agg_stride] = acc[d];
    }
}

// ---------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------
bool ab_resort_applicable(const ab_handle* h) {
    if (h->desc.kind != AB_SPACE_CONTINUOUS) return false;
    for (int d = 0; d < h->D; ++d) if (h->dims[d] < 3) return false;     // 3 distinct neighbour rows / cells
    return h->n > 0;
}

static int ensure_agg(ab_handle* h) {
    // one partial (count, Σ(pos - centre), Σ vel) per fine row of every reference cell: 4 = max K^(D-1)
    size_t need = (size_t)4 * (1 + 2 * h->D) * h->ncells * h->tsize;
    if (h->cell_agg_bytes >= need) return AB_OK;
    if (h->cell_agg) { AB_CUDA(h, cudaStreamSynchronize(h->stream)); AB_CUDA(h, cudaFree(h->cell_agg)); h->cell_agg = nullptr; }
    AB_CUDA(h, cudaMalloc(&h->cell_agg, need));
    h->cell_agg_bytes = need;
    h->agg_valid = false;
    return AB_OK;
}

template <class T, int D>
static int resort_t(ab_handle* h, const void* src_pos, const void* src_vel, bool want_agg) {
    const int cur = h->cur, nxt = cur ^ 1;
    SpaceDev<T, D> sp = make_space<T, D>(h);
    ResortArgs a;
    // tile width (fine cells of one fine row): ~RS_TARGET agents per tile on average, at most
    // RS_FC_MAX fine cells, a multiple of K so that tiles hold whole reference cells
    const int fd0 = h->dims[0] << h->kb;
    const double dens = (double)h->n / (double)h->nfine;
    int tw = (int)((double)h->rs_target / (dens > 1e-9 ? dens : 1e-9));
    if (tw > RS_FC_MAX) tw = RS_FC_MAX;
    if (tw < 8) tw = 8;
    if (tw > fd0) tw = fd0;
    a.tpr = (fd0 + tw - 1) / tw;
    a.tw = (fd0 + a.tpr - 1) / a.tpr;                 // even split of the row
    a.tw = (a.tw + h->K - 1) / h->K * h->K;
    a.tpr = (fd0 + a.tw - 1) / a.tw;
    a.nrows = (int)(h->nfine / fd0);
    // slab window: only the owned fine rows (they come first in key order) are re-sorted here; the halo
    // rows are replaced wholesale by the neighbours' boundary layers (ab_comm.cu)
    if (h->win_W) a.nrows = (int)((int64_t)a.nrows * h->win_nl / h->dims[D - 1]);
    a.ntiles = a.nrows * a.tpr;
    a.has_vel = h->desc.kind == AB_SPACE_CONTINUOUS;
    a.want_agg = want_agg ? 1 : 0;
    // the step kernel has counted the new keys into cell_count: scan them into the new cell_start
    AB_TRY(ab_run_scan(h, h->cell_start_alt));
    if (want_agg) AB_TRY(ensure_agg(h));
    const bool extra = h->fields.size() > 2u;
    {
        AB_LAUNCH(h, "k_resort");
        k_resort<T, D><<<a.ntiles, RS_THREADS, 0, h->stream>>>(
            sp, a, h->cell_start, h->newkey, h->id[cur], (const T*)src_pos, (const T*)src_vel, h->cap, (int)h->n,
            (T*)h->fields[0].buf[nxt], (T*)h->fields[1].buf[nxt], h->id[nxt], h->key[nxt], h->cell_start_alt,
            extra ? h->perm : nullptr, (T*)h->cell_agg, (int64_t)h->ncells, h->idtmp, (int32_t*)h->rank, h->d_err);
    }
    std::swap(h->cell_start, h->cell_start_alt);
    AB_TRY(ab_permute_extra_fields(h));
    h->cur = nxt;
    h->slot_map_valid = false;
    h->index_valid = true;
    h->agg_valid = want_agg;
    if (want_agg) h->agg_epoch++;
    h->pairs_n = -1;
    return AB_OK;
}

// Re-sorts from (src_pos, src_vel, h->newkey) — all in the CURRENT slot order —
// into the other buffer.  Requires the new keys counted into h->cell_count and
// that no agent moved further than one fine cell (checked on the device:
// AB_EF_OVERFLOW otherwise).
int ab_resort_tiled(ab_handle* h, const void* src_pos, const void* src_vel, bool want_agg) {
    if (h->dtype == AB_F64) return h->D == 2 ? resort_t<double, 2>(h, src_pos, src_vel, want_agg) : resort_t<double, 3>(h, src_pos, src_vel, want_agg);
    return h->D == 2 ? resort_t<float, 2>(h, src_pos, src_vel, want_agg) : resort_t<float, 3>(h, src_pos, src_vel, want_agg);
}

template <class T, int D>
static int agg_t(ab_handle* h, int64_t cell0, int64_t ncell) {
    if (ncell <= 0) return AB_OK;
    AB_LAUNCH(h, "k_cell_agg");
    const int nfr = h->D == 2 ? h->K : h->K * h->K;
    k_cell_agg<T, D><<<ab_blocks(ncell * nfr * 8, 256), 256, 0, h->stream>>>(
        make_space<T, D>(h), h->cell_start, (const T*)h->fields[0].buf[h->cur], (const T*)h->fields[1].buf[h->cur], h->cap,
        (uint32_t)cell0, (uint32_t)ncell, (T*)h->cell_agg, (int64_t)h->ncells);
    return AB_OK;
}

// aggregates of the reference cells [cell0, cell0 + ncell) of the current sorted state
int ab_agg_range(ab_handle* h, int64_t cell0, int64_t ncell) {
    AB_TRY(ensure_agg(h));
    h->agg_epoch++;
    if (h->dtype == AB_F64) return h->D == 2 ? agg_t<double, 2>(h, cell0, ncell) : agg_t<double, 3>(h, cell0, ncell);
    return h->D == 2 ? agg_t<float, 2>(h, cell0, ncell) : agg_t<float, 3>(h, cell0, ncell);
}

int ab_ensure_agg(ab_handle* h) {
    AB_TRY(ab_ensure_index(h));
    AB_TRY(ensure_agg(h));
    if (h->agg_valid) return AB_OK;
    AB_TRY(ab_agg_range(h, 0, h->ncells));
    h->agg_valid = true;
    return AB_OK;
}
