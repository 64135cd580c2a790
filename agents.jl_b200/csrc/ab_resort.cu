// ab_resort.cu — tile-based, deterministic, atomic-free re-sort of the population
// after a fused step, and the per-reference-cell aggregates of the sorted state.
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
//           tiles are independent of each other.
//   WARP  = one tile.  It streams the OLD slots of the 3^D neighbouring fine
//           rows of its tile (agents move less than one fine cell per step on
//           this path) as ONE concatenated sequence — rows in (dz, dy) order,
//           slots ascending — keeps those whose new key falls into the tile and
//           gives each its final slot on the spot:
//               slot = new cell_start[key] + (agents of that cell seen earlier
//                      in the sequence),
//           a STABLE rank computed with match.any inside the warp and a
//           per-cell cursor in shared memory.  No atomics, no staging, no
//           sorting pass: source reads are coalesced, each agent's row is read
//           once and written once.
//   order   inside a cell = order of the concatenated source sequence.  By
//           induction over steps (the index build orders cells by id) this is
//           a pure function of the trajectory: runs are bitwise reproducible
//           and independent of how the space is cut into slabs (the source
//           rows are enumerated geometrically, never by slot address).
//   then    (optionally) count, Σ(pos - centre), Σ vel of the tile's part of
//           every reference cell it touches (one partial per fine row of a
//           reference cell, combined in fixed order by k_cell_M).
//
// (History, profiles/README.md: a CTA-per-tile version bucketed the entries in
//  shared memory with atomics and ranked every cell by id with an O(n²) loop —
//  30 warp instructions per agent, 71 us per 1 M agents; this one needs ~6.)
//
// Nothing here is specific to the flock rule: any fused rule that bounds the
// per-step displacement by one fine cell can use it.
#include <algorithm>
#include <cstdlib>

#include "ab_internal.cuh"

#define RS_WARPS 4
#define RS_THREADS (RS_WARPS * 32)
#define RS_MAXSEG 27            // 9 source rows x (main run + 2 periodic wrap cells)
#define RS_TW_MAX 256           // fine cells per tile (cursor array per warp)
#define RS_RING 64              // (source slot, output slot) pairs waiting to be moved, per warp

struct ResortArgs {
    int tw, tpr, ntiles, nrows;
    int has_vel, want_agg;
    int scanfree;                    // tile bases from the per-row / per-tile counters instead of a scanned array
    const uint32_t* tile_cnt;        // this step's counters (scan-free path)
    const uint32_t* row_cnt;
    uint32_t* tile_other;            // the other parity: consumed by the previous re-sort, zeroed here
    uint32_t* row_other;
};

#ifdef RS_MINB
#define RS_BOUNDS __launch_bounds__(RS_THREADS, RS_MINB)
#else
#define RS_BOUNDS __launch_bounds__(RS_THREADS)
#endif
#ifndef RS_G2D
#define RS_G2D 4
#endif
#ifndef RS_AGG_UNROLL
#define RS_AGG_UNROLL 4
#endif
constexpr int kRsAggUnroll = RS_AGG_UNROLL;     // crowded-cell aggregate loop: loads of several iterations in flight
// Phases of a tile (one warp):
//   0  output range from the scanned tile totals; the tile's cell counts -> cursors (warp scan) -> new cell_start;
//      the counts are zeroed behind the read (the histogram is ready for the next step)
//   1  stream the keys of the 3^D source rows (G chunks of 32 in flight), rank the agents that fall into the tile
//      (match.any + per-cell cursor: final output slot at once) and queue (source slot, output slot) in a ring
//   2  whenever 32 pairs wait: move 32 agents with all lanes busy (gathered loads, stores clustered by cell)
//   3  aggregates of the tile's cells
//   HALO (slab window, no user columns): tiles of the W bottom / top owned layers also deliver what they place into
//      the boundary messages of the lower / upper neighbour (HaloOut) — there is no separate pack launch; those
//      tiles run first, and the last of them publishes the exchange number in the neighbours' flags
template <class T, int D, int G, bool HALO>
__global__ void RS_BOUNDS
k_resort(SpaceDev<T, D> sp, ResortArgs a, const __grid_constant__ HaloOut ho, const uint32_t* __restrict__ cs_old, const uint32_t* __restrict__ newkey,
         const int32_t* __restrict__ id_old, const T* __restrict__ src_pos, const T* __restrict__ src_vel,
         int64_t cap, T* dst_pos, T* dst_vel, int32_t* __restrict__ id_new,
         uint32_t* __restrict__ key_new, uint32_t* __restrict__ cell_count, const uint32_t* __restrict__ tile_base,
         uint32_t* __restrict__ cs_new, int32_t* __restrict__ perm,
         T* __restrict__ agg, int64_t agg_stride, int* __restrict__ err) {
    __shared__ uint32_t s_beg[RS_WARPS][RS_MAXSEG];
    __shared__ uint32_t s_pre[RS_WARPS][RS_MAXSEG + 1];
    __shared__ uint32_t s_cur[RS_WARPS][RS_TW_MAX];
    __shared__ uint2 s_ring[RS_WARPS][RS_RING];

    ab_dep_wait();
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const unsigned lt = (1u << lane) - 1u;
    int tile = blockIdx.x * RS_WARPS + w;
    if (tile >= a.ntiles) return;                                  // warp-uniform; no CTA barriers below
    const int kb = sp.kb;
    const int fd0 = sp.dims[0] << kb, fd1 = sp.dims[1] << kb, fd2 = D == 3 ? (sp.dims[2] << kb) : 1;
    int frow = tile / a.tpr;
    const int seg_x = tile - frow * a.tpr;                         // fine row, segment of it
    // HALO: nb fine rows per message; launch order = top boundary rows, bottom boundary rows, interior
    const int nb = HALO ? (int)(ho.ncs / (uint32_t)fd0) : 0;
    bool reg[2] = {false, false};                                  // this tile feeds the bottom / top message
    uint32_t s0[2] = {0, 0}, cntc[2] = {0, 0}, k0[2] = {0, 0};
    if (HALO) {
        frow = frow < nb ? a.nrows - nb + frow : frow - nb;
        tile = frow * a.tpr + seg_x;
        reg[0] = frow < nb && ho.buf[0] != nullptr;
        reg[1] = frow >= a.nrows - nb && ho.buf[1] != nullptr;
        const int t0b[2] = {0, (a.nrows - nb) * a.tpr}, t1b[2] = {nb * a.tpr, a.ntiles};
        k0[1] = ho.kown - ho.ncs;
#pragma unroll
        for (int dir = 0; dir < 2; ++dir) {
            if (!reg[dir]) continue;
            uint32_t cnt;
            if (a.scanfree) {
                uint32_t part0 = 0, part1 = 0;                      // rows before the region / rows of the region
                const int ra = dir == 0 ? 0 : a.nrows - nb, rb = dir == 0 ? nb : a.nrows;
                for (int r = lane; r < rb; r += 32) { const uint32_t v = a.row_cnt[r]; if (r < ra) part0 += v; else part1 += v; }
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) { part0 += __shfl_xor_sync(FULL, part0, o); part1 += __shfl_xor_sync(FULL, part1, o); }
                s0[dir] = part0;
                cnt = part1;
            } else {
                s0[dir] = tile_base[t0b[dir]];
                cnt = tile_base[t1b[dir]] - s0[dir];
            }
            cntc[dir] = min(cnt, (uint32_t)ho.m.C);
            if (tile == t0b[dir] && lane == 0) {                   // the region's first tile: header (and the capacity check)
                uint32_t* hdr = reinterpret_cast<uint32_t*>(ho.buf[dir]);
                hdr[0] = cntc[dir]; hdr[1] = k0[dir]; hdr[2] = ho.ncs; hdr[3] = 0xAB200u;
                if (cnt > (uint32_t)ho.m.C) atomicOr(err, AB_EF_COMM | AB_EF_COMM_CAP);
            }
            if (tile == t1b[dir] - 1 && lane == 0)                 // the region's last tile: closing entry of the cell_start slice
                reinterpret_cast<uint32_t*>(ho.buf[dir] + ho.m.off_cs)[ho.ncs] = cntc[dir];
        }
    }
    const int x0 = seg_x * a.tw;
    const int x1 = min(fd0, x0 + a.tw);          // exclusive, fine cells
    const uint32_t kbeg = (uint32_t)frow * (uint32_t)fd0 + (uint32_t)x0;
    const int fc = x1 - x0;
    const uint32_t kend = kbeg + (uint32_t)fc;
    const int fy0 = D == 2 ? frow : frow % fd1;
    const int fz0 = D == 2 ? 0 : frow / fd1;
    // ---- source segments: lane i describes segment i = 3*row + {main, left wrap, right wrap} ----
    // (loaded first: these two reads depend on nothing, and the key stream below depends on them)
    constexpr int NROW = D == 2 ? 3 : 9;
    constexpr int NSEG = 3 * NROW;
    uint32_t sb = 0, sl = 0;
    if (lane < NSEG) {
        const int r = lane / 3, j = lane - 3 * r;
        const int dy = r % 3 - 1, dz = D == 3 ? r / 3 - 1 : 0;
        int y = fy0 + dy, z = fz0 + dz;
        bool ok = true;
        if (y < 0 || y >= fd1) { if (!sp.periodic[1]) ok = false; y = y < 0 ? y + fd1 : y - fd1; }
        if (D == 3 && (z < 0 || z >= fd2)) { if (!sp.periodic[2]) ok = false; z = z < 0 ? z + fd2 : z - fd2; }
        const int lo = max(x0 - 1, 0), hi = min(x1, fd0 - 1);      // inclusive fine cells of the main run
        int c0 = lo, c1 = hi + 1;                                  // [c0, c1) cells of this segment
        if (j == 1) { ok = ok && sp.periodic[0] && x0 == 0 && hi < fd0 - 1; c0 = fd0 - 1; c1 = fd0; }
        if (j == 2) { ok = ok && sp.periodic[0] && x1 == fd0 && lo > 0; c0 = 0; c1 = 1; }
        if (ok) {
            const size_t rb = (size_t)(D == 3 ? (uint32_t)z * (uint32_t)fd1 + (uint32_t)y : (uint32_t)y) * (uint32_t)fd0;
            sb = cs_old[rb + c0];
            sl = cs_old[rb + c1] - sb;
        }
    }
    // sum of the row counters [r0, r1) over the warp (scan-free path)
    auto rows_sum = [&](int r0, int r1) -> uint32_t {
        uint32_t part = 0;
        for (int r = r0 + lane; r < r1; r += 32) part += a.row_cnt[r];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(FULL, part, o);
        return part;
    };
    uint32_t base, total;
    if (a.scanfree) {
        // agents in the fine rows before mine + in the tiles before me in my row
        uint32_t part = 0;
        for (int t = lane; t < seg_x; t += 32) part += a.tile_cnt[frow * a.tpr + t];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(FULL, part, o);
        base = rows_sum(0, frow) + part;
        total = a.tile_cnt[tile];
        if (lane == 0) { a.tile_other[tile] = 0; if (seg_x == 0) a.row_other[frow] = 0; }
    } else {
        base = tile_base[tile];
        total = tile_base[tile + 1] - base;
    }
    if (tile == a.ntiles - 1 && lane == 0) cs_new[kend] = base + total;      // end of the re-sorted key range
    bool lost = false;

    // HALO: entry of the message's cell_start slice for the tile's cell i, whose first output slot is `slot`
    auto msg_cs = [&](int i, uint32_t slot) {
#pragma unroll
        for (int dir = 0; dir < 2; ++dir)
            if (reg[dir]) reinterpret_cast<uint32_t*>(ho.buf[dir] + ho.m.off_cs)[kbeg + (uint32_t)i - k0[dir]] = min(slot - s0[dir], cntc[dir]);
    };
    if (total == 0) {
        for (int i = lane; i < fc; i += 32) {
            cs_new[(size_t)kbeg + i] = base;
            if (HALO) msg_cs(i, base);
        }
    } else {
    // ---- phase 0: lane l owns the cells [l * ch, (l + 1) * ch) of the tile ----
    {
        const int ch = (fc + 31) >> 5;                             // <= 8
        const int i0 = lane * ch;
        uint32_t cnt[RS_TW_MAX / 32];
        uint32_t sum = 0;
#pragma unroll
        for (int q = 0; q < RS_TW_MAX / 32; ++q) {
            cnt[q] = 0;
            if (q < ch && i0 + q < fc) cnt[q] = cell_count[(size_t)kbeg + i0 + q];
        }
#pragma unroll
        for (int q = 0; q < RS_TW_MAX / 32; ++q) sum += cnt[q];
        uint32_t inc = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(FULL, inc, o); if (lane >= o) inc += t; }
        uint32_t run = inc - sum;                                  // agents of the tile in the cells before mine
#pragma unroll
        for (int q = 0; q < RS_TW_MAX / 32; ++q) {
            if (q < ch && i0 + q < fc) {
                s_cur[w][i0 + q] = run;
                cs_new[(size_t)kbeg + i0 + q] = base + run;
                if (HALO) msg_cs(i0 + q, base + run);
                if (cnt[q]) cell_count[(size_t)kbeg + i0 + q] = 0;
                run += cnt[q];
            }
        }
    }
    uint32_t inc = sl;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(FULL, inc, o); if (lane >= o) inc += t; }
    const uint32_t L = __shfl_sync(FULL, inc, 31);                 // length of the concatenated source sequence
    if (lane < NSEG) { s_beg[w][lane] = sb; s_pre[w][lane] = inc - sl; }
    if (lane == 0) s_pre[w][NSEG] = L;
    __syncwarp();

    // moves the agents of ring entries [d0, d0 + nv): all lanes busy
    auto flush = [&](uint32_t d0, uint32_t nv) {
        if ((uint32_t)lane < nv) {
            const uint2 e = s_ring[w][(d0 + lane) & (RS_RING - 1)];
            const size_t si = e.x, oo = e.y;
            T pv[2 * D];
#pragma unroll
            for (int d = 0; d < D; ++d) pv[d] = src_pos[(size_t)d * cap + si];
            if (a.has_vel) {
#pragma unroll
                for (int d = 0; d < D; ++d) pv[D + d] = src_vel[(size_t)d * cap + si];
            }
            const int32_t idv = id_old[si];
            const uint32_t kv = newkey[si];
#pragma unroll
            for (int d = 0; d < D; ++d) dst_pos[(size_t)d * cap + oo] = pv[d];
            if (a.has_vel) {
#pragma unroll
                for (int d = 0; d < D; ++d) dst_vel[(size_t)d * cap + oo] = pv[D + d];
            }
            id_new[oo] = idv;
            key_new[oo] = kv;
            if (perm) perm[oo] = (int32_t)si;
            if (HALO) {
#pragma unroll
                for (int dir = 0; dir < 2; ++dir) {
                    const uint32_t t = (uint32_t)oo - s0[dir];
                    if (reg[dir] && t < cntc[dir]) {
                        char* mb = ho.buf[dir];
                        T* mpos = reinterpret_cast<T*>(mb + ho.m.off_pos);
                        T* mvel = reinterpret_cast<T*>(mb + ho.m.off_vel);
#pragma unroll
                        for (int d = 0; d < D; ++d) { mpos[(size_t)d * ho.m.C + t] = pv[d]; mvel[(size_t)d * ho.m.C + t] = pv[D + d]; }
                        reinterpret_cast<int32_t*>(mb + ho.m.off_id)[t] = idv;
                        reinterpret_cast<uint32_t*>(mb + ho.m.off_key)[t] = kv;
                    }
                }
            }
        }
        __syncwarp();                                              // ring entries may be overwritten from here on
    };

    uint32_t queued = 0, moved = 0;                                // ring counters (warp-uniform)
    int sg = 0;                                                    // this lane's current segment (monotone)
    for (uint32_t v0 = 0; v0 < L; v0 += 32u * G) {
        uint32_t k[G], s[G];
        // keys of G chunks (independent coalesced loads)
#pragma unroll
        for (int g = 0; g < G; ++g) {
            const uint32_t v = v0 + 32u * g + lane;
            k[g] = 0xffffffffu;
            s[g] = 0;
            if (v < L) {
                while (v >= s_pre[w][sg + 1]) ++sg;
                s[g] = s_beg[w][sg] + (v - s_pre[w][sg]);
                k[g] = newkey[s[g]];
            }
        }
        // stable rank inside the warp + per-cell cursor -> final slot; queue the pair
#pragma unroll
        for (int g = 0; g < G; ++g) {
            const bool in = k[g] >= kbeg && k[g] < kend;
            const unsigned bal = __ballot_sync(FULL, in);
            if (bal == 0) continue;                                // warp-uniform
            const uint32_t lc = in ? k[g] - kbeg : 0xffffffffu;
            const unsigned m = __match_any_sync(FULL, lc);
            if (in) {
                const uint32_t o = base + s_cur[w][lc] + __popc(m & lt);
                s_ring[w][(queued + __popc(bal & lt)) & (RS_RING - 1)] = make_uint2(s[g], o);
            }
            __syncwarp();
            if (in && (m >> lane) == 1u) s_cur[w][lc] += __popc(m);      // highest lane of the group advances the cursor
            queued += __popc(bal);
            __syncwarp();
            if (queued - moved >= 32u) { flush(moved, 32u); moved += 32u; }
        }
    }
    if (queued != moved) flush(moved, queued - moved);
    // every agent counted into this tile must have been found in the 3^D neighbourhood
    if (queued != total) {
        if (lane == 0) atomicOr(err, AB_EF_OVERFLOW);              // an agent left its 3^D neighbourhood
        lost = true;
    }
    }   // total != 0

    if (HALO && ho.done_ctr && (frow < nb || frow >= a.nrows - nb)) {
        // direct exchange: this boundary tile's part of the messages is stored (peer memory); the LAST boundary tile
        // to get here publishes the exchange number in the neighbours' flags.  __syncwarp orders the lanes' stores
        // before lane 0's system-scope fence (fences are cumulative).
        __syncwarp();
        if (lane == 0) {
            __threadfence_system();
            const uint32_t nbt = (uint32_t)(min(2 * nb, a.nrows) * a.tpr);
            if (atomicAdd(ho.done_ctr, 1u) == nbt - 1u) {
                __threadfence_system();
                *ho.done_ctr = 0;
                if (ho.flag[0]) asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(ho.flag[0]), "r"(ho.epoch) : "memory");
                if (ho.flag[1]) asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(ho.flag[1]), "r"(ho.epoch) : "memory");
            }
        }
    }
    if (lost) return;

    // ---- per reference cell: count, Σ(pos - centre), Σ vel of this fine row's part, canonical order ---
    // Canonical sum of a cell's n agents (shared with k_cell_agg): 8 strided partials a_g = Σ_i x[g + 8 i]
    // combined as ((a0+a4)+(a2+a6)) + ((a1+a5)+(a3+a7)).  Cells with n <= 8 (nearly all of them in a dilute
    // state) are summed by ONE lane — each partial is a single term — cells with more by an 8-lane group.
    if (a.want_agg) {
        __syncwarp();        // the tile's output slots were written by this warp; read them back through L2
        constexpr int NC = 1 + 2 * D;
        // component c (pos 0..D-1, vel D..2D-1) of output slot t
        auto row = [&](int c, uint32_t t) -> T {
            return c < D ? __ldcg(dst_pos + (size_t)c * cap + t) : __ldcg(dst_vel + (size_t)(c - D) * cap + t);
        };
        const int K = sp.K;
        const int ncell = fc >> kb;                           // x0 and tw are multiples of K
        const int cy = fy0 >> kb, cz = fz0 >> kb;
        const int part = (fy0 & (K - 1)) + K * (fz0 & (K - 1));
        const uint32_t refrow = (D == 3 ? ((uint32_t)cz * sp.dims[1] + cy) : (uint32_t)cy) * (uint32_t)sp.dims[0];
        T* aggp = agg + (size_t)part * NC * agg_stride + refrow;
        T ctr[D];                                             // centre of the row's cells in the other dimensions
        {
            int cc[3] = {0, cy, cz};
            cc[D - 1] = local_to_global_last<T, D>(sp, cc[D - 1]);     // centres are global (slab window)
#pragma unroll
            for (int d = 1; d < D; ++d) ctr[d] = sp.spacing * ((T)(cc[d] + 1) - T(0.5));
        }
        for (int c0 = 0; c0 < ncell; c0 += 32) {
            const int c = c0 + lane;
            const bool live = c < ncell;
            uint32_t b = 0, e = 0;
            if (live) {
                // after the placement the cursor of a fine cell is its end (= the next cell's start)
                const int f0 = c << kb, f1 = (c + 1) << kb;
                b = base + (total && f0 ? s_cur[w][f0 - 1] : 0u);
                e = base + (total ? s_cur[w][f1 - 1] : 0u);
            }
            const uint32_t n = e - b;
            const int cx = (x0 >> kb) + c;
            if (live && n <= 8u) {
                ctr[0] = sp.spacing * ((T)(cx + 1) - T(0.5));
                // pair(i) = x[i] + x[i + 4] with x[i] = 0 + term i (0 beyond n): the group path's partials
                auto pair = [&](int i, T* out) {
                    T u[2 * D], v[2 * D];
#pragma unroll
                    for (int d = 0; d < 2 * D; ++d) u[d] = v[d] = T(0);
                    if ((uint32_t)i < n) {
#pragma unroll
                        for (int d = 0; d < D; ++d) {
                            u[d] += row(d, b + i) - ctr[d];
                            u[D + d] += row(D + d, b + i);
                        }
                    }
                    if ((uint32_t)(i + 4) < n) {
#pragma unroll
                        for (int d = 0; d < D; ++d) {
                            v[d] += row(d, b + i + 4) - ctr[d];
                            v[D + d] += row(D + d, b + i + 4);
                        }
                    }
#pragma unroll
                    for (int d = 0; d < 2 * D; ++d) out[d] = u[d] + v[d];
                };
                T p0[2 * D], p1[2 * D], q[2 * D];
                pair(0, p0); pair(2, p1);
#pragma unroll
                for (int d = 0; d < 2 * D; ++d) q[d] = p0[d] + p1[d];
                pair(1, p0); pair(3, p1);
                aggp[cx] = (T)n;
#pragma unroll
                for (int d = 0; d < 2 * D; ++d) aggp[(size_t)(1 + d) * agg_stride + cx] = q[d] + (p0[d] + p1[d]);
            }
            // crowded cells: 8 lanes per cell, 4 cells at a time
            unsigned big = __ballot_sync(FULL, live && n > 8u);
            const int grp = lane >> 3, gl = lane & 7;
            while (big) {
                unsigned rest = big;
                int src = -1;
#pragma unroll
                for (int g = 0; g < 4; ++g) {
                    const int l = rest ? __ffs(rest) - 1 : -1;
                    if (l >= 0) rest &= rest - 1;
                    if (g == grp) src = l;
                }
                big = rest;
                const uint32_t gb = __shfl_sync(FULL, b, src < 0 ? 0 : src), ge = __shfl_sync(FULL, e, src < 0 ? 0 : src);
                const int gcx = (x0 >> kb) + c0 + src;
                ctr[0] = sp.spacing * ((T)(gcx + 1) - T(0.5));
                T acc[2 * D];
#pragma unroll
                for (int d = 0; d < 2 * D; ++d) acc[d] = T(0);
                if (src >= 0) {
#pragma unroll kRsAggUnroll
                    for (uint32_t t = gb + gl; t < ge; t += 8) {
#pragma unroll
                        for (int d = 0; d < D; ++d) {
                            acc[d] += row(d, t) - ctr[d];
                            acc[D + d] += row(D + d, t);
                        }
                    }
                }
#pragma unroll
                for (int d = 0; d < 2 * D; ++d) {
                    acc[d] += __shfl_xor_sync(FULL, acc[d], 4);
                    acc[d] += __shfl_xor_sync(FULL, acc[d], 2);
                    acc[d] += __shfl_xor_sync(FULL, acc[d], 1);
                }
                if (src >= 0 && gl == 0) {
                    aggp[gcx] = (T)(ge - gb);
#pragma unroll
                    for (int d = 0; d < 2 * D; ++d) aggp[(size_t)(1 + d) * agg_stride + gcx] = acc[d];
                }
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

// Tile geometry of the next tiled re-sort: ~rs_target agents per tile on average, at most RS_TW_MAX fine cells, a
// multiple of K so that tiles hold whole reference cells.  Fixed before the step kernel runs (it counts per tile).
int ab_resort_plan(ab_handle* h) {
    const int D = h->D;
    const int fd0 = h->dims[0] << h->kb;
    if (fd0 > 65536) return ab_fail(h, AB_ERR_INVALID, "more than 65536 fine cells per row");
    const double dens = (double)h->n / (double)h->nfine;
    int tw = (int)((double)h->rs_target / (dens > 1e-9 ? dens : 1e-9));
    if (tw > RS_TW_MAX) tw = RS_TW_MAX;
    if (tw < 4) tw = 4;
    if (tw > fd0) tw = fd0;
    int tpr = (fd0 + tw - 1) / tw;
    tw = (fd0 + tpr - 1) / tpr;                       // even split of the row
    tw = (tw + h->K - 1) / h->K * h->K;
    if (tw > RS_TW_MAX) tw = RS_TW_MAX / h->K * h->K;
    tpr = (fd0 + tw - 1) / tw;
    int64_t nrows = h->nfine / fd0;
    // slab window: only the owned fine rows (they come first in key order) are re-sorted here; the halo
    // rows are replaced wholesale by the neighbours' boundary layers (ab_comm.cu)
    if (h->win_W) nrows = nrows * h->win_nl / h->dims[D - 1];
    const int64_t ntiles = nrows * tpr;
    if (ntiles > 0x7fffffffLL) return ab_fail(h, AB_ERR_INVALID, "too many re-sort tiles");
    const size_t need = ((size_t)ntiles + SCAN_TILE - 1) / SCAN_TILE * SCAN_TILE;
    const size_t need_rows = ((size_t)nrows + 63) / 64 * 64;
    // few fine rows (2-D spaces): no scan launch — the re-sort tiles sum the per-row and per-tile counters themselves
    bool scanfree = nrows <= 4096 && tpr <= 4096;
    if (const char* e = getenv("AB_SCANFREE")) scanfree = scanfree && atoi(e) != 0;
    const bool changed = tw != h->rs_tw || tpr != h->rs_tpr || (int)nrows != h->rs_nrows || scanfree != h->rs_scanfree;
    if (need > h->tile_cap || need_rows > h->row_cap) {
        AB_CUDA(h, cudaStreamSynchronize(h->stream));
        if (h->tile_count) AB_CUDA(h, cudaFree(h->tile_count));
        if (h->tile_base) AB_CUDA(h, cudaFree(h->tile_base));
        if (h->row_count) AB_CUDA(h, cudaFree(h->row_count));
        h->tile_count = h->tile_base = h->row_count = nullptr;
        h->tile_cap = std::max(need, h->tile_cap);
        h->row_cap = std::max(need_rows, h->row_cap);
        AB_CUDA(h, cudaMalloc((void**)&h->tile_count, 2 * h->tile_cap * sizeof(uint32_t)));
        AB_CUDA(h, cudaMalloc((void**)&h->tile_base, (h->tile_cap + 1) * sizeof(uint32_t)));
        AB_CUDA(h, cudaMalloc((void**)&h->row_count, 2 * h->row_cap * sizeof(uint32_t)));
        if (h->tile_cap / SCAN_TILE + 1 > h->scan_status_len) return ab_fail(h, AB_ERR_INVALID, "tile scan larger than the handle's status buffer");
    }
    if (changed || !h->rs_ntiles) {
        // a new geometry: both halves of the counters start from zero (stream order: after everything queued so far)
        AB_CUDA(h, cudaMemsetAsync(h->tile_count, 0, 2 * h->tile_cap * sizeof(uint32_t), h->stream));
        AB_CUDA(h, cudaMemsetAsync(h->row_count, 0, 2 * h->row_cap * sizeof(uint32_t), h->stream));
        h->rs_par = 0;
    }
    h->rs_scanfree = scanfree;
    h->rs_tw = tw; h->rs_tpr = tpr; h->rs_nrows = (int)nrows; h->rs_ntiles = (int)ntiles;
    h->rs_magic = (uint32_t)((0x100000000ULL + (uint64_t)tw - 1) / (uint64_t)tw);
    return AB_OK;
}

template <class T, int D>
static int resort_t(ab_handle* h, const void* src_pos, const void* src_vel, bool want_agg, cudaEvent_t after_scan) {
    const int cur = h->cur, nxt = cur ^ 1;
    SpaceDev<T, D> sp = make_space<T, D>(h);
    ResortArgs a;
    a.tw = h->rs_tw; a.tpr = h->rs_tpr; a.nrows = h->rs_nrows; a.ntiles = h->rs_ntiles;
    if (a.ntiles <= 0 || a.tw <= 0) return ab_fail(h, AB_ERR_STATE, "ab_resort_plan must precede the step kernel");
    a.has_vel = h->desc.kind == AB_SPACE_CONTINUOUS;
    a.want_agg = want_agg ? 1 : 0;
    // the step kernel has counted the new keys per fine cell, per tile (and per fine row): every tile finds its output
    // range from those (scan-free path) or from the scanned tile totals, and turns its own cell counts into its slice
    // of the new cell_start
    a.scanfree = h->rs_scanfree ? 1 : 0;
    a.tile_cnt = h->tile_count + (size_t)h->rs_par * h->tile_cap;
    a.row_cnt = h->row_count + (size_t)h->rs_par * h->row_cap;
    a.tile_other = h->tile_count + (size_t)(h->rs_par ^ 1) * h->tile_cap;
    a.row_other = h->row_count + (size_t)(h->rs_par ^ 1) * h->row_cap;
    if (!h->rs_scanfree) AB_TRY(ab_scan_u32(h, h->tile_count, h->tile_base, (uint32_t)a.ntiles));
    if (after_scan) AB_CUDA(h, cudaEventRecord(after_scan, h->stream));     // tile_base is complete: a side-stream halo install may start
    if (want_agg) AB_TRY(ensure_agg(h));
    const bool extra = h->fields.size() > 2u;
    // slab window: the boundary tiles deliver the neighbours' messages while they place their agents
    HaloOut ho;
    ho.enabled = 0;
    if (h->win_W && h->comm) AB_TRY(ab_comm_pack_begin(h, &ho));
    h->halo_packed = ho.enabled != 0;
    {
        AB_LAUNCH(h, "k_resort");
        constexpr int G = (D == 2) ? RS_G2D : 2;           // source chunks in flight per warp
        auto kern = ho.enabled ? k_resort<T, D, G, true> : k_resort<T, D, G, false>;
        AB_CUDA(h, ab_launch_chain(h->pdl, kern, dim3((a.ntiles + RS_WARPS - 1) / RS_WARPS), dim3(RS_THREADS), h->stream,
            sp, a, ho, h->cell_start, h->newkey, h->id[cur], (const T*)src_pos, (const T*)src_vel, h->cap,
            (T*)h->fields[0].buf[nxt], (T*)h->fields[1].buf[nxt], h->id[nxt], h->key[nxt], h->cell_count, h->tile_base,
            h->cell_start_alt, extra ? h->perm : nullptr, (T*)h->cell_agg, (int64_t)h->ncells, h->d_err));
    }
    if (h->rs_scanfree) h->rs_par ^= 1;                // the next step accumulates into the other half
    h->unpack_rows = h->rs_scanfree && h->win_W;
    h->unpack_tb = !h->rs_scanfree && h->win_W;
    std::swap(h->cell_start, h->cell_start_alt);
    AB_TRY(ab_permute_extra_fields(h, true));
    h->cur = nxt;
    h->slot_map_valid = false;
    h->index_valid = true;
    h->agg_valid = want_agg;
    if (want_agg) h->agg_epoch++;
    h->pairs_n = -1;
    return AB_OK;
}

// Re-sorts from (src_pos, src_vel, h->newkey) — all in the CURRENT slot order —
// into the other buffer.  Requires the new keys counted into h->cell_count / h->tile_count (count_new_key,
// geometry of ab_resort_plan) and that no agent moved further than one fine cell (checked on the device:
// AB_EF_OVERFLOW otherwise).
int ab_resort_tiled(ab_handle* h, const void* src_pos, const void* src_vel, bool want_agg, cudaEvent_t after_scan) {
    if (h->dtype == AB_F64) return h->D == 2 ? resort_t<double, 2>(h, src_pos, src_vel, want_agg, after_scan) : resort_t<double, 3>(h, src_pos, src_vel, want_agg, after_scan);
    return h->D == 2 ? resort_t<float, 2>(h, src_pos, src_vel, want_agg, after_scan) : resort_t<float, 3>(h, src_pos, src_vel, want_agg, after_scan);
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
