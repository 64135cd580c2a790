// ab_comm.cu — multi-GPU slab decomposition of a ContinuousSpace (one process per GPU).
//
// The reference is single-task (src/simulations/step_standard.jl:9-23); this is the
// scale-out of its hot path.  The space is cut into slabs of whole
// reference-cell LAYERS along the last axis — the slowest axis of the column-major cell
// key (src/spaces/grid_multi.jl:73), so a slab is one contiguous key range and the owned
// agents are one contiguous slot range of the cell-sorted state.
//
// A rank's handle stores a WINDOW of nl + 2W layers: its nl owned layers, then the W
// layers above (upper halo, owned by rank+1), then the W layers below (lower halo, owned
// by rank-1); local layer = (global layer - first owned layer) mod (nl + 2W), so every
// kernel's ordinary periodic index arithmetic finds the halo (ab_internal.cuh, SpaceDev).
// Positions stay global: pos2cell (continuous.jl:112-113) is evaluated on the same
// numbers as on one GPU, and cell centres / minimum images use global coordinates.
//
// One step on a rank (flock_steps_t, ab_flock.cu):
//   sweep     owned birds AND the first halo layer on each side (a few % redundant work):
//             the halo layer's birds that fly into the slab are this step's immigrants,
//             computed locally from the halo instead of being received
//   scan + re-sort of the owned rows (emigrants fall out: no owned tile collects them)
//   exchange  (this file) the W top and W bottom owned layers of the NEW sorted state replace the
//             neighbours' halo layers wholesale — halo refresh and migration in a single message.
//             Default (ab_comm_ipc_export / _import): k_halo_pack stores them straight into the
//             neighbours' memory over NVLink (CUDA IPC mapping) and publishes a flag; k_halo_unpack
//             waits for the two flags.  Baseline: one NCCL group of two sends + two receives of the
//             fixed-capacity messages.
// There is no size round trip through the host; the whole step stays asynchronous on the stream.
// Because agents keep their global ids and the in-cell order is canonical, an N-GPU run is bitwise
// identical to the 1-GPU run.
#include <dlfcn.h>

#include <algorithm>
#include <cstdlib>
#include <cstring>

#include "ab_internal.cuh"

#define AB_HALO_W 3   // radius-2 offset table (approximate search) of the recomputed first halo layer

// ---- NCCL, bound at run time (dlopen: the process may already hold torch's libnccl.so.2) ----------
typedef struct { char internal[128]; } nccl_uid_t;
typedef void* nccl_comm_t;
enum { NCCL_UINT8 = 1 };
struct nccl_api {
    void* lib = nullptr;
    int (*GetUniqueId)(nccl_uid_t*) = nullptr;
    int (*CommInitRank)(nccl_comm_t*, int, nccl_uid_t, int) = nullptr;
    int (*CommDestroy)(nccl_comm_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    int (*Send)(const void*, size_t, int, int, nccl_comm_t, cudaStream_t) = nullptr;
    int (*Recv)(void*, size_t, int, int, nccl_comm_t, cudaStream_t) = nullptr;
    int (*AllReduce)(const void*, void*, size_t, int, int, nccl_comm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
};
static nccl_api g_nccl;
static std::string g_nccl_err;

static bool nccl_load() {
    if (g_nccl.lib) return true;
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    void* lib = nullptr;
    for (const char* nm : names) {
        lib = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
        if (lib) break;
    }
    if (!lib) { g_nccl_err = std::string("cannot load NCCL: ") + dlerror(); return false; }
#define SYM(field, name)                                                    \
    *(void**)(&g_nccl.field) = dlsym(lib, name);                            \
    if (!g_nccl.field) { g_nccl_err = std::string("NCCL symbol missing: ") + name; return false; }
    SYM(GetUniqueId, "ncclGetUniqueId");
    SYM(CommInitRank, "ncclCommInitRank");
    SYM(CommDestroy, "ncclCommDestroy");
    SYM(GroupStart, "ncclGroupStart");
    SYM(GroupEnd, "ncclGroupEnd");
    SYM(Send, "ncclSend");
    SYM(Recv, "ncclRecv");
    SYM(AllReduce, "ncclAllReduce");
    SYM(GetErrorString, "ncclGetErrorString");
#undef SYM
    g_nccl.lib = lib;
    return true;
}

struct ab_comm {
    nccl_comm_t comm = nullptr;
    int rank = 0, nranks = 1, lower = 0, upper = 0;
    int64_t cap_agents = 0;          // message capacity C (agents)
    int64_t want_cap = 0;            // requested by ab_comm_set_capacity (0 = derive from the population)
    size_t msg_bytes = 0;
    int msg_K = 0;
    void* sendbuf[2] = {nullptr, nullptr};   // [0] bottom layers -> lower neighbour, [1] top layers -> upper neighbour
    void* recvbuf[2] = {nullptr, nullptr};   // [0] from the upper neighbour (its bottom layers), [1] from the lower one
    int64_t n_own = 0;
    // in-process peers (ab_comm_link_local): slabs of one process, on the same or on different devices
    ab_handle *peer_lower = nullptr, *peer_upper = nullptr;
    bool linked_local = false;       // ab_comm_link_local done (an end slab of a non-periodic axis has one peer only)
    bool has_lower = true, has_upper = true;   // false on the end ranks of a non-periodic last axis
    bool no_fused_pack = false;      // AB_FUSED_PACK=0: always pack with the separate k_halo_pack launch (A/B experiments)
    size_t fused_pack_max_bytes = (size_t)8 << 20;   // AB_FUSED_PACK_MAXMB: largest message the re-sort delivers itself
    cudaEvent_t ev_packed = nullptr, ev_consumed = nullptr;
    uint32_t* h_counts = nullptr;    // pinned {n_own, n_window}
    int64_t exchanges = 0;
    // direct exchange over peer memory (ab_comm_ipc_export / _import): every rank owns ONE block of
    // [2 parities][2 sides] message slots + 4 flags; the neighbours' blocks are mapped through CUDA IPC and
    // k_halo_pack stores straight into them over NVLink (only the agents present travel, no NCCL call on
    // the data path); a flag per slot carries the exchange number.
    char* ipc_block = nullptr;
    size_t ipc_stride = 0;           // bytes per message slot (layout at the finest refinement)
    int64_t ipc_cap = 0;             // capacity the block was sized for
    char *ipc_lower = nullptr, *ipc_upper = nullptr;     // the neighbours' blocks
    bool ipc_on = false;
    int cur_par = 0;                 // parity of the exchange in flight (pack -> transfer -> unpack)
    unsigned long long* d_wait = nullptr;    // [0] ns spent waiting for the neighbours' flags (k_halo_unpack), [1] waits counted
    int32_t* d_align = nullptr;      // operand of the start-alignment all-reduce (ab_comm_align)
};
static inline char* ipc_slot(char* block, size_t stride, int par, int side) { return block + (size_t)(par * 2 + side) * stride; }
static inline uint32_t* ipc_flag(char* block, size_t stride, int par, int side) {
    return reinterpret_cast<uint32_t*>(block + 4 * stride) + (size_t)(par * 2 + side) * 32;      // 128 B apart
}

#define AB_NCCL(h, call)                                                                                   \
    do {                                                                                                   \
        int _r = (call);                                                                                   \
        if (_r != 0) return ab_fail((h), AB_ERR_COMM, std::string("NCCL error: ") + g_nccl.GetErrorString(_r) + " in " #call); \
    } while (0)

// slab layout: gdl layers over nranks ranks, the first (gdl % nranks) ranks hold one layer more
static void slab_of(int gdl, int nranks, int rank, int* first, int* count) {
    const int base = gdl / nranks, rem = gdl % nranks;
    *count = base + (rank < rem ? 1 : 0);
    *first = rank * base + (rank < rem ? rank : rem);
}

// fine keys of one reference-cell layer of the window
static int64_t keys_per_layer(const ab_handle* h) { return (h->nfine / h->dims[h->D - 1]); }
static int64_t cells_per_layer(const ab_handle* h) { return (h->ncells / h->dims[h->D - 1]); }

// ---------------------------------------------------------------------------------------------------
// message: [hdr 16 B][cs slice (ncs + 1) u32, padded to 16 B][pos D x C][vel D x C][id C][key C]
// ---------------------------------------------------------------------------------------------------
struct MsgCols { char* p[AB_MSG_MAXF]; };   // the columns' device arrays (current buffer), same order
static MsgLayout msg_layout(const ab_handle* h, int64_t C, uint32_t ncs) {
    const int D = h->D;
    const size_t tsize = h->tsize;
    MsgLayout m;
    m.ncs = ncs;
    m.C = C;
    m.off_cs = 16;
    m.off_pos = (m.off_cs + ((size_t)ncs + 1) * 4 + 15) / 16 * 16;
    m.off_vel = m.off_pos + (size_t)D * C * tsize;
    m.off_id = m.off_vel + (size_t)D * C * tsize;
    m.off_key = m.off_id + (size_t)C * 4;
    size_t end = (m.off_key + (size_t)C * 4 + 15) / 16 * 16;
    m.nx = 0;
    for (size_t f = 2; f < h->fields.size() && m.nx < AB_MSG_MAXF; ++f) {
        m.off_x[m.nx] = end;
        m.x_ncomp[m.nx] = h->fields[f].ncomp;
        m.x_esize[m.nx] = (int)h->fields[f].esize;
        end = (end + (size_t)h->fields[f].ncomp * C * h->fields[f].esize + 15) / 16 * 16;
        m.nx++;
    }
    for (int i = m.nx; i < AB_MSG_MAXF; ++i) { m.off_x[i] = 0; m.x_ncomp[i] = 0; m.x_esize[i] = 0; }
    m.bytes = end;
    return m;
}
static MsgCols msg_cols(const ab_handle* h) {
    MsgCols c;
    int k = 0;
    for (size_t f = 2; f < h->fields.size() && k < AB_MSG_MAXF; ++f) c.p[k++] = (char*)h->fields[f].buf[h->cur];
    for (; k < AB_MSG_MAXF; ++k) c.p[k] = nullptr;
    return c;
}
// one element of a user column: agent slot `s` of a [ncomp x cap] array <-> message row `t` of a [ncomp x C] array
__device__ __forceinline__ void copy_elem(char* dst, const char* src, int esize) {
    if (esize == 8) *reinterpret_cast<uint64_t*>(dst) = *reinterpret_cast<const uint64_t*>(src);
    else if (esize == 4) *reinterpret_cast<uint32_t*>(dst) = *reinterpret_cast<const uint32_t*>(src);
    else *dst = *src;
}

// Packs the boundary layers of the CURRENT sorted state: blockIdx.y = 0 bottom W owned layers (keys
// [0, ncs)), 1 = top W owned layers (keys [kown - ncs, kown)).
template <class E, int D>
__global__ void __launch_bounds__(256)
k_halo_pack(MsgLayout m, MsgCols xc, uint32_t kown, const uint32_t* __restrict__ cs, const E* __restrict__ pos,
            const E* __restrict__ vel, const int32_t* __restrict__ id, const uint32_t* __restrict__ key, int64_t cap,
            char* __restrict__ buf0, char* __restrict__ buf1, int* __restrict__ err,
            uint32_t* done_ctr, uint32_t* flag0, uint32_t* flag1, uint32_t epoch) {
    const int dir = blockIdx.y;
    char* buf = dir == 0 ? buf0 : buf1;                  // nullptr: no neighbour on this side (non-periodic last axis, end rank)
    const uint32_t k0 = dir == 0 ? 0u : kown - m.ncs;
    const uint32_t s0 = cs[k0], s1 = cs[k0 + m.ncs];
    uint32_t cnt = buf ? s1 - s0 : 0u;
    if (cnt > (uint32_t)m.C) {
        if (blockIdx.x == 0 && threadIdx.x == 0) atomicOr(err, AB_EF_COMM | AB_EF_COMM_CAP);
        cnt = (uint32_t)m.C;
    }
    const uint32_t t0 = blockIdx.x * blockDim.x + threadIdx.x, stride = gridDim.x * blockDim.x;
    if (t0 == 0 && buf) {
        uint32_t* hdr = reinterpret_cast<uint32_t*>(buf);
        hdr[0] = cnt; hdr[1] = k0; hdr[2] = m.ncs; hdr[3] = 0xAB200u;
    }
    uint32_t* mcs = reinterpret_cast<uint32_t*>(buf + m.off_cs);
    if (buf) for (uint32_t i = t0; i <= m.ncs; i += stride) mcs[i] = min(cs[k0 + i] - s0, cnt);
    E* mpos = reinterpret_cast<E*>(buf + m.off_pos);
    E* mvel = reinterpret_cast<E*>(buf + m.off_vel);
    int32_t* mid = reinterpret_cast<int32_t*>(buf + m.off_id);
    uint32_t* mkey = reinterpret_cast<uint32_t*>(buf + m.off_key);
    for (uint32_t t = t0; t < cnt; t += stride) {
        const size_t s = (size_t)s0 + t;
#pragma unroll
        for (int d = 0; d < D; ++d) {
            mpos[(size_t)d * m.C + t] = pos[(size_t)d * cap + s];
            mvel[(size_t)d * m.C + t] = vel[(size_t)d * cap + s];
        }
        mid[t] = id[s];
        mkey[t] = key[s];
        for (int f = 0; f < m.nx; ++f) {
            const int es = m.x_esize[f];
            for (int c = 0; c < m.x_ncomp[f]; ++c)
                copy_elem(buf + m.off_x[f] + ((size_t)c * m.C + t) * es, xc.p[f] + ((size_t)c * cap + s) * es, es);
        }
    }
    if (done_ctr) {
        // direct exchange: the LAST CTA to finish publishes the exchange number in the neighbours' flags.
        // The CTA barrier orders every thread's peer-memory stores before thread 0's system-scope fence
        // (fences are cumulative), so one fence per CTA — not one per thread — precedes the count.
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence_system();
            const uint32_t total = gridDim.x * gridDim.y;
            if (atomicAdd(done_ctr, 1u) == total - 1) {
                __threadfence_system();
                *done_ctr = 0;
                if (flag0) asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(flag0), "r"(epoch) : "memory");
                if (flag1) asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(flag1), "r"(epoch) : "memory");
            }
        }
    }
}

// How long a rank waits for a neighbour's message before it calls the peer dead.  A neighbour that is merely busy on the
// host (a checkpoint, a collect, the chunked scatter of a 100 M-agent state) must not trip it, so the default is generous;
// AB_HALO_TIMEOUT_S (seconds, read when the direct exchange is set up) overrides it.
__device__ unsigned long long g_halo_timeout_ns = 120000000000ull;

// Spins until *f == epoch (a neighbour's k_halo_pack has delivered); gives up after g_halo_timeout_ns (dead peer).
__device__ __forceinline__ bool wait_flag(const uint32_t* f, uint32_t epoch) {
    const unsigned long long limit = g_halo_timeout_ns;
    unsigned long long t0, t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    for (;;) {
        uint32_t v;
        asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(f) : "memory");
        if (v == epoch) return true;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        if (t - t0 > limit) return false;
        __nanosleep(100);
    }
}

// Side-stream install: ONE thread waits for the neighbours' flags, so that the wait occupies one SM and not a CTA on every
// SM (a grid of spinning CTAs next to the re-sort that must publish the neighbour's flag dead-locked two ranks: the
// resident spinners kept the SMs from taking the re-sort's CTAs).  k_halo_unpack then finds both flags set.
__global__ void k_halo_wait(const uint32_t* wflag0, const uint32_t* wflag1, uint32_t epoch, int* __restrict__ err,
                            unsigned long long* wait_stat) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    unsigned long long t0, t1;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    const bool ok = (!wflag0 || wait_flag(wflag0, epoch)) && (!wflag1 || wait_flag(wflag1, epoch));
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
    if (wait_stat) { atomicAdd(&wait_stat[0], t1 - t0); atomicAdd(&wait_stat[1], 1ull); }
    if (!ok) atomicOr(err, AB_EF_COMM | AB_EF_COMM_TIMEOUT);
}

// Installs the received layers as this rank's halo.  Blocks [0, nbc) copy the upper neighbour's bottom layers
// (keys [kown, kown + ncs)), blocks [nbc, 2 nbc) the lower neighbour's top layers (keys [kown + ncs, kown + 2 ncs)).
// Halo slots follow the owned ones: [n_own, n_own + cnt_u) and [n_own + cnt_u, n_own + cnt_u + cnt_l).
// Blocks from 2 nbc on (approximate search) build the halo cells' aggregates — count, Σ(pos - centre), Σ vel per
// fine row of a reference cell, the canonical 8-strided sum of k_resort / k_cell_agg — straight from the messages,
// so no separate aggregate launch follows the install.
template <class T, int D>
__global__ void __launch_bounds__(256)
k_halo_unpack(SpaceDev<T, D> sp, MsgLayout m, MsgCols xc, uint32_t kown, uint32_t* __restrict__ cs, T* __restrict__ pos, T* __restrict__ vel,
              int32_t* __restrict__ id, uint32_t* __restrict__ key, int64_t cap, const char* __restrict__ buf0,
              const char* __restrict__ buf1, int* __restrict__ err, const uint32_t* wflag0, const uint32_t* wflag1,
              uint32_t epoch, int nbc, int nba, T* __restrict__ agg, int64_t agg_stride, uint32_t halo_cell0,
              uint32_t cells_per_side, unsigned long long* wait_stat, const uint32_t* __restrict__ row_cnt, int nrows,
              const uint32_t* __restrict__ nown_ptr) {
    if (wflag0 || wflag1) {
        // direct exchange: both messages were stored into our memory by the neighbours' pack kernels; they
        // never wait for us in return (their pack depends only on their own steps; slots alternate by parity)
        __shared__ int s_ok;
        if (threadIdx.x == 0) {
            unsigned long long t0, t1;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
            s_ok = (!wflag0 || wait_flag(wflag0, epoch)) && (!wflag1 || wait_flag(wflag1, epoch));
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
            if (blockIdx.x == 0 && wait_stat) { atomicAdd(&wait_stat[0], t1 - t0); atomicAdd(&wait_stat[1], 1ull); }
        }
        __syncthreads();
        if (!s_ok) {
            if (blockIdx.x == 0 && threadIdx.x == 0) atomicOr(err, AB_EF_COMM | AB_EF_COMM_TIMEOUT);
            return;
        }
    }
    // buf0 / buf1 == nullptr: no neighbour on that side (non-periodic last axis, end rank) — its halo layers stay empty
    const uint32_t* h0 = reinterpret_cast<const uint32_t*>(buf0);
    const uint32_t* h1 = reinterpret_cast<const uint32_t*>(buf1);
    // owned agents after this step's re-sort: the cell_start entry behind the owned keys — or, when this launch runs
    // CONCURRENTLY with that re-sort (side stream), the sum of the step's per-row counters, complete since the sweep
    uint32_t n_own;
    if (row_cnt) {
        __shared__ uint32_t s_nown;
        if (threadIdx.x == 0) s_nown = 0;
        __syncthreads();
        uint32_t part = 0;
        for (int r = threadIdx.x; r < nrows; r += blockDim.x) part += row_cnt[r];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
        if ((threadIdx.x & 31) == 0 && part) atomicAdd(&s_nown, part);
        __syncthreads();
        n_own = s_nown;
    } else n_own = nown_ptr ? *nown_ptr : cs[kown];      // nown_ptr: the scanned tile totals' last entry (scan path)
    uint32_t cnt_u = buf0 ? h0[0] : 0u, cnt_l = buf1 ? h1[0] : 0u;
    bool bad = (buf0 && (h0[3] != 0xAB200u || h0[2] != m.ncs)) || (buf1 && (h1[3] != 0xAB200u || h1[2] != m.ncs));
    const bool nofit = (uint64_t)n_own + cnt_u + cnt_l > (uint64_t)cap;   // window does not fit the handle's capacity
    if (bad || nofit) {
        if (blockIdx.x == 0 && threadIdx.x == 0) atomicOr(err, AB_EF_COMM | (bad ? AB_EF_COMM_HDR : 0) | (nofit ? AB_EF_COMM_WINDOW : 0));
        bad = true;
        cnt_u = cnt_l = 0;
    }
    const bool agg_role = (int)blockIdx.x >= 2 * nbc;
    const int rb = agg_role ? (int)blockIdx.x - 2 * nbc : (int)blockIdx.x;
    const int per = agg_role ? nba : nbc;
    const int src = rb >= per ? 1 : 0;
    const int bx = rb - src * per;                       // block index within the role and side
    const char* buf = src == 0 ? buf0 : buf1;
    const uint32_t* hdr = src == 0 ? h0 : h1;
    const uint32_t cnt = src == 0 ? cnt_u : cnt_l;
    const uint32_t kdst = src == 0 ? kown : kown + m.ncs;
    const uint32_t* mcs = reinterpret_cast<const uint32_t*>(buf + m.off_cs);
    const T* mpos = reinterpret_cast<const T*>(buf + m.off_pos);
    const T* mvel = reinterpret_cast<const T*>(buf + m.off_vel);
    if (!agg_role) {
        const uint32_t base = src == 0 ? n_own : n_own + cnt_u;
        const uint32_t t0 = bx * blockDim.x + threadIdx.x, stride = per * blockDim.x;
        if (!buf) {                                          // no neighbour: empty halo cells
            for (uint32_t i = t0 + 1; i <= m.ncs; i += stride) cs[kdst + i] = base;
            return;
        }
        const uint32_t kdelta = kdst - hdr[1];               // sender's key frame -> this window's (mod 2^32)
        for (uint32_t i = t0 + 1; i <= m.ncs; i += stride) cs[kdst + i] = base + (bad ? 0u : mcs[i]);
        const int32_t* mid = reinterpret_cast<const int32_t*>(buf + m.off_id);
        const uint32_t* mkey = reinterpret_cast<const uint32_t*>(buf + m.off_key);
        for (uint32_t t = t0; t < cnt; t += stride) {
            const size_t s = (size_t)base + t;
#pragma unroll
            for (int d = 0; d < D; ++d) {
                pos[(size_t)d * cap + s] = mpos[(size_t)d * m.C + t];
                vel[(size_t)d * cap + s] = mvel[(size_t)d * m.C + t];
            }
            id[s] = mid[t];
            key[s] = mkey[t] + kdelta;
            for (int f = 0; f < m.nx; ++f) {
                const int es = m.x_esize[f];
                for (int c = 0; c < m.x_ncomp[f]; ++c)
                    copy_elem(xc.p[f] + ((size_t)c * cap + s) * es, buf + m.off_x[f] + ((size_t)c * m.C + t) * es, es);
            }
        }
        return;
    }
    // ---- aggregates of the halo cells of this side: one LANE per (fine row of a reference cell, reference cell) ----
    // Canonical sum of a cell's n agents (k_resort / k_cell_agg): 8 strided partials a_g = Σ_i x[g + 8 i] combined as
    // ((a0+a4)+(a2+a6)) + ((a1+a5)+(a3+a7)).  Cells with n <= 8 — nearly all of them in a dilute state, 1.1 agents per
    // item at F3D's density — are summed by their own lane (each partial is one term); fuller cells by lanes 0..7 of
    // the warp, one cell at a time.
    constexpr int NC = 1 + 2 * D;
    const int K = sp.K, kb = sp.kb;
    const int nfr = D == 2 ? K : K * K;
    const int lane = threadIdx.x & 31;
    const uint32_t nitems = cells_per_side * (uint32_t)nfr;
    for (uint32_t g0 = (uint32_t)bx * blockDim.x + (threadIdx.x & ~31u); g0 < nitems; g0 += (uint32_t)per * blockDim.x) {
        const uint32_t g = g0 + (uint32_t)lane;                     // the loop condition is warp-uniform
        const bool live = g < nitems;
        const int part = live ? (int)(g / cells_per_side) : 0;      // cells fastest: neighbouring lanes, neighbouring cells
        const uint32_t cl = live ? g - (uint32_t)part * cells_per_side : 0u;
        const uint32_t c = halo_cell0 + (uint32_t)src * cells_per_side + cl;
        int cc[D];
        unlin_cell<D>(sp.dims, c, cc);
        const FineRows fr = fine_rows<D>(sp.dims, kb, cc[1], D == 3 ? cc[2] : 0);
        const uint32_t i0 = fr.base[part] + ((uint32_t)cc[0] << kb) - kdst;     // index into the message's cell_start slice
        uint32_t b = 0, e = 0;
        if (live && !bad && buf) { b = min(mcs[i0], cnt); e = min(mcs[i0 + (uint32_t)K], cnt); }
        cc[D - 1] = local_to_global_last<T, D>(sp, cc[D - 1]);                   // centres are global
        T ctr[D];
#pragma unroll
        for (int d = 0; d < D; ++d) ctr[d] = sp.spacing * ((T)(cc[d] + 1) - T(0.5));
        const uint32_t n = e - b;
        T* dst = agg + (size_t)part * NC * agg_stride + c;
        if (live && n <= 8u) {
            // term i of component d (0 beyond n): the 8-lane path's partial a_i
            auto term = [&](int d, uint32_t i) -> T {
                if (i >= n) return T(0);
                return d < D ? T(0) + (mpos[(size_t)d * m.C + b + i] - ctr[d]) : T(0) + mvel[(size_t)(d - D) * m.C + b + i];
            };
            dst[0] = (T)n;
#pragma unroll
            for (int d = 0; d < 2 * D; ++d) {
                const T q = (term(d, 0) + term(d, 4)) + (term(d, 2) + term(d, 6));
                dst[(size_t)(1 + d) * agg_stride] = q + ((term(d, 1) + term(d, 5)) + (term(d, 3) + term(d, 7)));
            }
        }
        unsigned big = __ballot_sync(0xffffffffu, live && n > 8u);
        while (big) {
            const int srcl = __ffs(big) - 1;
            big &= big - 1;
            const uint32_t gb = __shfl_sync(0xffffffffu, b, srcl), ge = __shfl_sync(0xffffffffu, e, srcl);
            T gctr[D];
#pragma unroll
            for (int d = 0; d < D; ++d) gctr[d] = __shfl_sync(0xffffffffu, ctr[d], srcl);
            T acc[2 * D];
#pragma unroll
            for (int d = 0; d < 2 * D; ++d) acc[d] = T(0);
            if (lane < 8) {
                for (uint32_t t = gb + (uint32_t)lane; t < ge; t += 8) {
#pragma unroll
                    for (int d = 0; d < D; ++d) {
                        acc[d] += mpos[(size_t)d * m.C + t] - gctr[d];
                        acc[D + d] += mvel[(size_t)d * m.C + t];
                    }
                }
            }
#pragma unroll
            for (int d = 0; d < 2 * D; ++d) {
                acc[d] += __shfl_xor_sync(0xffffffffu, acc[d], 4);
                acc[d] += __shfl_xor_sync(0xffffffffu, acc[d], 2);
                acc[d] += __shfl_xor_sync(0xffffffffu, acc[d], 1);
            }
            // the owner lane holds the destination; lane 0 holds the sums
            const unsigned long long dptr = __shfl_sync(0xffffffffu, (unsigned long long)dst, srcl);
            if (lane == 0) {
                T* gd = reinterpret_cast<T*>(dptr);
                gd[0] = (T)(ge - gb);
#pragma unroll
                for (int d = 0; d < 2 * D; ++d) gd[(size_t)(1 + d) * agg_stride] = acc[d];
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------
void ab_comm_free(ab_handle* h) {
    ab_comm* c = h->comm;
    if (!c) return;
    for (int i = 0; i < 2; ++i) {
        if (c->sendbuf[i]) cudaFree(c->sendbuf[i]);
        if (c->recvbuf[i]) cudaFree(c->recvbuf[i]);
    }
    if (c->ipc_lower) cudaIpcCloseMemHandle(c->ipc_lower);
    if (c->ipc_upper && c->ipc_upper != c->ipc_lower) cudaIpcCloseMemHandle(c->ipc_upper);
    if (c->ipc_block) cudaFree(c->ipc_block);
    if (c->h_counts) cudaFreeHost(c->h_counts);
    if (c->d_wait) cudaFree(c->d_wait);
    if (c->d_align) cudaFree(c->d_align);
    if (c->ev_packed) cudaEventDestroy(c->ev_packed);
    if (c->ev_consumed) cudaEventDestroy(c->ev_consumed);
    if (c->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(c->comm);
    delete c;
    h->comm = nullptr;
}

static int ensure_buffers(ab_handle* h) {
    ab_comm* c = h->comm;
    const uint32_t ncs = (uint32_t)(AB_HALO_W * keys_per_layer(h));
    int64_t C = c->want_cap;
    if (C <= 0) {
        // default: 3x the mean population of W layers of this window (all ranks hold the same number of
        // agents per layer on average; ranks MUST agree on C — callers with skewed states set it explicitly)
        return ab_fail(h, AB_ERR_STATE, "ab_comm_set_capacity must be called before stepping a sharded handle");
    }
    C = (C + 63) / 64 * 64;
    if (c->ipc_on) {
        if (C != c->ipc_cap) return ab_fail(h, AB_ERR_STATE, "message capacity changed after ab_comm_ipc_export");
        const MsgLayout mi = msg_layout(h, C, ncs);
        if (mi.bytes > c->ipc_stride) return ab_fail(h, AB_ERR_STATE, "message does not fit the exported block");
        c->cap_agents = C; c->msg_bytes = mi.bytes; c->msg_K = h->K;
        return AB_OK;
    }
    const MsgLayout m = msg_layout(h, C, ncs);
    if (c->cap_agents == C && c->msg_K == h->K && c->sendbuf[0] && c->msg_bytes == m.bytes) return AB_OK;
    AB_CUDA(h, cudaStreamSynchronize(h->stream));
    for (int i = 0; i < 2; ++i) {
        if (c->sendbuf[i]) { AB_CUDA(h, cudaFree(c->sendbuf[i])); c->sendbuf[i] = nullptr; }
        if (c->recvbuf[i]) { AB_CUDA(h, cudaFree(c->recvbuf[i])); c->recvbuf[i] = nullptr; }
        AB_CUDA(h, cudaMalloc(&c->sendbuf[i], m.bytes));
        AB_CUDA(h, cudaMalloc(&c->recvbuf[i], m.bytes));
    }
    c->cap_agents = C;
    c->msg_bytes = m.bytes;
    c->msg_K = h->K;
    return AB_OK;
}

// Destinations of this step's boundary messages (parity slot of the neighbours' blocks, or the send buffers), the
// waits that protect them, and — direct exchange — the flags to publish.  out->enabled says whether the tiled
// re-sort may deliver the messages itself (no user columns travel: those are permuted after the re-sort).
int ab_comm_pack_begin(ab_handle* h, HaloOut* out) {
    ab_comm* c = h->comm;
    out->enabled = 0;
    if (!c || !h->win_W) return ab_fail(h, AB_ERR_STATE, "handle is not sharded");
    if (h->halo_stale)
        return ab_fail(h, AB_ERR_COMM, "a halo message timed out earlier: the window's halo layers are stale; every rank must call ab_comm_exchange before stepping again");
    AB_TRY(ensure_buffers(h));
    if (h->fields.size() > 2u + AB_MSG_MAXF) return ab_fail(h, AB_ERR_INVALID, "too many user columns for the halo message");
    const uint32_t ncs = (uint32_t)(AB_HALO_W * keys_per_layer(h));
    out->ncs = ncs;
    out->kown = (uint32_t)(h->win_nl * keys_per_layer(h));
    out->m = msg_layout(h, c->cap_agents, ncs);
    // in-process peers: they must have consumed the previous message before it is overwritten
    if (c->linked_local && c->peer_lower) AB_CUDA(h, cudaStreamWaitEvent(h->stream, c->peer_lower->comm->ev_consumed, 0));
    if (c->linked_local && c->peer_upper) AB_CUDA(h, cudaStreamWaitEvent(h->stream, c->peer_upper->comm->ev_consumed, 0));
    // my bottom layers are the lower neighbour's "from upper" message (side 0), my top layers the upper
    // neighbour's "from lower" one (side 1)
    c->cur_par = (int)(c->exchanges & 1);
    out->buf[0] = !c->has_lower ? nullptr : c->ipc_on ? ipc_slot(c->ipc_lower, c->ipc_stride, c->cur_par, 0) : (char*)c->sendbuf[0];
    out->buf[1] = !c->has_upper ? nullptr : c->ipc_on ? ipc_slot(c->ipc_upper, c->ipc_stride, c->cur_par, 1) : (char*)c->sendbuf[1];
    out->done_ctr = c->ipc_on ? ipc_flag(c->ipc_block, c->ipc_stride, 2, 0) : nullptr;
    out->flag[0] = c->ipc_on && c->has_lower ? ipc_flag(c->ipc_lower, c->ipc_stride, c->cur_par, 0) : nullptr;
    out->flag[1] = c->ipc_on && c->has_upper ? ipc_flag(c->ipc_upper, c->ipc_stride, c->cur_par, 1) : nullptr;
    out->epoch = (uint32_t)(c->exchanges + 1);
    // The re-sort delivers the messages itself when they are SMALL (2-D slabs: a few thousand agents per side — the pack
    // launch and its latency are what there is to save).  Large messages (3-D: tens of MB per side) go through
    // k_halo_pack: it reads the sorted state sequentially and stores 256-byte runs over NVLink, where the re-sort's
    // boundary tiles would store 8 bytes at a time (measured, F3D-12.5M per GPU: k_resort 1.66 ms with the fused
    // delivery against 1.04 ms unsharded).
    const bool small = out->m.bytes <= c->fused_pack_max_bytes;
    out->enabled = (out->m.nx == 0 && !c->no_fused_pack && small) ? 1 : 0;
    return AB_OK;
}

template <class E, int D>
static int pack_t(ab_handle* h) {
    ab_comm* c = h->comm;
    if (h->halo_packed) {                                // delivered by the re-sort of this step (HaloOut)
        h->halo_packed = false;
        if (c->linked_local) AB_CUDA(h, cudaEventRecord(c->ev_packed, h->stream));
        return AB_OK;
    }
    HaloOut ho;
    AB_TRY(ab_comm_pack_begin(h, &ho));
    const int cur = h->cur;
    // one item per thread (up to 592 CTAs per side): the stores are posted writes, the cost is latency
    const int nb = (int)std::min<int64_t>(std::max<int64_t>((std::max<int64_t>(c->cap_agents, ho.ncs) + 255) / 256, 1), 592);
    {
        AB_LAUNCH(h, "k_halo_pack");
        k_halo_pack<E, D><<<dim3(nb, 2), 256, 0, h->stream>>>(ho.m, msg_cols(h), ho.kown, h->cell_start, (const E*)h->fields[0].buf[cur],
                                                             (const E*)h->fields[1].buf[cur], h->id[cur], h->key[cur], h->cap,
                                                             ho.buf[0], ho.buf[1], h->d_err, ho.done_ctr, ho.flag[0], ho.flag[1], ho.epoch);
    }
    if (c->linked_local) AB_CUDA(h, cudaEventRecord(c->ev_packed, h->stream));
    return AB_OK;
}

// Moves the two boundary messages: NCCL between processes, device-to-device copies between in-process peers.
static int transfer(ab_handle* h) {
    ab_comm* c = h->comm;
    const size_t bytes = c->msg_bytes;
    if (c->linked_local) {
        ab_handle *lo = c->peer_lower, *up = c->peer_upper;
        if (up) AB_CUDA(h, cudaStreamWaitEvent(h->stream, up->comm->ev_packed, 0));
        if (lo) AB_CUDA(h, cudaStreamWaitEvent(h->stream, lo->comm->ev_packed, 0));
        if ((up && up->comm->msg_bytes != bytes) || (lo && lo->comm->msg_bytes != bytes)) return ab_fail(h, AB_ERR_COMM, "peers disagree on the message capacity");
        // my upper halo = the upper peer's bottom layers (its sendbuf[0]); my lower halo = the lower peer's top layers
        if (up) {
            if (up->device == h->device) AB_CUDA(h, cudaMemcpyAsync(c->recvbuf[0], up->comm->sendbuf[0], bytes, cudaMemcpyDeviceToDevice, h->stream));
            else AB_CUDA(h, cudaMemcpyPeerAsync(c->recvbuf[0], h->device, up->comm->sendbuf[0], up->device, bytes, h->stream));
        }
        if (lo) {
            if (lo->device == h->device) AB_CUDA(h, cudaMemcpyAsync(c->recvbuf[1], lo->comm->sendbuf[1], bytes, cudaMemcpyDeviceToDevice, h->stream));
            else AB_CUDA(h, cudaMemcpyPeerAsync(c->recvbuf[1], h->device, lo->comm->sendbuf[1], lo->device, bytes, h->stream));
        }
        AB_CUDA(h, cudaEventRecord(c->ev_consumed, h->stream));
    } else if (c->ipc_on) {
        // nothing to move: the neighbours' pack kernels store into our block; k_halo_unpack waits for their flags
    } else {
        if (!c->comm) return ab_fail(h, AB_ERR_STATE, "sharded handle without a communicator or in-process peers");
        // one group: bottom layers -> lower neighbour, top layers -> upper neighbour; receives in the matching
        // order (with 2 ranks both neighbours are the same peer: first send pairs with first receive)
        AB_NCCL(h, g_nccl.GroupStart());
        if (c->has_lower) AB_NCCL(h, g_nccl.Send(c->sendbuf[0], bytes, NCCL_UINT8, c->lower, c->comm, h->stream));
        if (c->has_upper) AB_NCCL(h, g_nccl.Send(c->sendbuf[1], bytes, NCCL_UINT8, c->upper, c->comm, h->stream));
        if (c->has_upper) AB_NCCL(h, g_nccl.Recv(c->recvbuf[0], bytes, NCCL_UINT8, c->upper, c->comm, h->stream));
        if (c->has_lower) AB_NCCL(h, g_nccl.Recv(c->recvbuf[1], bytes, NCCL_UINT8, c->lower, c->comm, h->stream));
        AB_NCCL(h, g_nccl.GroupEnd());
    }
    c->exchanges++;
    return AB_OK;
}

template <class T, int D>
static int unpack_t(ab_handle* h, bool want_agg) {
    ab_comm* c = h->comm;
    const int cur = h->cur;
    const uint32_t ncs = (uint32_t)(AB_HALO_W * keys_per_layer(h));
    const uint32_t kown = (uint32_t)(h->win_nl * keys_per_layer(h));
    const MsgLayout m = msg_layout(h, c->cap_agents, ncs);
    const int nbc = (int)std::min<int64_t>(std::max<int64_t>((std::max<int64_t>(c->cap_agents / 4, ncs) + 255) / 256, 1), 592);
    const int64_t cpl = cells_per_layer(h);
    const int nfr = D == 2 ? h->K : h->K * h->K;
    int nba = 0;
    if (want_agg) {
        if (!h->cell_agg) return ab_fail(h, AB_ERR_STATE, "halo aggregates requested before the re-sort produced the owned ones");
        nba = (int)std::min<int64_t>(std::max<int64_t>((AB_HALO_W * cpl * nfr + 255) / 256, 1), 1184);
        h->agg_epoch++;
    }
    const uint32_t* wf0 = c->ipc_on && c->has_upper ? ipc_flag(c->ipc_block, c->ipc_stride, c->cur_par, 0) : nullptr;
    const uint32_t* wf1 = c->ipc_on && c->has_lower ? ipc_flag(c->ipc_block, c->ipc_stride, c->cur_par, 1) : nullptr;
    const bool prewait = (h->unpack_rows || h->unpack_tb) && (wf0 || wf1) && h->side_unpack_now;
    if (prewait) {
        AB_LAUNCH(h, "k_halo_wait");
        k_halo_wait<<<1, 32, 0, h->stream>>>(wf0, wf1, (uint32_t)c->exchanges, h->d_err, c->d_wait);
    }
    {
        AB_LAUNCH(h, "k_halo_unpack");
        k_halo_unpack<T, D><<<dim3(2 * nbc + 2 * nba), 256, 0, h->stream>>>(make_space<T, D>(h), m, msg_cols(h), kown, h->cell_start, (T*)h->fields[0].buf[cur],
                                                               (T*)h->fields[1].buf[cur], h->id[cur], h->key[cur], h->cap,
                                                               !c->has_upper ? nullptr : c->ipc_on ? ipc_slot(c->ipc_block, c->ipc_stride, c->cur_par, 0) : (const char*)c->recvbuf[0],
                                                               !c->has_lower ? nullptr : c->ipc_on ? ipc_slot(c->ipc_block, c->ipc_stride, c->cur_par, 1) : (const char*)c->recvbuf[1], h->d_err,
                                                               wf0, wf1,
                                                               (uint32_t)c->exchanges, nbc, nba, (T*)h->cell_agg, (int64_t)h->ncells,
                                                               (uint32_t)(h->win_nl * cpl), (uint32_t)(AB_HALO_W * cpl), prewait ? nullptr : c->d_wait,
                                                               h->unpack_rows ? h->row_count + (size_t)(h->rs_par ^ 1) * h->row_cap : nullptr, h->rs_nrows,
                                                               h->unpack_tb ? h->tile_base + h->rs_ntiles : nullptr);
        h->unpack_rows = h->unpack_tb = false;
    }
    return AB_OK;
}

int ab_comm_pack(ab_handle* h) {
    if (!h->comm || !h->win_W) return ab_fail(h, AB_ERR_STATE, "handle is not sharded");
    if (h->tsize == 8) return h->D == 2 ? pack_t<uint64_t, 2>(h) : pack_t<uint64_t, 3>(h);
    return h->D == 2 ? pack_t<uint32_t, 2>(h) : pack_t<uint32_t, 3>(h);
}
int ab_comm_transfer(ab_handle* h) { return transfer(h); }
int ab_comm_unpack(ab_handle* h, bool want_agg) {
    if (h->tsize == 8) return h->D == 2 ? unpack_t<double, 2>(h, want_agg) : unpack_t<double, 3>(h, want_agg);
    return h->D == 2 ? unpack_t<float, 2>(h, want_agg) : unpack_t<float, 3>(h, want_agg);
}
bool ab_comm_local(const ab_handle* h) { return h->comm && h->comm->linked_local; }
bool ab_comm_direct_on(const ab_handle* h) { return h->comm && h->comm->ipc_on; }

int ab_comm_refresh_counts(ab_handle* h) {
    ab_comm* c = h->comm;
    if (!c) return AB_OK;
    AB_TRY(ab_ensure_index(h));
    const int64_t kown = h->win_nl * keys_per_layer(h);
    AB_CUDA(h, cudaMemcpyAsync(&c->h_counts[0], h->cell_start + kown, 4, cudaMemcpyDeviceToHost, h->stream));
    AB_CUDA(h, cudaMemcpyAsync(&c->h_counts[1], h->cell_start + h->nfine, 4, cudaMemcpyDeviceToHost, h->stream));
    AB_CUDA(h, cudaStreamSynchronize(h->stream));
    c->n_own = c->h_counts[0];
    h->n = c->h_counts[1];
    return AB_OK;
}

__global__ void k_rows_to_soa(const char* __restrict__ rows, char* __restrict__ dst, int64_t cap, int n, int ncomp, int esize) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= (int64_t)n * ncomp) return;
    const int i = (int)(t / ncomp), c = (int)(t - (int64_t)i * ncomp);
    for (int b = 0; b < esize; ++b) dst[((size_t)c * cap + i) * esize + b] = rows[(size_t)t * esize + b];
}
__global__ void k_soa_to_rows(const char* __restrict__ src, char* __restrict__ rows, int64_t cap, int n, int ncomp, int esize) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= (int64_t)n * ncomp) return;
    const int i = (int)(t / ncomp), c = (int)(t - (int64_t)i * ncomp);
    for (int b = 0; b < esize; ++b) rows[(size_t)t * esize + b] = src[((size_t)c * cap + i) * esize + b];
}
__global__ void k_gids_in(const int64_t* __restrict__ gids, int32_t* __restrict__ id, int n, int* __restrict__ err) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int64_t g = gids[i];
    if (g < 1 || g > 2147483647LL) { atomicOr(err, AB_EF_OVERFLOW); id[i] = 0; return; }
    id[i] = (int32_t)(g - 1);
}
__global__ void k_gids_out(const int32_t* __restrict__ id, int64_t* __restrict__ gids, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) gids[i] = (int64_t)id[i] + 1;
}

extern "C" {

int ab_comm_unique_id(void* out128) {
    if (!out128) return AB_ERR_INVALID;
    if (!nccl_load()) return ab_fail(nullptr, AB_ERR_COMM, g_nccl_err);
    nccl_uid_t id;
    int r = g_nccl.GetUniqueId(&id);
    if (r != 0) return ab_fail(nullptr, AB_ERR_COMM, std::string("ncclGetUniqueId: ") + g_nccl.GetErrorString(r));
    memcpy(out128, &id, 128);
    return AB_OK;
}

// Turns a freshly created ContinuousSpace handle (no agents yet) into rank `rank` of `nranks` slabs along
// the last axis.  nccl_unique_id == NULL builds the window without a communicator (layout tests).
int ab_comm_init(ab_handle* h, int rank, int nranks, const void* nccl_unique_id) {
    if (!h) return AB_ERR_INVALID;
    if (h->desc.kind != AB_SPACE_CONTINUOUS) return ab_fail(h, AB_ERR_INVALID, "slab decomposition is for ContinuousSpace handles");
    if (h->comm || h->win_W) return ab_fail(h, AB_ERR_STATE, "handle is already sharded");
    if (h->n != 0) return ab_fail(h, AB_ERR_STATE, "ab_comm_init must come before the agents are added");
    if (nranks < 2 || rank < 0 || rank >= nranks) return ab_fail(h, AB_ERR_INVALID, "need nranks >= 2 and 0 <= rank < nranks");
    const int D = h->D, gdl = h->dims[D - 1];
    const bool per_last = h->desc.periodic[D - 1] != 0;
    int first, nl;
    slab_of(gdl, nranks, rank, &first, &nl);
    const int minl = gdl / nranks;
    // every slab holds its own boundary layers (W) and the window must not overlap itself
    if (minl < AB_HALO_W + 1 || (per_last && minl + 2 * AB_HALO_W + (gdl % nranks ? 1 : 0) > gdl))
        return ab_fail(h, AB_ERR_INVALID, "too few cell layers per rank for the halo width (need >= 4 layers per rank and >= 2 ranks' worth of space)");
    AB_CUDA(h, cudaSetDevice(h->device));
    ab_comm* c = new ab_comm();
    c->rank = rank; c->nranks = nranks;
    c->lower = (rank + nranks - 1) % nranks;
    c->upper = (rank + 1) % nranks;
    // non-periodic last axis: the end slabs have one neighbour; the halo layers beyond the space stay empty
    if (const char* e = getenv("AB_FUSED_PACK")) c->no_fused_pack = atoi(e) == 0;
    if (const char* e = getenv("AB_FUSED_PACK_MAXMB")) c->fused_pack_max_bytes = (size_t)atol(e) << 20;
    c->has_lower = per_last || rank > 0;
    c->has_upper = per_last || rank < nranks - 1;
    if (cudaMallocHost((void**)&c->h_counts, 2 * sizeof(uint32_t)) != cudaSuccess) { delete c; return ab_fail(h, AB_ERR_CUDA, "cudaMallocHost failed"); }
    if (cudaMalloc((void**)&c->d_wait, 2 * sizeof(unsigned long long)) != cudaSuccess || cudaMemset(c->d_wait, 0, 2 * sizeof(unsigned long long)) != cudaSuccess ||
        cudaMalloc((void**)&c->d_align, 2 * sizeof(int32_t)) != cudaSuccess || cudaMemset(c->d_align, 0, 2 * sizeof(int32_t)) != cudaSuccess) {
        cudaFreeHost(c->h_counts); if (c->d_wait) cudaFree(c->d_wait); if (c->d_align) cudaFree(c->d_align); delete c;
        return ab_fail(h, AB_ERR_CUDA, "cudaMalloc failed");
    }
    if (nccl_unique_id) {
        if (!nccl_load()) { cudaFreeHost(c->h_counts); cudaFree(c->d_wait); cudaFree(c->d_align); delete c; return ab_fail(h, AB_ERR_COMM, g_nccl_err); }
        nccl_uid_t id;
        memcpy(&id, nccl_unique_id, 128);
        int r = g_nccl.CommInitRank(&c->comm, nranks, id, rank);
        if (r != 0) { cudaFreeHost(c->h_counts); cudaFree(c->d_wait); cudaFree(c->d_align); delete c; return ab_fail(h, AB_ERR_COMM, std::string("ncclCommInitRank: ") + g_nccl.GetErrorString(r)); }
    }
    h->comm = c;
    h->win_gdl = gdl; h->win_goff = first; h->win_nl = nl; h->win_W = AB_HALO_W;
    h->dims[D - 1] = nl + 2 * AB_HALO_W;
    int64_t ncells = 1;
    for (int d = 0; d < D; ++d) ncells *= h->dims[d];
    h->ncells = ncells;
    h->K = 1; h->kb = 0; h->ksh = 0;
    h->nfine = ncells;
    h->index_valid = false; h->agg_valid = false; h->pairs_n = -1;
    return AB_OK;
}

// In-process slabs: hs[r] is rank r of n (each created with ab_comm_init(h, r, n, NULL)); neighbours
// exchange their boundary layers by device-to-device copies ordered with CUDA events.  Such handles are
// stepped together with ab_step_flock_group.
int ab_comm_link_local(ab_handle** hs, int n) {
    if (!hs || n < 2) return AB_ERR_INVALID;
    for (int r = 0; r < n; ++r) {
        ab_handle* h = hs[r];
        if (!h || !h->comm || !h->win_W) return ab_fail(h, AB_ERR_STATE, "ab_comm_init(h, rank, n, NULL) first");
        if (h->comm->rank != r || h->comm->nranks != n) return ab_fail(h, AB_ERR_INVALID, "handles must be passed in rank order");
        if (h->comm->comm) return ab_fail(h, AB_ERR_STATE, "handle already has an NCCL communicator");
    }
    for (int r = 0; r < n; ++r) {
        ab_handle* h = hs[r];
        ab_comm* c = h->comm;
        AB_CUDA(h, cudaSetDevice(h->device));
        c->peer_lower = c->has_lower ? hs[(r + n - 1) % n] : nullptr;
        c->peer_upper = c->has_upper ? hs[(r + 1) % n] : nullptr;
        c->linked_local = true;
        if (!c->ev_packed) AB_CUDA(h, cudaEventCreateWithFlags(&c->ev_packed, cudaEventDisableTiming));
        if (!c->ev_consumed) AB_CUDA(h, cudaEventCreateWithFlags(&c->ev_consumed, cudaEventDisableTiming));
        AB_CUDA(h, cudaEventRecord(c->ev_consumed, h->stream));      // "nothing pending" for the first pack
        for (ab_handle* peer : {c->peer_lower, c->peer_upper})
            if (peer && peer->device != h->device) {
                int can = 0;
                cudaDeviceCanAccessPeer(&can, h->device, peer->device);
                if (can) { cudaError_t e = cudaDeviceEnablePeerAccess(peer->device, 0); if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError(); }
            }
    }
    return AB_OK;
}

// Direct exchange over peer memory.  Export: allocates this rank's receive block for the capacity set by
// ab_comm_set_capacity and returns its 64-byte CUDA IPC handle; the host exchanges the handles (any
// transport) and imports the two neighbours'.  From then on the boundary messages are stored by
// k_halo_pack straight into the neighbour's memory (NVLink) and NCCL is not used on the data path.
int ab_comm_ipc_export(ab_handle* h, void* out64) {
    if (!h || !h->comm || !h->win_W) return ab_fail(h, AB_ERR_STATE, "handle is not sharded");
    if (!out64) return ab_fail(h, AB_ERR_INVALID, "null output");
    ab_comm* c = h->comm;
    if (c->linked_local) return ab_fail(h, AB_ERR_STATE, "in-process slabs exchange by device copies");
    if (c->want_cap <= 0) return ab_fail(h, AB_ERR_STATE, "ab_comm_set_capacity first");
    if (c->ipc_block) return ab_fail(h, AB_ERR_STATE, "already exported");
    AB_CUDA(h, cudaSetDevice(h->device));
    const int64_t C = (c->want_cap + 63) / 64 * 64;
    int ksh_max = ((h->ncells << h->D) <= h->nfine_cap) ? h->D : 0;                  // finest refinement a step may choose
    if (h->D == 2 && (h->ncells << 4) <= h->nfine_cap) ksh_max = 4;
    const uint32_t ncs_max = (uint32_t)(AB_HALO_W * ((h->ncells << ksh_max) / h->dims[h->D - 1]));
    c->ipc_stride = (msg_layout(h, C, ncs_max).bytes + 255) / 256 * 256;       // user columns must be registered before the export
    c->ipc_cap = C;
    AB_CUDA(h, cudaMalloc((void**)&c->ipc_block, 4 * c->ipc_stride + 4096));
    AB_CUDA(h, cudaMemset(c->ipc_block, 0, 4 * c->ipc_stride + 4096));
    cudaIpcMemHandle_t mh;
    AB_CUDA(h, cudaIpcGetMemHandle(&mh, c->ipc_block));
    static_assert(sizeof(mh) == 64, "CUDA IPC handle size");
    memcpy(out64, &mh, 64);
    return AB_OK;
}
int ab_comm_ipc_import(ab_handle* h, const void* lower64, const void* upper64) {
    if (!h || !h->comm || !h->win_W) return ab_fail(h, AB_ERR_STATE, "handle is not sharded");
    ab_comm* c = h->comm;
    if (!c->ipc_block) return ab_fail(h, AB_ERR_STATE, "ab_comm_ipc_export first");
    if ((c->has_lower && !lower64) || (c->has_upper && !upper64)) return ab_fail(h, AB_ERR_INVALID, "null handle");
    if (c->ipc_on) return ab_fail(h, AB_ERR_STATE, "already imported");
    AB_CUDA(h, cudaSetDevice(h->device));
    cudaIpcMemHandle_t ml, mu;
    void* p = nullptr;
    if (c->has_lower) {
        memcpy(&ml, lower64, 64);
        AB_CUDA(h, cudaIpcOpenMemHandle(&p, ml, cudaIpcMemLazyEnablePeerAccess));
        c->ipc_lower = (char*)p;
    }
    if (c->has_upper) {
        if (c->has_lower && c->lower == c->upper) c->ipc_upper = c->ipc_lower;          // two ranks, periodic: one neighbour, one mapping
        else {
            memcpy(&mu, upper64, 64);
            AB_CUDA(h, cudaIpcOpenMemHandle(&p, mu, cudaIpcMemLazyEnablePeerAccess));
            c->ipc_upper = (char*)p;
        }
    }
    if (const char* e = getenv("AB_HALO_TIMEOUT_S")) {
        const double s = atof(e);
        if (s > 0) {
            const unsigned long long ns = (unsigned long long)(s * 1e9);
            AB_CUDA(h, cudaMemcpyToSymbol(g_halo_timeout_ns, &ns, sizeof ns));
        }
    }
    c->ipc_on = true;
    return AB_OK;
}
int ab_comm_direct(const ab_handle* h) { return h && h->comm && h->comm->ipc_on ? 1 : 0; }

// Halo refresh outside a step (collective: every rank calls it): after ab_upload_local replaced the owned agents'
// columns from the host, the window's halo layers are rebuilt from the neighbours' boundary layers — index build on
// what the handle holds, then the same pack / transfer / unpack as inside a step.
int ab_comm_exchange(ab_handle* h) {
    if (!h || !h->comm || !h->win_W) return ab_fail(h, AB_ERR_STATE, "handle is not sharded");
    if (ab_comm_local(h)) return ab_fail(h, AB_ERR_STATE, "in-process slabs exchange inside ab_step_flock_group");
    AB_CUDA(h, cudaSetDevice(h->device));
    h->index_valid = false;
    h->unpack_rows = h->unpack_tb = false;
    h->halo_stale = false;             // set again below if this exchange times out as well
    AB_TRY(ab_ensure_index(h));
    AB_TRY(ab_comm_pack(h));
    AB_TRY(ab_comm_transfer(h));
    AB_TRY(ab_comm_unpack(h, false));
    h->agg_valid = false;
    h->pairs_n = -1;
    AB_TRY(ab_check_device_errors_now(h));
    return ab_comm_refresh_counts(h);
}

// Start alignment for timed multi-rank runs: queues a one-word all-reduce on the handle's stream, so whatever the
// caller queues next starts on every rank when the LAST rank has reached this point on the device (a host-side
// barrier leaves the ranks' launch skew inside the timed region).  No-op without a communicator.
int ab_comm_align(ab_handle* h) {
    if (!h || !h->comm || !h->win_W) return ab_fail(h, AB_ERR_STATE, "handle is not sharded");
    ab_comm* c = h->comm;
    if (!c->comm) return AB_OK;
    AB_CUDA(h, cudaSetDevice(h->device));
    AB_NCCL(h, g_nccl.AllReduce(c->d_align, c->d_align + 1, 1, 2 /* ncclInt32 */, 0 /* ncclSum */, c->comm, h->stream));
    return AB_OK;
}

// Time this rank's k_halo_unpack launches have spent waiting for the neighbours' "message delivered" flags since
// the last reset (direct exchange only): tells rank skew from work in a multi-GPU step time.
int ab_comm_wait_stats(ab_handle* h, double* wait_us, int64_t* exchanges, int reset) {
    if (!h || !h->comm || !h->win_W) return ab_fail(h, AB_ERR_STATE, "handle is not sharded");
    ab_comm* c = h->comm;
    AB_CUDA(h, cudaSetDevice(h->device));
    unsigned long long v[2] = {0, 0};
    AB_CUDA(h, cudaMemcpyAsync(v, c->d_wait, sizeof(v), cudaMemcpyDeviceToHost, h->stream));
    if (reset) AB_CUDA(h, cudaMemsetAsync(c->d_wait, 0, sizeof(v), h->stream));
    AB_CUDA(h, cudaStreamSynchronize(h->stream));
    if (wait_us) *wait_us = (double)v[0] * 1e-3;
    if (exchanges) *exchanges = (int64_t)v[1];
    return AB_OK;
}

int ab_comm_set_capacity(ab_handle* h, int64_t halo_agents) {
    if (!h || !h->comm) return ab_fail(h, AB_ERR_STATE, "handle is not sharded");
    if (halo_agents < 1 || halo_agents > 2000000000LL) return ab_fail(h, AB_ERR_INVALID, "bad halo capacity");
    h->comm->want_cap = halo_agents;
    return AB_OK;
}

int ab_comm_slab(const ab_handle* h, int64_t* first_layer, int64_t* n_layers) {
    if (!h || !h->win_W) return AB_ERR_STATE;
    if (first_layer) *first_layer = h->win_goff;
    if (n_layers) *n_layers = h->win_nl;
    return AB_OK;
}
int ab_comm_halo(const ab_handle* h) { return h ? h->win_W : 0; }

int64_t ab_nlocal(const ab_handle* h) {
    if (!h) return 0;
    return h->comm ? h->comm->n_own : h->n;
}

// Rows in slot order (row i -> slot i).  Field 0 (pos) defines the population: n agents with the given
// 1-based GLOBAL ids, all inside this rank's window; the other fields must follow with the same rows.
int ab_upload_local(ab_handle* h, int field, const void* host, const int64_t* global_ids, int64_t n) {
    if (!h || !h->win_W) return ab_fail(h, AB_ERR_STATE, "handle is not sharded");
    if (field < 0 || field >= (int)h->fields.size()) return ab_fail(h, AB_ERR_INVALID, "bad field id");
    if (n < 0 || n > h->cap) return ab_fail(h, AB_ERR_INVALID, "more rows than the handle's capacity (ab_resize first)");
    if (field != 0 && n != h->n) return ab_fail(h, AB_ERR_INVALID, "upload pos (field 0) first; other fields must have the same rows");
    if (n && (!host || (field == 0 && !global_ids))) return ab_fail(h, AB_ERR_INVALID, "null pointer");
    AB_CUDA(h, cudaSetDevice(h->device));
    ab_field& f = h->fields[field];
    if (field == 0) { h->n = n; h->comm->n_own = 0; }
    if (n == 0) return AB_OK;
    const size_t bytes = (size_t)n * f.ncomp * f.esize, idb = field == 0 ? (size_t)n * 8 : 0;
    char* stage;
    AB_TRY(ab_scratch(h, bytes + idb + 512, (void**)&stage));
    AB_CUDA(h, cudaMemcpyAsync(stage, host, bytes, cudaMemcpyHostToDevice, h->stream));
    {
        AB_LAUNCH(h, "k_rows_to_soa");
        k_rows_to_soa<<<ab_blocks(n * f.ncomp, 256), 256, 0, h->stream>>>(stage, (char*)f.buf[h->cur], h->cap, (int)n, f.ncomp, (int)f.esize);
    }
    if (field == 0) {
        char* ids_dev = stage + (bytes + 255) / 256 * 256;
        AB_CUDA(h, cudaMemcpyAsync(ids_dev, global_ids, idb, cudaMemcpyHostToDevice, h->stream));
        AB_LAUNCH(h, "k_gids_in");
        k_gids_in<<<ab_blocks(n, 256), 256, 0, h->stream>>>((const int64_t*)ids_dev, h->id[h->cur], (int)n, h->d_err);
        h->index_valid = false; h->pairs_n = -1;
    }
    if (field <= 1) h->agg_valid = false;                      // Σ(pos - centre) / Σ vel per cell are stale
    AB_CUDA(h, cudaStreamSynchronize(h->stream));
    return ab_check_device_errors(h);
}

// Owned agents (slot order) with their global ids.  host: cap rows of the field; ids_out: cap int64.
int ab_download_local(ab_handle* h, int field, void* host, int64_t* ids_out, int64_t cap, int64_t* n_out) {
    if (!h || !h->win_W) return ab_fail(h, AB_ERR_STATE, "handle is not sharded");
    if (field < 0 || field >= (int)h->fields.size()) return ab_fail(h, AB_ERR_INVALID, "bad field id");
    AB_CUDA(h, cudaSetDevice(h->device));
    AB_TRY(ab_comm_refresh_counts(h));
    const int64_t n = h->comm->n_own;
    if (n_out) *n_out = n;
    if (n > cap) return ab_fail(h, AB_ERR_INVALID, "host buffer smaller than the owned population");
    if (n == 0) return AB_OK;
    ab_field& f = h->fields[field];
    const size_t bytes = (size_t)n * f.ncomp * f.esize, idb = ids_out ? (size_t)n * 8 : 0;
    char* stage;
    AB_TRY(ab_scratch(h, bytes + idb + 512, (void**)&stage));
    {
        AB_LAUNCH(h, "k_soa_to_rows");
        k_soa_to_rows<<<ab_blocks(n * f.ncomp, 256), 256, 0, h->stream>>>((const char*)f.buf[h->cur], stage, h->cap, (int)n, f.ncomp, (int)f.esize);
    }
    if (host) AB_CUDA(h, cudaMemcpyAsync(host, stage, bytes, cudaMemcpyDeviceToHost, h->stream));
    if (ids_out) {
        char* ids_dev = stage + (bytes + 255) / 256 * 256;
        AB_LAUNCH(h, "k_gids_out");
        k_gids_out<<<ab_blocks(n, 256), 256, 0, h->stream>>>(h->id[h->cur], (int64_t*)ids_dev, (int)n);
        AB_CUDA(h, cudaMemcpyAsync(ids_out, ids_dev, idb, cudaMemcpyDeviceToHost, h->stream));
    }
    AB_CUDA(h, cudaStreamSynchronize(h->stream));
    return AB_OK;
}

}  // extern "C"
