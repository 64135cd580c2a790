// ab_collect.cu — on-device aggregates of agent columns for data collection.
//
// run!(model, n; adata = [(field, aggregator[, filter]), ...]) calls collect_agent_data! after every
// collected tick (src/simulations/collect.jl:154-183); the tuple form evaluates
// aggregator(get_data(a, field) for a in allagents(model) if filter(a))  (collect.jl:466-502, 520-560).
// With device-resident columns that would mean a full column download per tick; here each
// (field, aggregator, filter) triple is reduced on the device and only one double (+ a count) per triple
// crosses PCIe.  Supported aggregators: sum, mean, minimum, maximum, count; filters: one comparison of a
// column component with a constant (the common `a -> a.field == v` / `>` / `<` forms).
//
// The reduction order is fixed (slot-strided per thread, tree per block, partials in block order), so a
// result is reproducible for a given state; it is NOT the reference's left fold in id order, so
// floating-point sums agree with it to rounding (tests: 1e-12 relative), integer / min / max / count
// results exactly.  On a sharded handle the aggregate covers the OWNED agents of the rank (the host adds
// the ranks' partial results).
#include <algorithm>
#include <mutex>

#include "ab_internal.cuh"

#define RED_THREADS 256
#define RED_ITEMS 16
#define RED_CHUNK (RED_THREADS * RED_ITEMS)

struct RedSpecDev {
    const void* col;      // column component base (element 0 of that component)
    int dtype;
    const void* fcol;     // filter column component base or nullptr
    int fdtype, fcmp;
    double fval;
    int op;
};

__device__ __forceinline__ double red_load(const void* p, int dtype, int s) {
    switch (dtype) {
        case AB_F64: return ((const double*)p)[s];
        case AB_F32: return (double)((const float*)p)[s];
        case AB_I32: return (double)((const int32_t*)p)[s];
        case AB_I64: return (double)((const long long*)p)[s];
        default: return (double)((const uint8_t*)p)[s];
    }
}
__device__ __forceinline__ bool red_pass(const RedSpecDev& sp, int s) {
    if (!sp.fcol) return true;
    const double v = red_load(sp.fcol, sp.fdtype, s);
    switch (sp.fcmp) {
        case AB_CMP_EQ: return v == sp.fval;
        case AB_CMP_NE: return v != sp.fval;
        case AB_CMP_LT: return v < sp.fval;
        case AB_CMP_LE: return v <= sp.fval;
        case AB_CMP_GT: return v > sp.fval;
        case AB_CMP_GE: return v >= sp.fval;
    }
    return true;
}
__device__ __forceinline__ double red_combine(int op, double a, double b) {
    if (op == AB_RED_MIN) return fmin(a, b);
    if (op == AB_RED_MAX) return fmax(a, b);
    return a + b;
}
__device__ __forceinline__ double red_identity(int op) {
    if (op == AB_RED_MIN) return INFINITY;
    if (op == AB_RED_MAX) return -INFINITY;
    return 0.0;
}

// blockIdx.y = spec; blockIdx.x = chunk of RED_CHUNK slots.  partial[spec][chunk] = {value, count}
__global__ void __launch_bounds__(RED_THREADS)
k_reduce_chunks(const RedSpecDev* __restrict__ specs, int n, int nchunks, double* __restrict__ pval,
                unsigned long long* __restrict__ pcnt) {
    __shared__ double s_v[RED_THREADS / 32];
    __shared__ unsigned long long s_c[RED_THREADS / 32];
    const RedSpecDev sp = specs[blockIdx.y];
    const int base = blockIdx.x * RED_CHUNK + threadIdx.x;
    double acc = red_identity(sp.op);
    unsigned long long cnt = 0;
#pragma unroll 4
    for (int i = 0; i < RED_ITEMS; ++i) {
        const int s = base + i * RED_THREADS;
        if (s < n && red_pass(sp, s)) {
            ++cnt;
            if (sp.op != AB_RED_COUNT) acc = red_combine(sp.op, acc, red_load(sp.col, sp.dtype, s));
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        acc = red_combine(sp.op, acc, __shfl_xor_sync(0xffffffffu, acc, o));
        cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    }
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (lane == 0) { s_v[wid] = acc; s_c[wid] = cnt; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double a = s_v[0];
        unsigned long long c = s_c[0];
        for (int w = 1; w < RED_THREADS / 32; ++w) { a = red_combine(sp.op, a, s_v[w]); c += s_c[w]; }
        pval[(size_t)blockIdx.y * nchunks + blockIdx.x] = a;
        pcnt[(size_t)blockIdx.y * nchunks + blockIdx.x] = c;
    }
}

// one block per spec: partials in chunk order
__global__ void __launch_bounds__(RED_THREADS)
k_reduce_final(const RedSpecDev* __restrict__ specs, int nchunks, const double* __restrict__ pval,
               const unsigned long long* __restrict__ pcnt, double* __restrict__ out, long long* __restrict__ ocnt) {
    __shared__ double s_v[RED_THREADS];
    __shared__ unsigned long long s_c[RED_THREADS];
    const RedSpecDev sp = specs[blockIdx.x];
    double acc = red_identity(sp.op);
    unsigned long long cnt = 0;
    for (int i = threadIdx.x; i < nchunks; i += RED_THREADS) {
        acc = red_combine(sp.op, acc, pval[(size_t)blockIdx.x * nchunks + i]);
        cnt += pcnt[(size_t)blockIdx.x * nchunks + i];
    }
    s_v[threadIdx.x] = acc;
    s_c[threadIdx.x] = cnt;
    __syncthreads();
    for (int o = RED_THREADS / 2; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o) {
            s_v[threadIdx.x] = red_combine(sp.op, s_v[threadIdx.x], s_v[threadIdx.x + o]);
            s_c[threadIdx.x] += s_c[threadIdx.x + o];
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        double v = s_v[0];
        const unsigned long long c = s_c[0];
        if (sp.op == AB_RED_COUNT) v = (double)c;
        else if (sp.op == AB_RED_MEAN) v = c ? v / (double)c : __longlong_as_double(0x7ff8000000000000LL);   // mean of nothing: NaN
        out[blockIdx.x] = v;
        ocnt[blockIdx.x] = (long long)c;
    }
}

extern "C" int ab_reduce(ab_handle* h, const ab_reduce_spec* specs, int nspec, double* out, int64_t* counts) {
    if (!h || !specs || !out || nspec < 1 || nspec > 64) return ab_fail(h, AB_ERR_INVALID, "ab_reduce: 1..64 specs and an output array");
    AB_CUDA(h, cudaSetDevice(h->device));
    if (h->grid_cols_dirty) AB_TRY(ab_grid_sync_columns(h));
    int64_t n = h->n;
    if (h->win_W) { AB_TRY(ab_comm_refresh_counts(h)); n = ab_nlocal(h); }     // owned agents come first
    std::vector<RedSpecDev> dev(nspec);
    for (int i = 0; i < nspec; ++i) {
        const ab_reduce_spec& s = specs[i];
        if (s.op < AB_RED_SUM || s.op > AB_RED_COUNT) return ab_fail(h, AB_ERR_INVALID, "ab_reduce: unknown aggregator");
        if (s.field < 0 || s.field >= (int)h->fields.size() || s.comp < 0 || s.comp >= h->fields[s.field].ncomp)
            return ab_fail(h, AB_ERR_INVALID, "ab_reduce: bad field / component");
        const ab_field& f = h->fields[s.field];
        dev[i].col = (const char*)f.buf[h->cur] + (size_t)s.comp * h->cap * f.esize;
        dev[i].dtype = f.dtype;
        dev[i].op = s.op;
        dev[i].fcol = nullptr; dev[i].fdtype = 0; dev[i].fcmp = AB_CMP_NONE; dev[i].fval = s.filter_value;
        if (s.filter_cmp != AB_CMP_NONE) {
            if (s.filter_cmp < AB_CMP_EQ || s.filter_cmp > AB_CMP_GE) return ab_fail(h, AB_ERR_INVALID, "ab_reduce: unknown filter comparison");
            if (s.filter_field < 0 || s.filter_field >= (int)h->fields.size() || s.filter_comp < 0 ||
                s.filter_comp >= h->fields[s.filter_field].ncomp)
                return ab_fail(h, AB_ERR_INVALID, "ab_reduce: bad filter field / component");
            const ab_field& ff = h->fields[s.filter_field];
            dev[i].fcol = (const char*)ff.buf[h->cur] + (size_t)s.filter_comp * h->cap * ff.esize;
            dev[i].fdtype = ff.dtype;
            dev[i].fcmp = s.filter_cmp;
        }
    }
    const int nchunks = (int)std::max<int64_t>((n + RED_CHUNK - 1) / RED_CHUNK, 1);
    const size_t o_spec = 0, o_pval = (sizeof(RedSpecDev) * nspec + 255) / 256 * 256;
    const size_t o_pcnt = o_pval + (size_t)nspec * nchunks * 8, o_out = o_pcnt + (size_t)nspec * nchunks * 8;
    const size_t o_cnt = o_out + (size_t)nspec * 8, total = o_cnt + (size_t)nspec * 8;
    char* base;
    AB_TRY(ab_scratch2(h, total + 256, (void**)&base));
    AB_CUDA(h, cudaMemcpyAsync(base + o_spec, dev.data(), sizeof(RedSpecDev) * nspec, cudaMemcpyHostToDevice, h->stream));
    {
        AB_LAUNCH(h, "k_reduce_chunks");
        k_reduce_chunks<<<dim3(nchunks, nspec), RED_THREADS, 0, h->stream>>>((const RedSpecDev*)(base + o_spec), (int)n, nchunks,
                                                                           (double*)(base + o_pval), (unsigned long long*)(base + o_pcnt));
    }
    {
        AB_LAUNCH(h, "k_reduce_final");
        k_reduce_final<<<nspec, RED_THREADS, 0, h->stream>>>((const RedSpecDev*)(base + o_spec), nchunks, (const double*)(base + o_pval),
                                                            (const unsigned long long*)(base + o_pcnt), (double*)(base + o_out),
                                                            (long long*)(base + o_cnt));
    }
    std::vector<long long> cnt(nspec);
    AB_CUDA(h, cudaMemcpyAsync(out, base + o_out, (size_t)nspec * 8, cudaMemcpyDeviceToHost, h->stream));
    AB_CUDA(h, cudaMemcpyAsync(cnt.data(), base + o_cnt, (size_t)nspec * 8, cudaMemcpyDeviceToHost, h->stream));
    AB_CUDA(h, cudaStreamSynchronize(h->stream));
    if (counts) for (int i = 0; i < nspec; ++i) counts[i] = cnt[i];
    return AB_OK;
}

// ---------------------------------------------------------------------------------------------------
// Schedulers.ByProperty(field)(model) — src/submodules/schedulers.jl:130-151: the agent ids in `sortperm` order of
// a property column (ascending, stable: ties keep id order).  Under the synchronous update an ordering of ALL agents
// changes nothing, so this is a helper for host-side loops: the sort runs on the device — a bitonic network over
// (sortable key, id) pairs, the id in the low word making the order total and therefore stable — and only the n
// sorted ids cross PCIe.
// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned long long sort_key_f64(double v) {
    if (v != v) return 0xFFFFFFFFFFFFFFF0ull;                      // NaN last (isless), below the padding
    unsigned long long b = (unsigned long long)__double_as_longlong(v);
    return (b & 0x8000000000000000ull) ? ~b : (b | 0x8000000000000000ull);
}
__global__ void k_sort_fill(const void* __restrict__ col, int dtype, const int32_t* __restrict__ id, int n, int npad,
                            ulonglong2* __restrict__ P) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= npad) return;
    if (s >= n) { P[s] = make_ulonglong2(~0ull, ~0ull); return; }  // padding sorts to the end
    const int i = id ? id[s] : s;                                  // grids keep their columns in id order
    unsigned long long k;
    if (dtype == AB_I64) k = (unsigned long long)(((const long long*)col)[s]) ^ 0x8000000000000000ull;
    else k = sort_key_f64(red_load(col, dtype, s));                // f64 / f32 / i32 / u8: exact in a double
    P[s] = make_ulonglong2(k, (unsigned long long)i);
}
__global__ void k_sort_step(ulonglong2* __restrict__ P, int npad, int k, int j) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= npad) return;
    const int l = i ^ j;
    if (l <= i) return;
    const ulonglong2 a = P[i], b = P[l];
    const bool a_gt_b = a.x > b.x || (a.x == b.x && a.y > b.y);
    const bool up = (i & k) == 0;
    if (a_gt_b == up) { P[i] = b; P[l] = a; }
}
__global__ void k_sort_out(const ulonglong2* __restrict__ P, int n, int descending, int64_t* __restrict__ out) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    out[descending ? n - 1 - s : s] = (int64_t)P[s].y + 1;
}

extern "C" int ab_ids_by_property(ab_handle* h, int field, int comp, int descending, int64_t* out_ids, int64_t n) {
    if (!h || field < 0 || field >= (int)h->fields.size()) return ab_fail(h, AB_ERR_INVALID, "bad field id");
    if (h->win_W) return ab_fail(h, AB_ERR_STATE, "not available on a sharded handle");
    if (n != h->n) return ab_fail(h, AB_ERR_INVALID, "length must equal the agent count");
    const ab_field& f = h->fields[field];
    if (comp < 0 || comp >= f.ncomp) return ab_fail(h, AB_ERR_INVALID, "bad component");
    if (n == 0) return AB_OK;
    if (!out_ids) return ab_fail(h, AB_ERR_INVALID, "null output");
    if (n > ((int64_t)1 << 30)) return ab_fail(h, AB_ERR_INVALID, "ab_ids_by_property sorts at most 2^30 agents (32-bit padded network)");
    AB_CUDA(h, cudaSetDevice(h->device));
    if (h->grid_cols_dirty) AB_TRY(ab_grid_sync_columns(h));
    int npad = 1;
    while (npad < n) npad <<= 1;
    char* base;
    AB_TRY(ab_scratch2(h, (size_t)npad * sizeof(ulonglong2) + (size_t)n * sizeof(int64_t) + 256, (void**)&base));
    ulonglong2* P = (ulonglong2*)base;
    int64_t* dout = (int64_t*)(base + (size_t)npad * sizeof(ulonglong2));
    const char* col = (const char*)f.buf[h->cur] + (size_t)comp * h->cap * f.esize;
    const bool cont = h->desc.kind == AB_SPACE_CONTINUOUS;         // continuous handles keep columns in slot order
    {
        AB_LAUNCH(h, "k_sort_fill");
        k_sort_fill<<<ab_blocks(npad, 256), 256, 0, h->stream>>>(col, f.dtype, cont ? h->id[h->cur] : nullptr, (int)n, npad, P);
    }
    for (int k = 2; k <= npad; k <<= 1)
        for (int j = k >> 1; j > 0; j >>= 1) {
            AB_LAUNCH(h, "k_sort_step");
            k_sort_step<<<ab_blocks(npad, 256), 256, 0, h->stream>>>(P, npad, k, j);
        }
    {
        AB_LAUNCH(h, "k_sort_out");
        k_sort_out<<<ab_blocks(n, 256), 256, 0, h->stream>>>(P, (int)n, descending, dout);
    }
    AB_CUDA(h, cudaMemcpyAsync(out_ids, dout, (size_t)n * sizeof(int64_t), cudaMemcpyDeviceToHost, h->stream));
    return ab_check_device_errors_now(h);
}

// ---------------------------------------------------------------------------------------------------
// ab_reduce_static — the chunked reduction above for callers that must not allocate device memory.
//
// Who that is: a Julia session whose CUDA.jl pool (or a Python process whose torch caching allocator)
// already owns nearly all of the GPU, so that a cudaMalloc made behind the pool's back can fail although
// the pool still has free blocks; an ensemble driver with several models in flight, where the implicit
// device-wide synchronisation of cudaMalloc / cudaFree would stall every other model's stream; a process
// confined to a fixed memory budget (a MIG slice, an MPS client with a memory limit).  Such a caller
// holds its columns in memory it got from its own pool and cannot create an ab_handle (ab_create and the
// handle's grow-only scratch allocate).  Here it passes device pointers for the column and for the two
// results; the per-chunk partial results — the only scratch this reduction needs — live in the array
// below, which is part of the library image: every GPU that loads libagents_b200.so carries these
// 256 KiB from the moment the module is loaded until the process exits, whether or not this entry point
// is ever called.
//
// The array is shared by every call on a device.  It keeps whatever the previous call left there; that
// is harmless because the first kernel writes partial[0 .. nchunks) and the second reads exactly that
// range — nothing has to start at zero, nothing is cleared.  Two calls whose kernels can run at the same
// time on one GPU (different streams, or different host threads using different streams) would write
// the same partials: the caller has to order them (same stream, or an event between them).
// ---------------------------------------------------------------------------------------------------
#define AB_STATIC_MAX_N (1 << 26)                                   // 67 108 864 values per call
#define AB_STATIC_MAX_CHUNKS (AB_STATIC_MAX_N / RED_CHUNK)          // 16 384 partials
#define AB_STATIC_BYTES ((size_t)AB_STATIC_MAX_CHUNKS * 16)         // value + count per partial: 256 KiB
#define AB_STATIC_MAX_DEVICES 64

static __device__ __align__(256) unsigned char g_reduce_static_scratch[AB_STATIC_BYTES];

// In host code the name g_reduce_static_scratch is only a token the runtime maps to the device copies; its
// host-side address is not device memory.  The address a kernel may use is asked for once per device, with
// that device current, and remembered here.
static void* g_static_addr[AB_STATIC_MAX_DEVICES];
static std::mutex g_static_mu;

static int static_scratch_address(int device, void** out) {
    std::lock_guard<std::mutex> lock(g_static_mu);
    if (!g_static_addr[device]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_reduce_static_scratch);   // the current device's copy
        if (e != cudaSuccess) return ab_cuda_fail(nullptr, e, "cudaGetSymbolAddress(g_reduce_static_scratch)");
        if (!p) return ab_fail(nullptr, AB_ERR_CUDA, "ab_reduce_static: the scratch array has no device address");
        g_static_addr[device] = p;
    }
    *out = g_static_addr[device];
    return AB_OK;
}

// k_reduce_chunks for one triple handed over as a kernel argument (no spec table in device memory)
__global__ void __launch_bounds__(RED_THREADS)
k_reduce_chunks_static(const RedSpecDev sp, int n, double* __restrict__ pval, unsigned long long* __restrict__ pcnt) {
    __shared__ double s_v[RED_THREADS / 32];
    __shared__ unsigned long long s_c[RED_THREADS / 32];
    const int base = blockIdx.x * RED_CHUNK + threadIdx.x;
    double acc = red_identity(sp.op);
    unsigned long long cnt = 0;
#pragma unroll 4
    for (int i = 0; i < RED_ITEMS; ++i) {
        const int s = base + i * RED_THREADS;
        if (s < n && red_pass(sp, s)) {
            ++cnt;
            if (sp.op != AB_RED_COUNT) acc = red_combine(sp.op, acc, red_load(sp.col, sp.dtype, s));
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        acc = red_combine(sp.op, acc, __shfl_xor_sync(0xffffffffu, acc, o));
        cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    }
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (lane == 0) { s_v[wid] = acc; s_c[wid] = cnt; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double a = s_v[0];
        unsigned long long c = s_c[0];
        for (int w = 1; w < RED_THREADS / 32; ++w) { a = red_combine(sp.op, a, s_v[w]); c += s_c[w]; }
        pval[blockIdx.x] = a;
        pcnt[blockIdx.x] = c;
    }
}

// k_reduce_final for that triple: one block, partials in chunk order, results to the caller's device words
__global__ void __launch_bounds__(RED_THREADS)
k_reduce_final_static(int op, int nchunks, const double* __restrict__ pval, const unsigned long long* __restrict__ pcnt,
                      double* __restrict__ out, long long* __restrict__ ocnt) {
    __shared__ double s_v[RED_THREADS];
    __shared__ unsigned long long s_c[RED_THREADS];
    double acc = red_identity(op);
    unsigned long long cnt = 0;
    for (int i = threadIdx.x; i < nchunks; i += RED_THREADS) {
        acc = red_combine(op, acc, pval[i]);
        cnt += pcnt[i];
    }
    s_v[threadIdx.x] = acc;
    s_c[threadIdx.x] = cnt;
    __syncthreads();
    for (int o = RED_THREADS / 2; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o) {
            s_v[threadIdx.x] = red_combine(op, s_v[threadIdx.x], s_v[threadIdx.x + o]);
            s_c[threadIdx.x] += s_c[threadIdx.x + o];
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        double v = s_v[0];
        const unsigned long long c = s_c[0];
        if (op == AB_RED_COUNT) v = (double)c;
        else if (op == AB_RED_MEAN) v = c ? v / (double)c : __longlong_as_double(0x7ff8000000000000LL);
        *out = v;
        if (ocnt) *ocnt = (long long)c;
    }
}

// Everything is queued on `stream` (NULL: the default stream) and the call returns without waiting: *d_out and
// *d_count are valid once the stream has reached this point.  The first call on a device looks up the device
// address of the scratch array (see above) — make that first call before starting any stream capture.  No
// cudaMalloc, cudaFree or other allocating routine is called; messages of failures are read with
// ab_last_error(NULL).
extern "C" int ab_reduce_static(int device, void* stream, const void* d_col, int dtype, int64_t n, int op,
                                const void* d_filter_col, int filter_dtype, int filter_cmp, double filter_value,
                                double* d_out, int64_t* d_count) {
    if (!d_out) return ab_fail(nullptr, AB_ERR_INVALID, "ab_reduce_static: d_out is NULL");
    if (device < 0 || device >= AB_STATIC_MAX_DEVICES) return ab_fail(nullptr, AB_ERR_INVALID, "ab_reduce_static: device must be 0..63");
    if (n < 0) return ab_fail(nullptr, AB_ERR_INVALID, "ab_reduce_static: negative length");
    if (n > AB_STATIC_MAX_N)
        return ab_fail(nullptr, AB_ERR_INVALID, "ab_reduce_static: " + std::to_string(n) + " values exceed the fixed scratch of this entry point (" +
                       std::to_string((long long)AB_STATIC_MAX_N) + " values, 2^26); split the column or use ab_reduce on a handle");
    if (op < AB_RED_SUM || op > AB_RED_COUNT) return ab_fail(nullptr, AB_ERR_INVALID, "ab_reduce_static: unknown aggregator");
    if (dtype < AB_F64 || dtype > AB_U8) return ab_fail(nullptr, AB_ERR_INVALID, "ab_reduce_static: unknown column dtype");
    if (n > 0 && !d_col && op != AB_RED_COUNT) return ab_fail(nullptr, AB_ERR_INVALID, "ab_reduce_static: d_col is NULL");
    if (filter_cmp != AB_CMP_NONE) {
        if (filter_cmp < AB_CMP_EQ || filter_cmp > AB_CMP_GE) return ab_fail(nullptr, AB_ERR_INVALID, "ab_reduce_static: unknown filter comparison");
        if (filter_dtype < AB_F64 || filter_dtype > AB_U8) return ab_fail(nullptr, AB_ERR_INVALID, "ab_reduce_static: unknown filter dtype");
        if (n > 0 && !d_filter_col) return ab_fail(nullptr, AB_ERR_INVALID, "ab_reduce_static: filter comparison without a filter column");
    }
    int prev = -1;
    AB_CUDA(nullptr, cudaGetDevice(&prev));
    if (prev != device) AB_CUDA(nullptr, cudaSetDevice(device));
    void* scratch = nullptr;
    int rc = static_scratch_address(device, &scratch);
    if (rc == AB_OK) {
        RedSpecDev sp;
        sp.col = d_col; sp.dtype = dtype; sp.op = op;
        sp.fcol = filter_cmp != AB_CMP_NONE ? d_filter_col : nullptr;
        sp.fdtype = filter_dtype; sp.fcmp = filter_cmp; sp.fval = filter_value;
        const int nchunks = (int)std::max<int64_t>((n + RED_CHUNK - 1) / RED_CHUNK, 1);
        double* pval = (double*)scratch;
        unsigned long long* pcnt = (unsigned long long*)((char*)scratch + (size_t)AB_STATIC_MAX_CHUNKS * 8);
        const cudaStream_t st = (cudaStream_t)stream;
        k_reduce_chunks_static<<<nchunks, RED_THREADS, 0, st>>>(sp, (int)n, pval, pcnt);
        k_reduce_final_static<<<1, RED_THREADS, 0, st>>>(op, nchunks, pval, pcnt, d_out, (long long*)d_count);
        const cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) rc = ab_cuda_fail(nullptr, e, "ab_reduce_static launches");
    }
    if (prev != device) cudaSetDevice(prev);
    return rc;
}
