// ab_field.cu — spatial fields over a ContinuousSpace:
//   get_spatial_index / get_spatial_property  (src/spaces/continuous.jl:536-573) for all agents
//   (or for a batch of query positions), and the force-field form of the `update_vel!` hook
//   (continuous.jl:86-91, :186-194): vel += scale * field[get_spatial_index(pos)] before a move.
// The reference's index rule (:566-573):  εs = spacesize[1:Dp] ./ size(property);
//   idxs = floor.(Int, pos[1:Dp] ./ εs) .+ 1  — property may have fewer dimensions than the space
// (the front dimensions of pos are used); an index outside the array is the reference's BoundsError
// (pos == extent on a non-periodic axis) and is reported as AB_ERR_OUTSIDE.
#include "ab_internal.cuh"

struct FieldDims {
    int pd;                 // dimensions of the property array (1..D)
    int64_t size[3];
    double eps[3];          // spacesize[d] / size[d], computed in the space's element type like the reference
};

template <class T, int D>
__device__ __forceinline__ bool field_index(const FieldDims& fd, const T* p, int64_t* idx) {
    bool ok = true;
#pragma unroll
    for (int d = 0; d < D; ++d) {
        if (d >= fd.pd) break;
        const T q = p[d] / (T)fd.eps[d];
        const double fl = floor((double)q);
        if (!(fl >= 0.0 && fl < (double)fd.size[d])) { ok = false; idx[d] = 1; }
        else idx[d] = (int64_t)fl + 1;
    }
    return ok;
}

// out_idx: nq x pd (1-based, row per query) and/or out_val: nq x ncomp gathered from the column-major property
template <class T, int D>
__global__ void k_field_lookup(FieldDims fd, const T* __restrict__ pos, int64_t cap, const int32_t* __restrict__ id,
                               const T* __restrict__ qpos, int nq, const double* __restrict__ prop, int ncomp,
                               int64_t* __restrict__ out_idx, double* __restrict__ out_val, int* __restrict__ err) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= nq) return;
    T p[D];
    int64_t row = t;
    if (qpos) {
#pragma unroll
        for (int d = 0; d < D; ++d) p[d] = qpos[(size_t)t * D + d];
    } else {
#pragma unroll
        for (int d = 0; d < D; ++d) p[d] = pos[(size_t)d * cap + t];
        row = id[t];                                       // slot -> id order
    }
    int64_t idx[3] = {1, 1, 1};
    if (!field_index<T, D>(fd, p, idx)) atomicOr(err, AB_EF_OUTSIDE);
    if (out_idx)
        for (int d = 0; d < fd.pd; ++d) out_idx[(size_t)row * fd.pd + d] = idx[d];
    if (out_val) {
        int64_t lin = 0, total = 1;
        for (int d = fd.pd - 1; d >= 0; --d) lin = lin * fd.size[d] + (idx[d] - 1);
        for (int d = 0; d < fd.pd; ++d) total *= fd.size[d];
        for (int c = 0; c < ncomp; ++c) out_val[(size_t)row * ncomp + c] = prop[(size_t)c * total + lin];
    }
}

// vel[d] += scale * field[d][index(pos)] for the active agents (all when mask == nullptr)
template <class T, int D>
__global__ void k_field_kick(FieldDims fd, const T* __restrict__ pos, T* __restrict__ vel, int64_t cap, int n,
                             const double* __restrict__ field, double scale, const int32_t* __restrict__ id,
                             const uint8_t* __restrict__ active, int* __restrict__ err) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    if (active && !active[id[s]]) return;
    T p[D];
#pragma unroll
    for (int d = 0; d < D; ++d) p[d] = pos[(size_t)d * cap + s];
    int64_t idx[3] = {1, 1, 1};
    if (!field_index<T, D>(fd, p, idx)) { atomicOr(err, AB_EF_OUTSIDE); return; }
    int64_t lin = 0, total = 1;
    for (int d = fd.pd - 1; d >= 0; --d) lin = lin * fd.size[d] + (idx[d] - 1);
    for (int d = 0; d < fd.pd; ++d) total *= fd.size[d];
#pragma unroll
    for (int d = 0; d < D; ++d)
        vel[(size_t)d * cap + s] = (T)((double)vel[(size_t)d * cap + s] + scale * field[(size_t)d * total + lin]);
}

static int make_dims(ab_handle* h, const int64_t* dims, int pd, FieldDims* fd, int64_t* total) {
    if (h->desc.kind != AB_SPACE_CONTINUOUS) return ab_fail(h, AB_ERR_INVALID, "spatial fields need a ContinuousSpace handle");
    if (h->win_W) return ab_fail(h, AB_ERR_STATE, "not available on a sharded handle");
    if (!dims || pd < 1 || pd > h->D) return ab_fail(h, AB_ERR_INVALID, "property must have between 1 and D dimensions");
    fd->pd = pd;
    *total = 1;
    for (int d = 0; d < 3; ++d) { fd->size[d] = 1; fd->eps[d] = 1.0; }
    for (int d = 0; d < pd; ++d) {
        if (dims[d] < 1) return ab_fail(h, AB_ERR_INVALID, "empty property dimension");
        fd->size[d] = dims[d];
        // usize ./ propertysize in the space's element type (continuous.jl:570)
        fd->eps[d] = h->dtype == AB_F64 ? h->desc.extent[d] / (double)dims[d] : (double)((float)h->desc.extent[d] / (float)dims[d]);
        *total *= dims[d];
    }
    return AB_OK;
}

extern "C" {

int ab_spatial_lookup(ab_handle* h, const int64_t* prop_dims, int pd, const double* property, int ncomp,
                      const void* qpos, int64_t nq, int64_t* out_index, double* out_value) {
    FieldDims fd;
    int64_t total;
    AB_TRY(make_dims(h, prop_dims, pd, &fd, &total));
    if (out_value && (!property || ncomp < 1)) return ab_fail(h, AB_ERR_INVALID, "property values requested without a property array");
    if (!out_index && !out_value) return ab_fail(h, AB_ERR_INVALID, "nothing to compute");
    if (!qpos) nq = h->n;
    if (nq < 0) return ab_fail(h, AB_ERR_INVALID, "negative query count");
    if (nq == 0) return AB_OK;
    AB_CUDA(h, cudaSetDevice(h->device));
    auto up = [](size_t b) { return (b + 255) / 256 * 256; };
    const size_t bq = qpos ? (size_t)nq * h->D * h->tsize : 0, bp = out_value ? (size_t)total * ncomp * 8 : 0;
    const size_t bi = out_index ? (size_t)nq * pd * 8 : 0, bv = out_value ? (size_t)nq * ncomp * 8 : 0;
    char* dev;
    AB_TRY(ab_scratch(h, up(bq) + up(bp) + up(bi) + up(bv) + 256, (void**)&dev));
    char* dq = dev;
    double* dp = (double*)(dev + up(bq));
    int64_t* di = (int64_t*)(dev + up(bq) + up(bp));
    double* dv = (double*)(dev + up(bq) + up(bp) + up(bi));
    if (qpos) AB_CUDA(h, cudaMemcpyAsync(dq, qpos, bq, cudaMemcpyHostToDevice, h->stream));
    if (out_value) AB_CUDA(h, cudaMemcpyAsync(dp, property, bp, cudaMemcpyHostToDevice, h->stream));
    {
        AB_LAUNCH(h, "k_field_lookup");
        const int cur = h->cur, nb = ab_blocks(nq, 256);
#define FL(T, DD) k_field_lookup<T, DD><<<nb, 256, 0, h->stream>>>(fd, (const T*)h->fields[0].buf[cur], h->cap, h->id[cur], qpos ? (const T*)dq : nullptr, (int)nq, out_value ? dp : nullptr, ncomp, out_index ? di : nullptr, out_value ? dv : nullptr, h->d_err)
        if (h->dtype == AB_F64) { if (h->D == 2) FL(double, 2); else FL(double, 3); }
        else { if (h->D == 2) FL(float, 2); else FL(float, 3); }
#undef FL
    }
    if (out_index) AB_CUDA(h, cudaMemcpyAsync(out_index, di, bi, cudaMemcpyDeviceToHost, h->stream));
    if (out_value) AB_CUDA(h, cudaMemcpyAsync(out_value, dv, bv, cudaMemcpyDeviceToHost, h->stream));
    return ab_check_device_errors_now(h);                 // synchronises; BoundsError of the reference -> AB_ERR_OUTSIDE
}

int ab_update_vel_field(ab_handle* h, const int64_t* prop_dims, int pd, const double* field, double scale,
                        const int64_t* active_ids, int64_t nactive) {
    FieldDims fd;
    int64_t total;
    AB_TRY(make_dims(h, prop_dims, pd, &fd, &total));
    if (!field) return ab_fail(h, AB_ERR_INVALID, "null field");
    if (h->n == 0) return AB_OK;
    AB_CUDA(h, cudaSetDevice(h->device));
    const size_t bf = (size_t)total * h->D * 8, bm = active_ids ? (size_t)h->n : 0;
    char* dev;
    AB_TRY(ab_scratch(h, (bf + 255) / 256 * 256 + bm + 1024, (void**)&dev));
    double* dfield = (double*)dev;
    uint8_t* dmask = active_ids ? (uint8_t*)(dev + (bf + 255) / 256 * 256) : nullptr;
    AB_CUDA(h, cudaMemcpyAsync(dfield, field, bf, cudaMemcpyHostToDevice, h->stream));
    std::vector<uint8_t> mask;
    if (active_ids) {
        mask.assign((size_t)h->n, 0);
        for (int64_t k = 0; k < nactive; ++k) {
            if (active_ids[k] < 1 || active_ids[k] > h->n) return ab_fail(h, AB_ERR_INVALID, "active id out of range");
            mask[(size_t)active_ids[k] - 1] = 1;
        }
        AB_CUDA(h, cudaMemcpyAsync(dmask, mask.data(), bm, cudaMemcpyHostToDevice, h->stream));
    }
    {
        AB_LAUNCH(h, "k_field_kick");
        const int cur = h->cur, nb = ab_blocks(h->n, 256);
#define FK(T, DD) k_field_kick<T, DD><<<nb, 256, 0, h->stream>>>(fd, (const T*)h->fields[0].buf[cur], (T*)h->fields[1].buf[cur], h->cap, (int)h->n, dfield, scale, h->id[cur], dmask, h->d_err)
        if (h->dtype == AB_F64) { if (h->D == 2) FK(double, 2); else FK(double, 3); }
        else { if (h->D == 2) FK(float, 2); else FK(float, 3); }
#undef FK
    }
    h->agg_valid = false;                                 // Σ vel per cell changed
    return ab_check_device_errors_now(h);                 // also keeps `mask` alive until the copy is done
}

}  // extern "C"
