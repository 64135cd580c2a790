// ab_collide.cu — elastic_collision!(a, b, f) (src/spaces/continuous.jl:333-366) for a list of
// agent pairs, the consumer of interacting_pairs(model, r, :nearest) in the social-distancing /
// billiard model family (test/continuous_space_tests.jl:382-460).
//
// The reference resolves the pairs one after the other; a pair list from true_pairs! is almost
// disjoint, but an id may survive in two pairs (continuous.jl:481-496 removes one entry per
// doubled id).  The list is therefore cut into ROUNDS on the host — round(k) = 1 + the latest
// round of an earlier pair sharing an agent with pair k — and every round is one launch, one
// thread per pair: the result equals the sequential loop whatever the list looks like.
#include <algorithm>
#include <unordered_map>

#include "ab_internal.cuh"

enum { EC_ASSERT = 64 };     // device error bit: ∞ mass with nonzero velocity (continuous.jl:350, :355)

template <class T, class M>
__global__ void k_elastic(const int64_t* __restrict__ pa, const int64_t* __restrict__ pb, const int32_t* __restrict__ pidx,
                          int np, const int32_t* __restrict__ slot_of_id, const T* __restrict__ pos, T* __restrict__ vel,
                          const M* __restrict__ mass, int64_t cap, uint8_t* __restrict__ happened, int* __restrict__ err) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= np) return;
    const int sa = slot_of_id[pa[t] - 1], sb = slot_of_id[pb[t] - 1];
    const double x1[2] = {(double)pos[sa], (double)pos[cap + sa]}, x2[2] = {(double)pos[sb], (double)pos[cap + sb]};
    double v1[2] = {(double)vel[sa], (double)vel[cap + sa]}, v2[2] = {(double)vel[sb], (double)vel[cap + sb]};
    const double r1[2] = {x1[0] - x2[0], x1[1] - x2[1]};                 // B to A (:340)
    const double nrm = sqrt(r1[0] * r1[0] + r1[1] * r1[1]);
    const double n = nrm * nrm;                                          // norm(r1)^2 (:341)
    uint8_t hap = 0;
    if (n != 0.0) {
        const double dv[2] = {v1[0] - v2[0], v1[1] - v2[1]};
        const double r2[2] = {x2[0] - x1[0], x2[1] - x1[1]};             // A to B (:344)
        const double m1 = mass ? (double)mass[sa] : 1.0, m2 = mass ? (double)mass[sb] : 1.0;
        const bool i1 = isinf(m1) && m1 > 0, i2 = isinf(m2) && m2 > 0;
        double f1 = 0, f2 = 0;
        bool go = true;
        if (i1 && i2) go = false;                                        // :347
        else if (i1) {
            if (v1[0] != 0.0 || v1[1] != 0.0) { atomicOr(err, EC_ASSERT); go = false; }
            else if (r1[0] * v2[0] + r1[1] * v2[1] <= 0.0) go = false;   // :351
            else { v1[0] = v1[1] = 0.0; f1 = 0.0; f2 = 2.0; }
        } else if (i2) {
            if (v2[0] != 0.0 || v2[1] != 0.0) { atomicOr(err, EC_ASSERT); go = false; }
            else if (r2[0] * v1[0] + r2[1] * v1[1] <= 0.0) go = false;   // :356
            else { v2[0] = v2[1] = 0.0; f1 = 2.0; f2 = 0.0; }
        } else {
            if (dv[0] * r2[0] + dv[1] * r2[1] <= 0.0) go = false;        // disks do not face each other (:361)
            else { f1 = 2.0 * m2 / (m1 + m2); f2 = 2.0 * m1 / (m1 + m2); }
        }
        if (go) {
            const double q1 = (dv[0] * r1[0] + dv[1] * r1[1]) / n, q2 = (dv[0] * r2[0] + dv[1] * r2[1]) / n;
#pragma unroll
            for (int d = 0; d < 2; ++d) {
                vel[(size_t)d * cap + sa] = (T)(v1[d] - f1 * q1 * r1[d]);      // :365
                vel[(size_t)d * cap + sb] = (T)(v2[d] + f2 * q2 * r2[d]);      // :366
            }
            hap = 1;
        }
    }
    if (happened) happened[pidx[t]] = hap;
}

template <class T>
static int launch_round(ab_handle* h, const int64_t* da, const int64_t* db, const int32_t* dix, int np, int mass_field,
                        uint8_t* dhap) {
    const int cur = h->cur;
    const T* pos = (const T*)h->fields[0].buf[cur];
    T* vel = (T*)h->fields[1].buf[cur];
    AB_LAUNCH(h, "k_elastic");
    h->agg_valid = false;                                     // velocities change in place: Σ vel per cell is stale
    const int nb = (int)ab_blocks(np, 128);
    if (mass_field < 0)
        k_elastic<T, double><<<nb, 128, 0, h->stream>>>(da, db, dix, np, h->slot_of_id, pos, vel, (const double*)nullptr, h->cap, dhap, h->d_err);
    else if (h->fields[mass_field].dtype == AB_F64)
        k_elastic<T, double><<<nb, 128, 0, h->stream>>>(da, db, dix, np, h->slot_of_id, pos, vel, (const double*)h->fields[mass_field].buf[cur], h->cap, dhap, h->d_err);
    else
        k_elastic<T, float><<<nb, 128, 0, h->stream>>>(da, db, dix, np, h->slot_of_id, pos, vel, (const float*)h->fields[mass_field].buf[cur], h->cap, dhap, h->d_err);
    return AB_OK;
}

extern "C" {

int ab_elastic_collisions(ab_handle* h, const int64_t* a, const int64_t* b, int64_t npairs, int mass_field,
                          uint8_t* happened, int64_t* ncollided) {
    if (h->win_W) return ab_fail(h, AB_ERR_STATE, "not available on a sharded handle");
    if (h->desc.kind != AB_SPACE_CONTINUOUS) return ab_fail(h, AB_ERR_INVALID, "elastic_collision! needs a ContinuousSpace handle");
    if (h->D != 2) return ab_fail(h, AB_ERR_INVALID, "This function works only for two dimensions.");      // continuous.jl:339
    if (npairs < 0 || (npairs && (!a || !b))) return ab_fail(h, AB_ERR_INVALID, "bad pair list");
    if (mass_field >= (int)h->fields.size()) return ab_fail(h, AB_ERR_INVALID, "bad mass field");
    if (mass_field >= 0) {
        const ab_field& f = h->fields[mass_field];
        if (f.ncomp != 1 || (f.dtype != AB_F64 && f.dtype != AB_F32)) return ab_fail(h, AB_ERR_INVALID, "mass must be a scalar f64/f32 column");
    }
    if (ncollided) *ncollided = 0;
    if (npairs == 0) return AB_OK;
    AB_CUDA(h, cudaSetDevice(h->device));
    // rounds: pairs sharing an agent keep the reference's sequential order
    std::unordered_map<int64_t, int> last;
    std::vector<int> round((size_t)npairs);
    int nrounds = 0;
    for (int64_t k = 0; k < npairs; ++k) {
        if (a[k] < 1 || a[k] > h->n || b[k] < 1 || b[k] > h->n || a[k] == b[k]) return ab_fail(h, AB_ERR_INVALID, "pair id out of range");
        int r = 0;
        auto ia = last.find(a[k]), ib = last.find(b[k]);
        if (ia != last.end()) r = std::max(r, ia->second);
        if (ib != last.end()) r = std::max(r, ib->second);
        round[(size_t)k] = r;
        last[a[k]] = last[b[k]] = r + 1;
        nrounds = std::max(nrounds, r + 1);
    }
    std::vector<int32_t> idx((size_t)npairs);
    for (int64_t k = 0; k < npairs; ++k) idx[(size_t)k] = (int32_t)k;
    std::stable_sort(idx.begin(), idx.end(), [&](int32_t x, int32_t y) { return round[(size_t)x] < round[(size_t)y]; });
    std::vector<int64_t> sa((size_t)npairs), sb((size_t)npairs);
    std::vector<int> rbeg((size_t)nrounds + 1, 0);
    for (int64_t k = 0; k < npairs; ++k) {
        sa[(size_t)k] = a[idx[(size_t)k]];
        sb[(size_t)k] = b[idx[(size_t)k]];
        rbeg[(size_t)round[(size_t)idx[(size_t)k]] + 1]++;
    }
    for (int r = 0; r < nrounds; ++r) rbeg[(size_t)r + 1] += rbeg[(size_t)r];
    AB_TRY(ab_ensure_slot_map(h));
    char* dev;
    const size_t np = (size_t)npairs;
    AB_TRY(ab_scratch(h, np * 21 + 1024, (void**)&dev));
    int64_t* da = (int64_t*)dev;
    int64_t* db = da + np;
    int32_t* dix = (int32_t*)(db + np);
    uint8_t* dhap = (uint8_t*)(dix + np);
    AB_CUDA(h, cudaMemcpyAsync(da, sa.data(), np * 8, cudaMemcpyHostToDevice, h->stream));
    AB_CUDA(h, cudaMemcpyAsync(db, sb.data(), np * 8, cudaMemcpyHostToDevice, h->stream));
    AB_CUDA(h, cudaMemcpyAsync(dix, idx.data(), np * 4, cudaMemcpyHostToDevice, h->stream));
    for (int r = 0; r < nrounds; ++r) {
        const int b0 = rbeg[(size_t)r], cnt = rbeg[(size_t)r + 1] - b0;
        if (cnt <= 0) continue;
        if (h->dtype == AB_F64) AB_TRY(launch_round<double>(h, da + b0, db + b0, dix + b0, cnt, mass_field, dhap));
        else AB_TRY(launch_round<float>(h, da + b0, db + b0, dix + b0, cnt, mass_field, dhap));
    }
    std::vector<uint8_t> hap(np);
    AB_CUDA(h, cudaMemcpyAsync(hap.data(), dhap, np, cudaMemcpyDeviceToHost, h->stream));
    AB_CUDA(h, cudaMemcpyAsync(h->h_err, h->d_err, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    AB_CUDA(h, cudaStreamSynchronize(h->stream));
    if (*h->h_err & EC_ASSERT) {
        AB_CUDA(h, cudaMemsetAsync(h->d_err, 0, sizeof(int), h->stream));
        return ab_fail(h, AB_ERR_INVALID, "AssertionError: An agent with ∞ mass cannot have nonzero velocity");
    }
    int64_t c = 0;
    for (size_t k = 0; k < np; ++k) { c += hap[k]; if (happened) happened[k] = hap[k]; }
    if (ncollided) *ncollided = c;
    return AB_OK;
}

}  // extern "C"
