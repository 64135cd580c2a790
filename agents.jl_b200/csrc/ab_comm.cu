// ab_comm.cu — multi-GPU slab decomposition (one process per GPU).
#include "ab_internal.cuh"

void ab_comm_free(ab_handle* h) {}
int ab_comm_active(const ab_handle* h) { return 0; }
int ab_comm_step_flock(ab_handle* h, const ab_flock_params*, int64_t) { return ab_fail(h, AB_ERR_COMM, "not implemented yet"); }

extern "C" {
int ab_comm_unique_id(void* out128) { return AB_ERR_COMM; }
int ab_comm_init(ab_handle* h, int rank, int nranks, const void* id) { return ab_fail(h, AB_ERR_COMM, "multi-GPU: not implemented yet"); }
int ab_comm_slab(const ab_handle* h, int64_t* first_layer, int64_t* n_layers) { return AB_ERR_COMM; }
int64_t ab_nlocal(const ab_handle* h) { return h ? h->n : 0; }
int ab_download_local(ab_handle* h, int field, void* host, int64_t* ids_out, int64_t cap, int64_t* n_out) { return ab_fail(h, AB_ERR_COMM, "multi-GPU: not implemented yet"); }
int ab_upload_local(ab_handle* h, int field, const void* host, const int64_t* global_ids, int64_t n) { return ab_fail(h, AB_ERR_COMM, "multi-GPU: not implemented yet"); }
}
