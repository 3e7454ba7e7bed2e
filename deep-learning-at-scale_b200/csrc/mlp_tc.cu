// tcgen05 / TMEM MLP training kernel (DLAS_MODE_TC).  Placeholder until the tensor-core kernel lands:
// reports "unsupported" for every shape so that callers get DLAS_ERR_UNSUPPORTED_SHAPE, never a
// silent fallback.
#include "common.cuh"
namespace dlas {
size_t simt_workspace_bytes(const int* dims, int n_layers);
bool tc_supported(const int*, int) { return false; }
size_t tc_workspace_bytes(const int* dims, int n_layers) { return simt_workspace_bytes(dims, n_layers); }
int tc_grad(const int*, int, int, const float*, const float*, const float*, size_t, float*, void*, size_t,
            cudaStream_t) {
  set_error("DLAS_MODE_TC is not available in this build");
  return DLAS_ERR_UNSUPPORTED_SHAPE;
}
}  // namespace dlas
