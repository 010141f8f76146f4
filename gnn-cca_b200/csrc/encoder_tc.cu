// tcgen05 (kind::tf32) path of the wide node-encoder layer.  Placeholder until the TMA/TMEM kernel lands:
// reports "unsupported" (never falls back silently).
#include "common.cuh"

int gnncca_linear_tf32(const float* x, const float* W, const float* b, float* y, int64_t rows, int32_t in_dim,
                       int32_t out_dim, int32_t relu, void* stream) {
  (void)x; (void)W; (void)b; (void)y; (void)rows; (void)in_dim; (void)out_dim; (void)relu; (void)stream;
  return GNNCCA_ERR_UNSUPPORTED;
}
