// Error slot, version and small host-side queries of the bsgl_b200 C ABI.
#include <cstring>

#include "common.cuh"

namespace bsgl {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what) {
  set_error("CUDA error %d (%s) at: %s", (int)e, cudaGetErrorString(e), what);
  return BSGL_ERR_CUDA;
}

}  // namespace bsgl

extern "C" {

int bsgl_version(void) { return BSGL_VERSION; }

const char* bsgl_last_error(void) { return bsgl::g_err; }

int bsgl_sm_count(int* out_sm_count) {
  BSGL_REQUIRE(out_sm_count != nullptr, "bsgl_sm_count: null output");
  int dev = 0;
  BSGL_CUDA_TRY(cudaGetDevice(&dev));
  BSGL_CUDA_TRY(cudaDeviceGetAttribute(out_sm_count, cudaDevAttrMultiProcessorCount, dev));
  return BSGL_OK;
}

int bsgl_heads_per_group(int n_kv_heads) {
  if (n_kv_heads <= 0) return 0;
  if (n_kv_heads % 4 == 0) return 4;
  if (n_kv_heads % 2 == 0) return 2;
  return 1;
}

int bsgl_partial_sizes(int n_ctas, int n_q_heads, int n_kv_heads, int head_dim, int64_t* part_o_floats,
                       int64_t* part_lse_floats) {
  BSGL_REQUIRE(n_ctas > 0 && n_q_heads > 0 && n_kv_heads > 0 && head_dim > 0, "bsgl_partial_sizes: bad sizes");
  BSGL_REQUIRE(n_q_heads % n_kv_heads == 0, "bsgl_partial_sizes: n_q_heads %% n_kv_heads != 0");
  int hpb = bsgl_heads_per_group(n_kv_heads);
  int64_t heads = (int64_t)hpb * (n_q_heads / n_kv_heads);
  if (part_o_floats) *part_o_floats = 2ll * n_ctas * heads * head_dim;
  if (part_lse_floats) *part_lse_floats = 2ll * n_ctas * heads;
  return BSGL_OK;
}

}  // extern "C"
