// K3 placeholder until the tcgen05 extend kernel lands: the entry point exists (the header
// declares it) and fails loudly.  Extend currently runs row-wise through bsgl_paged_attention.
#include "common.cuh"

extern "C" int bsgl_extend_attention(const void*, const void*, const void*, const int32_t*, int64_t, const int64_t*,
                                     const int64_t*, const int64_t*, const int32_t*, int, int64_t, int, int, int,
                                     float, int, void*, void*) {
  bsgl::set_error("bsgl_extend_attention: tcgen05 extend kernel not built yet");
  return BSGL_ERR_UNSUPPORTED;
}
