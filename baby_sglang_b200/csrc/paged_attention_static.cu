// bsgl_paged_attention_static: paged GQA decode attention (K2 split-KV pass + K2c LSE merge) for callers
// that cannot allocate device memory.
//
// Who that is.  bsgl_paged_attention needs two scratch buffers for the partial results of sequences that are
// split over several CTAs (part_o, part_lse), and it borrows them from the caller.  A torch caller takes them
// from the caching allocator.  Some hosts of this library have nothing like that to take them from:
//   * a serving worker whose memory budget was fixed when it started — the KV pool is sized to whatever HBM is
//     left after the weights (memory_pool.py's `size` is exactly that), so a cudaMalloc in the middle of serving
//     either fails or eats into a neighbour's budget, and every cudaMalloc/cudaFree synchronises the whole
//     device, stalling the overlap worker's other stream;
//   * a thin C / cgo / JNI binding that only passes through pointers it was handed and has no device allocator
//     of its own to call;
//   * code that runs where allocation is forbidden outright (inside a region that must not synchronise).
// For them the partials live here, in a fixed array in the library's own device image.  The caller still passes
// every input and the output; only the scratch is ours.
//
// What to know before using it:
//   * The array is BSGL_STATIC_SCRATCH_BYTES (8 MiB) on EVERY GPU whose context loads this library, from the
//     moment the module is loaded until the process exits.  A request that needs more (the need grows with
//     n_ctas * q-heads-per-group * head_dim) is refused with BSGL_ERR_INVALID and a message saying how much it
//     wanted; nothing is launched.  148 CTAs at the largest supported geometry need 4.8 MiB.
//   * The first call on each device looks the array's device address up (cudaGetSymbolAddress, with that device
//     current) and remembers it.  That look-up may make the driver load this module, so make the first call on a
//     device before you begin any stream capture on it; later calls only launch the two kernels.
//   * No cudaMalloc, cudaFree, or allocating library routine is called here or in the two launchers it uses.
//   * All calls on one device share the one array.  It keeps whatever the last call left in it; nothing in it
//     has to start at zero, because the merge kernel reads only the slots the attention kernel of the SAME call
//     wrote (the plan decides both), so there is nothing to clear.  Two calls whose kernels can overlap on the
//     GPU — different streams without an ordering between them, or two host threads on one device — race on the
//     partials and produce wrong output.  Keep calls on one device stream-ordered.
//
// The array's name below is, in host code, only a handle for the runtime: passing `g_static_scratch` itself to a
// kernel would hand it a host address.  Host code must go through static_scratch_for_current_device().
#include <mutex>

#include "common.cuh"

namespace bsgl {

static __device__ __align__(256) unsigned char g_static_scratch[BSGL_STATIC_SCRATCH_BYTES];

constexpr int kMaxDevices = 64;
static std::mutex g_static_mu;
static unsigned char* g_static_ptr[kMaxDevices] = {};   // device address of g_static_scratch, per device ordinal

// Device address of the scratch array on the current device; resolved once per device.
static int static_scratch_for_current_device(unsigned char** out) {
  int dev = 0;
  BSGL_CUDA_TRY(cudaGetDevice(&dev));
  BSGL_REQUIRE(dev >= 0 && dev < kMaxDevices, "bsgl_paged_attention_static: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_static_mu);
  if (g_static_ptr[dev] == nullptr) {
    void* p = nullptr;
    BSGL_CUDA_TRY(cudaGetSymbolAddress(&p, g_static_scratch));
    BSGL_REQUIRE(p != nullptr && (uintptr_t)p % 256 == 0,
                 "bsgl_paged_attention_static: scratch symbol resolved to a bad address");
    g_static_ptr[dev] = static_cast<unsigned char*>(p);
  }
  *out = g_static_ptr[dev];
  return BSGL_OK;
}

}  // namespace bsgl

extern "C" int bsgl_static_scratch_bytes(void) { return BSGL_STATIC_SCRATCH_BYTES; }

extern "C" int bsgl_paged_attention_static(const void* q, const void* k_slab, const void* v_slab,
                                           const int32_t* idx_base, const int64_t* row_idx_off,
                                           const int32_t* row_kv_len, const int32_t* plan_hdr,
                                           const int32_t* tile_indptr, const int32_t* cta_first_seq, int n_rows,
                                           int n_q_heads, int n_kv_heads, int head_dim, float sm_scale, int dtype,
                                           void* out, int n_ctas, void* stream) {
  using namespace bsgl;
  BSGL_REQUIRE(n_rows >= 0, "bsgl_paged_attention_static: negative row count");
  if (n_rows == 0) return BSGL_OK;
  int64_t o_floats = 0, lse_floats = 0;
  int rc = bsgl_partial_sizes(n_ctas, n_q_heads, n_kv_heads, head_dim, &o_floats, &lse_floats);
  if (rc != BSGL_OK) return rc;
  // part_o first, part_lse behind it on the next 256-byte boundary
  const int64_t o_bytes = (o_floats * (int64_t)sizeof(float) + 255) / 256 * 256;
  const int64_t need = o_bytes + lse_floats * (int64_t)sizeof(float);
  BSGL_REQUIRE(need <= (int64_t)BSGL_STATIC_SCRATCH_BYTES,
               "bsgl_paged_attention_static: n_ctas=%d with %d q heads over %d kv heads of dim %d needs %lld bytes of "
               "partials; the built-in scratch holds %d (use bsgl_paged_attention with your own buffers, or fewer CTAs)",
               n_ctas, n_q_heads, n_kv_heads, head_dim, (long long)need, (int)BSGL_STATIC_SCRATCH_BYTES);
  unsigned char* base = nullptr;
  rc = static_scratch_for_current_device(&base);
  if (rc != BSGL_OK) return rc;
  float* part_o = reinterpret_cast<float*>(base);
  float* part_lse = reinterpret_cast<float*>(base + o_bytes);
  rc = bsgl_paged_attention(q, k_slab, v_slab, idx_base, row_idx_off, row_kv_len, plan_hdr, tile_indptr,
                            cta_first_seq, n_rows, n_q_heads, n_kv_heads, head_dim, sm_scale, dtype, out, part_o,
                            part_lse, n_ctas, stream);
  if (rc != BSGL_OK) return rc;
  return bsgl_merge_partials(part_o, part_lse, plan_hdr, tile_indptr, cta_first_seq, n_ctas, n_rows, n_q_heads,
                             n_kv_heads, head_dim, dtype, out, stream);
}
