// K7  page-table writes and the free-slot FIFO on the device (sm_100a).  Integer work, bit-exact.
//
// Replaces, in the reference:
//   * prefill row construction  scheduler.py:423-442  (zeros([B, C]) + two slice writes and one
//     `.item()` per request + a whole-row index-put): ONE launch writes every row of the batch as
//     [reused prefix slots | newly allocated slots | 0 ...].
//   * decode append             scheduler.py:486-488  (B scalar device writes): ONE launch.
//   * slot allocator            memory_pool.py:68-80  (CPU tensor FIFO; `alloc` = slice + H2D copy,
//     `free` = blocking `.cpu()` + concat): a ring in HBM; pop / push are stream-ordered copies with
//     wrap-around, the FIFO order is the reference's (head out, tail in, no sort, no de-dup).
//     Head, tail and count stay on the host: every size is known there without a device sync.
//
// All of it is tiny HBM-bound traffic (the NS page table is 4.3 MB); the point is launch count and
// host syncs, not bandwidth: coalesced 4-byte accesses, one thread per cell.
#include "common.cuh"

namespace bsgl {

// desc[5][B] int64 rows: req_pool_idx, prefix_len, seq_len, new_off (offset of the request's first new
// slot in `new_slots`), prefix_off (offset of its first reused slot in `prefix_flat`).
__global__ void __launch_bounds__(256)
page_table_write_rows_kernel(int32_t* __restrict__ req_to_token, int64_t stride, int64_t n_rows, int row_len,
                             const int64_t* __restrict__ desc, int batch, const int32_t* __restrict__ prefix_flat,
                             const int32_t* __restrict__ new_slots) {
  const int i = blockIdx.y;
  const int64_t row = desc[i];
  if (row < 0 || row >= n_rows) return;   // never write outside the table
  const int prefix = (int)desc[(int64_t)batch + i];
  const int seq = (int)desc[2ll * batch + i];
  const int64_t new_off = desc[3ll * batch + i];
  const int64_t prefix_off = desc[4ll * batch + i];
  int32_t* dst = req_to_token + row * stride;
  for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < row_len; c += gridDim.x * blockDim.x) {
    int32_t v = 0;
    if (c < prefix) v = prefix_flat[prefix_off + c];
    else if (c < seq) v = new_slots[new_off + (c - prefix)];
    dst[c] = v;
  }
}

// req_to_token[req_pool_idx[i], seq_len[i] - 1] = slots[i]   (seq_len already counts the new token)
__global__ void __launch_bounds__(256)
page_table_append_kernel(int32_t* __restrict__ req_to_token, int64_t stride, int64_t n_rows, int row_len,
                         const int64_t* __restrict__ req_pool_indices, const int64_t* __restrict__ seq_lens,
                         const int32_t* __restrict__ slots, int batch) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= batch) return;
  const int64_t row = req_pool_indices[i], pos = seq_lens[i] - 1;
  if (row < 0 || row >= n_rows || pos < 0 || pos >= row_len) return;
  req_to_token[row * stride + pos] = slots[i];
}

__global__ void __launch_bounds__(256)
ring_pop_kernel(const int32_t* __restrict__ ring, int64_t cap, int64_t head, int64_t n, int32_t* __restrict__ out) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    int64_t p = head + i;
    if (p >= cap) p -= cap;
    out[i] = ring[p];
  }
}

template <typename V>
__global__ void __launch_bounds__(256)
ring_push_kernel(int32_t* __restrict__ ring, int64_t cap, int64_t tail, int64_t n, const V* __restrict__ vals,
                 int64_t val_stride) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    int64_t p = tail + i;
    if (p >= cap) p -= cap;
    ring[p] = (int32_t)vals[i * val_stride];
  }
}

static inline unsigned blocks_for(int64_t n, int threads, int64_t cap) {
  int64_t b = (n + threads - 1) / threads;
  return (unsigned)(b < 1 ? 1 : (b > cap ? cap : b));
}

}  // namespace bsgl

extern "C" int bsgl_page_table_write_rows(int32_t* req_to_token, int64_t page_table_stride, int64_t n_rows, int row_len,
                                          const int64_t* desc, int batch, const int32_t* prefix_flat,
                                          const int32_t* new_slots, void* stream) {
  using namespace bsgl;
  BSGL_REQUIRE(batch >= 0 && row_len >= 0, "bsgl_page_table_write_rows: negative sizes");
  if (batch == 0 || row_len == 0) return BSGL_OK;
  BSGL_REQUIRE(req_to_token && desc, "bsgl_page_table_write_rows: null pointer");
  BSGL_REQUIRE(page_table_stride >= row_len, "bsgl_page_table_write_rows: stride %lld < row_len %d",
               (long long)page_table_stride, row_len);
  BSGL_REQUIRE(batch <= 65535, "bsgl_page_table_write_rows: batch %d > 65535", batch);
  dim3 grid(blocks_for(row_len, 256, 64), (unsigned)batch);
  page_table_write_rows_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(req_to_token, page_table_stride, n_rows,
                                                                        row_len, desc, batch, prefix_flat, new_slots);
  return check_launch("page_table_write_rows_kernel");
}

extern "C" int bsgl_page_table_append(int32_t* req_to_token, int64_t page_table_stride, int64_t n_rows, int row_len,
                                      const int64_t* req_pool_indices, const int64_t* seq_lens, const int32_t* slots,
                                      int batch, void* stream) {
  using namespace bsgl;
  BSGL_REQUIRE(batch >= 0, "bsgl_page_table_append: negative batch");
  if (batch == 0) return BSGL_OK;
  BSGL_REQUIRE(req_to_token && req_pool_indices && seq_lens && slots, "bsgl_page_table_append: null pointer");
  page_table_append_kernel<<<blocks_for(batch, 256, 1 << 20), 256, 0, (cudaStream_t)stream>>>(
      req_to_token, page_table_stride, n_rows, row_len, req_pool_indices, seq_lens, slots, batch);
  return check_launch("page_table_append_kernel");
}

extern "C" int bsgl_slot_ring_pop(const int32_t* ring, int64_t capacity, int64_t head, int64_t n, int32_t* out,
                                  void* stream) {
  using namespace bsgl;
  BSGL_REQUIRE(n >= 0 && capacity > 0 && head >= 0 && head < capacity && n <= capacity,
               "bsgl_slot_ring_pop: bad ring arguments (cap %lld head %lld n %lld)", (long long)capacity,
               (long long)head, (long long)n);
  if (n == 0) return BSGL_OK;
  BSGL_REQUIRE(ring && out, "bsgl_slot_ring_pop: null pointer");
  ring_pop_kernel<<<blocks_for(n, 256, 1184), 256, 0, (cudaStream_t)stream>>>(ring, capacity, head, n, out);
  return check_launch("ring_pop_kernel");
}

extern "C" int bsgl_slot_ring_push(int32_t* ring, int64_t capacity, int64_t tail, int64_t n, const void* vals,
                                   int64_t val_stride, int vals_are_int64, void* stream) {
  using namespace bsgl;
  BSGL_REQUIRE(n >= 0 && capacity > 0 && tail >= 0 && tail < capacity && n <= capacity,
               "bsgl_slot_ring_push: bad ring arguments (cap %lld tail %lld n %lld)", (long long)capacity,
               (long long)tail, (long long)n);
  if (n == 0) return BSGL_OK;
  BSGL_REQUIRE(ring && vals, "bsgl_slot_ring_push: null pointer");
  const unsigned g = blocks_for(n, 256, 1184);
  if (vals_are_int64)
    ring_push_kernel<int64_t><<<g, 256, 0, (cudaStream_t)stream>>>(ring, capacity, tail, n, (const int64_t*)vals,
                                                                    val_stride);
  else
    ring_push_kernel<int32_t><<<g, 256, 0, (cudaStream_t)stream>>>(ring, capacity, tail, n, (const int32_t*)vals,
                                                                    val_stride);
  return check_launch("ring_push_kernel");
}
