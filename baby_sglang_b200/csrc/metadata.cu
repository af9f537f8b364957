// K4  batch metadata built on the device, once per batch (not per layer).
// Replaces the host loops of the reference: SimplifiedForwardBatch.create_prefill_batch's
// per-request `.item()` pairs (forward_batch.py:40-45), create_decode_batch (:61-84) and the
// per-request `.item()` + page-table slicing inside the attention loops (attn_backend.py:43-57,
// 117-125).  Also emits the split-KV plan consumed by the paged-attention kernel.
//
// Integer work on a few KB: one 1024-thread CTA, warp-shuffle block scans, no host sync.
#include <cstdlib>

#include "common.cuh"

namespace bsgl {

constexpr int kMetaThreads = 1024;
constexpr int kMaxSortedBatch = 4096;   // larger batches keep request order (B^2 / 1024 comparisons per thread)

// Exclusive block scan of a packed (hi = a, lo = b) pair of non-negative 32-bit counts.
__device__ __forceinline__ unsigned long long block_excl_scan(unsigned long long v, unsigned long long* warp_tot,
                                                              unsigned long long& block_total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  unsigned long long incl = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    unsigned long long n = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += n;
  }
  __syncthreads();   // protect warp_tot from the previous round
  if (lane == 31) warp_tot[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    unsigned long long w = warp_tot[lane];
    unsigned long long wi = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      unsigned long long n = __shfl_up_sync(0xffffffffu, wi, o);
      if (lane >= o) wi += n;
    }
    warp_tot[lane] = wi - w;               // exclusive per-warp offsets
    if (lane == 31) warp_tot[32] = wi;     // block total
  }
  __syncthreads();
  block_total = warp_tot[32];
  return warp_tot[warp] + incl - v;
}

__device__ __forceinline__ unsigned long long pack(unsigned hi, unsigned lo) {
  return ((unsigned long long)hi << 32) | lo;
}

// Writes plan_hdr and cta_first_seq once tile_indptr[0..n_seqs] is complete (caller synced).
__device__ void finish_plan(int n_tiles, int n_seqs, int n_ctas, int min_tiles, int32_t* plan_hdr,
                            const int32_t* tile_indptr, int32_t* cta_first_seq) {
  int n_eff = n_tiles / min_tiles;
  if (n_eff > n_ctas) n_eff = n_ctas;
  if (n_eff < 1) n_eff = 1;
  if (threadIdx.x == 0) {
    plan_hdr[0] = n_tiles;
    plan_hdr[1] = n_eff;
    plan_hdr[2] = n_seqs;
    plan_hdr[3] = 0;
  }
  for (int c = threadIdx.x; c <= n_ctas; c += blockDim.x) {
    int j = n_seqs;
    if (c <= n_eff && n_tiles > 0) {
      const int t = cta_tile_begin(c, n_tiles, n_eff);
      // smallest j with tile_indptr[j+1] > t
      int lo = 0, hi = n_seqs;
      while (lo < hi) {
        int mid = (lo + hi) >> 1;
        if (tile_indptr[mid + 1] > t) hi = mid; else lo = mid + 1;
      }
      j = lo;
    }
    cta_first_seq[c] = j;
  }
}

__global__ void __launch_bounds__(kMetaThreads) decode_metadata_kernel(
    int64_t C, const int64_t* __restrict__ rpi, const int64_t* __restrict__ seq_lens, int B, int nhg, int n_ctas,
    int64_t* __restrict__ positions, int32_t* __restrict__ kv_indptr, int32_t* __restrict__ qo_indptr,
    int64_t* __restrict__ row_idx_off, int32_t* __restrict__ row_kv_len, int32_t* __restrict__ plan_hdr,
    int32_t* __restrict__ tile_indptr, int32_t* __restrict__ cta_first_seq, int min_tiles) {
  __shared__ unsigned long long warp_tot[33];
  unsigned carry_kv = 0, carry_tiles = 0;
  for (int base = 0; base < B; base += kMetaThreads) {
    const int i = base + threadIdx.x;
    const int seq = i < B ? (int)seq_lens[i] : 0;
    const int tiles = (seq + BSGL_TILE_TOKENS - 1) / BSGL_TILE_TOKENS;
    unsigned long long tot;
    const unsigned long long ex = block_excl_scan(pack(seq, tiles), warp_tot, tot);
    if (i < B) {
      kv_indptr[i] = (int32_t)(carry_kv + (unsigned)(ex >> 32));
      qo_indptr[i] = i;
      positions[i] = (int64_t)seq - 1;
      row_kv_len[i] = seq;
      row_idx_off[i] = rpi[i] * C;
      const unsigned t0 = (carry_tiles + (unsigned)(ex & 0xffffffffu)) * nhg;
      for (int g = 0; g < nhg; ++g) tile_indptr[(int64_t)i * nhg + g] = (int32_t)(t0 + g * tiles);
    }
    carry_kv += (unsigned)(tot >> 32);
    carry_tiles += (unsigned)(tot & 0xffffffffu);
  }
  const int n_tiles = (int)(carry_tiles * nhg);
  if (threadIdx.x == 0) {
    kv_indptr[B] = (int32_t)carry_kv;
    qo_indptr[B] = B;
    tile_indptr[(int64_t)B * nhg] = n_tiles;
  }
  __syncthreads();
  finish_plan(n_tiles, B * nhg, n_ctas, min_tiles, plan_hdr, tile_indptr, cta_first_seq);
}

__global__ void __launch_bounds__(kMetaThreads) extend_metadata_kernel(
    int64_t C, const int64_t* __restrict__ rpi, const int64_t* __restrict__ seq_lens,
    const int64_t* __restrict__ prefix_lens, int B, int64_t T, int nhg, int n_ctas,
    int64_t* __restrict__ extended_lens, int64_t* __restrict__ positions, int32_t* __restrict__ qo_indptr,
    int32_t* __restrict__ kv_indptr, int64_t* __restrict__ row_idx_off, int32_t* __restrict__ row_kv_len,
    int32_t* __restrict__ plan_hdr, int32_t* __restrict__ tile_indptr, int32_t* __restrict__ cta_first_seq,
    int32_t* __restrict__ heavy_order, int min_tiles) {
  __shared__ unsigned long long warp_tot[33];
  // phase 0: requests by descending seq_len (rank by counting; ties keep index order)
  if (heavy_order != nullptr) {
    __shared__ int32_t keys[kMaxSortedBatch];
    if (B <= kMaxSortedBatch) {
      for (int i = threadIdx.x; i < B; i += kMetaThreads) keys[i] = (int32_t)seq_lens[i];
      __syncthreads();
      for (int i = threadIdx.x; i < B; i += kMetaThreads) {
        const int32_t k = keys[i];
        int rank = 0;
        for (int j = 0; j < B; ++j) rank += (keys[j] > k) || (keys[j] == k && j < i);
        heavy_order[rank] = i;
      }
    } else {
      for (int i = threadIdx.x; i < B; i += kMetaThreads) heavy_order[i] = i;
    }
  }
  // phase 1: per request
  unsigned carry_q = 0, carry_kv = 0;
  for (int base = 0; base < B; base += kMetaThreads) {
    const int i = base + threadIdx.x;
    const int seq = i < B ? (int)seq_lens[i] : 0;
    const int pre = i < B ? (int)prefix_lens[i] : 0;
    const int ext = seq - pre;
    unsigned long long tot;
    const unsigned long long ex = block_excl_scan(pack(ext, seq), warp_tot, tot);
    if (i < B) {
      extended_lens[i] = ext;
      qo_indptr[i] = (int32_t)(carry_q + (unsigned)(ex >> 32));
      kv_indptr[i] = (int32_t)(carry_kv + (unsigned)(ex & 0xffffffffu));
    }
    carry_q += (unsigned)(tot >> 32);
    carry_kv += (unsigned)(tot & 0xffffffffu);
  }
  if (threadIdx.x == 0) {
    qo_indptr[B] = (int32_t)carry_q;
    kv_indptr[B] = (int32_t)carry_kv;
  }
  __syncthreads();
  // phase 2: per q token (row)
  unsigned carry_tiles = 0;
  for (int64_t base = 0; base < T; base += kMetaThreads) {
    const int64_t t = base + threadIdx.x;
    int tiles = 0;
    int kvl = 0;
    if (t < T) {
      // request i with qo_indptr[i] <= t < qo_indptr[i+1]
      int lo = 0, hi = B;          // find first i with qo_indptr[i+1] > t
      while (lo < hi) {
        int mid = (lo + hi) >> 1;
        if ((int64_t)qo_indptr[mid + 1] > t) hi = mid; else lo = mid + 1;
      }
      const int i = lo < B ? lo : B - 1;
      const int64_t r = t - qo_indptr[i];
      const int64_t pre = prefix_lens[i];
      positions[t] = pre + r;
      kvl = (int)(pre + r + 1);
      row_kv_len[t] = kvl;
      row_idx_off[t] = rpi[i] * C;
      tiles = (kvl + BSGL_TILE_TOKENS - 1) / BSGL_TILE_TOKENS;
    }
    if (tile_indptr != nullptr) {
      unsigned long long tot;
      const unsigned long long ex = block_excl_scan(pack(0, tiles), warp_tot, tot);
      if (t < T) {
        const unsigned t0 = (carry_tiles + (unsigned)(ex & 0xffffffffu)) * nhg;
        for (int g = 0; g < nhg; ++g) tile_indptr[t * nhg + g] = (int32_t)(t0 + g * tiles);
      }
      carry_tiles += (unsigned)(tot & 0xffffffffu);
    }
  }
  if (tile_indptr != nullptr) {
    const int n_tiles = (int)(carry_tiles * nhg);
    if (threadIdx.x == 0) tile_indptr[T * nhg] = n_tiles;
    __syncthreads();
    finish_plan(n_tiles, (int)(T * nhg), n_ctas, min_tiles, plan_hdr, tile_indptr, cta_first_seq);
  }
}

__global__ void __launch_bounds__(256) gather_kv_indices_kernel(const int32_t* __restrict__ req_to_token, int64_t C,
                                                                const int64_t* __restrict__ rpi,
                                                                const int32_t* __restrict__ kv_indptr,
                                                                int32_t* __restrict__ kv_indices) {
  const int i = blockIdx.x;
  const int begin = kv_indptr[i], end = kv_indptr[i + 1];
  const int32_t* row = req_to_token + rpi[i] * C;
  for (int p = blockIdx.y * blockDim.x + threadIdx.x; p < end - begin; p += gridDim.y * blockDim.x)
    kv_indices[begin + p] = row[p];
}

// A CTA of the split-KV kernel gets at least this many 16-token tiles, so short batches use fewer CTAs and their
// sequences are cut into fewer partials.  BSGL_MIN_TILES_PER_CTA (16) is the measured default; the environment
// variable of the same name overrides it for tuning runs (read once).
static int min_tiles_per_cta() {
  static const int v = [] {
    const char* e = getenv("BSGL_MIN_TILES_PER_CTA");
    const int x = e ? atoi(e) : BSGL_MIN_TILES_PER_CTA;
    return x >= 1 ? x : BSGL_MIN_TILES_PER_CTA;
  }();
  return v;
}

}  // namespace bsgl

extern "C" {

int bsgl_min_tiles_per_cta(void) { return bsgl::min_tiles_per_cta(); }

int bsgl_build_decode_metadata(int64_t page_table_stride, const int64_t* req_pool_indices, const int64_t* seq_lens,
                               int batch, int n_head_groups, int n_ctas, int64_t* positions, int32_t* kv_indptr,
                               int32_t* qo_indptr, int64_t* row_idx_off, int32_t* row_kv_len, int32_t* plan_hdr,
                               int32_t* tile_indptr, int32_t* cta_first_seq, void* stream) {
  using namespace bsgl;
  BSGL_REQUIRE(batch >= 0 && n_head_groups > 0 && n_ctas > 0, "bsgl_build_decode_metadata: bad sizes");
  BSGL_REQUIRE(positions && kv_indptr && qo_indptr && row_idx_off && row_kv_len && plan_hdr && tile_indptr &&
                   cta_first_seq,
               "bsgl_build_decode_metadata: null output");
  BSGL_REQUIRE(batch == 0 || (req_pool_indices && seq_lens), "bsgl_build_decode_metadata: null input");
  decode_metadata_kernel<<<1, kMetaThreads, 0, (cudaStream_t)stream>>>(
      page_table_stride, req_pool_indices, seq_lens, batch, n_head_groups, n_ctas, positions, kv_indptr, qo_indptr,
      row_idx_off, row_kv_len, plan_hdr, tile_indptr, cta_first_seq, min_tiles_per_cta());
  return check_launch("decode_metadata_kernel");
}

int bsgl_build_extend_metadata(int64_t page_table_stride, const int64_t* req_pool_indices, const int64_t* seq_lens,
                               const int64_t* prefix_lens, int batch, int64_t n_tokens, int n_head_groups, int n_ctas,
                               int64_t* extended_lens, int64_t* positions, int32_t* qo_indptr, int32_t* kv_indptr,
                               int64_t* row_idx_off, int32_t* row_kv_len, int32_t* plan_hdr, int32_t* tile_indptr,
                               int32_t* cta_first_seq, int32_t* heavy_order, void* stream) {
  using namespace bsgl;
  BSGL_REQUIRE(batch >= 0 && n_tokens >= 0 && n_head_groups > 0 && n_ctas > 0,
               "bsgl_build_extend_metadata: bad sizes");
  BSGL_REQUIRE(extended_lens && positions && qo_indptr && kv_indptr && row_idx_off && row_kv_len,
               "bsgl_build_extend_metadata: null output");
  BSGL_REQUIRE((tile_indptr == nullptr) == (plan_hdr == nullptr) && (tile_indptr == nullptr) == (cta_first_seq == nullptr),
               "bsgl_build_extend_metadata: plan outputs must be all set or all NULL");
  BSGL_REQUIRE(batch == 0 || (req_pool_indices && seq_lens && prefix_lens), "bsgl_build_extend_metadata: null input");
  BSGL_REQUIRE(n_tokens * n_head_groups < (1ll << 31), "bsgl_build_extend_metadata: too many rows");
  extend_metadata_kernel<<<1, kMetaThreads, 0, (cudaStream_t)stream>>>(
      page_table_stride, req_pool_indices, seq_lens, prefix_lens, batch, n_tokens, n_head_groups, n_ctas,
      extended_lens, positions, qo_indptr, kv_indptr, row_idx_off, row_kv_len, plan_hdr, tile_indptr, cta_first_seq,
      heavy_order, min_tiles_per_cta());
  return check_launch("extend_metadata_kernel");
}

int bsgl_gather_kv_indices(const int32_t* req_to_token, int64_t page_table_stride, const int64_t* req_pool_indices,
                           const int32_t* kv_indptr, int batch, int32_t* kv_indices, void* stream) {
  using namespace bsgl;
  BSGL_REQUIRE(batch >= 0, "bsgl_gather_kv_indices: negative batch");
  if (batch == 0) return BSGL_OK;
  BSGL_REQUIRE(req_to_token && req_pool_indices && kv_indptr && kv_indices, "bsgl_gather_kv_indices: null pointer");
  dim3 grid((unsigned)batch, 8);
  gather_kv_indices_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(req_to_token, page_table_stride, req_pool_indices,
                                                                  kv_indptr, kv_indices);
  return check_launch("gather_kv_indices_kernel");
}

}  // extern "C"
