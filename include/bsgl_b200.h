/*
 * bsgl_b200 — C ABI of the B200-native paged-attention path for baby-sglang.
 *
 * This is the drop-in boundary: plain pointers, sizes and a CUDA stream handle; no torch
 * types.  Every entry point returns 0 on success and a non-zero code on failure, in which
 * case bsgl_last_error() (thread-local) describes it.  All device pointers are BORROWED for
 * stream-ordered work only: the library never allocates, frees or retains them, holds no
 * process-global device scratch (one exception, opt-in: bsgl_paged_attention_static below keeps
 * its partials in an 8 MiB array inside the library's device image), and launches on the `stream`
 * it is given (the reference runs this path on a worker thread under a non-default stream,
 * overlap_worker.py:104-106).
 *
 * What each entry point replaces in the reference (paths relative to baby-llm/baby-sglang):
 *
 *   bsgl_kv_store               MHATokenToKVPool.set_kv_buffer            memory_pool.py:152-171
 *   bsgl_build_decode_metadata  SimplifiedForwardBatch.create_decode_batch forward_batch.py:61-84
 *                               + the per-request .item()/slice chain      attn_backend.py:117-125
 *   bsgl_build_extend_metadata  SimplifiedForwardBatch.create_prefill_batch forward_batch.py:24-59
 *                               + the per-request .item()/slice chain      attn_backend.py:43-57
 *   bsgl_gather_kv_indices      req_to_token[r, :seq] flattening           attn_backend.py:55-57,123-125
 *   bsgl_paged_attention        _decode_attention (gather, GQA expand,     attn_backend.py:112-149
 *                               SDPA, cat) — split-KV partial pass; also
 *                               row-wise extend for short extends          attn_backend.py:34-110
 *   bsgl_merge_partials         (no counterpart: LSE merge of the split-KV partials)
 *   bsgl_paged_attention_static (no counterpart: K2 + K2c for callers that cannot allocate scratch)
 *   bsgl_extend_attention       _prefill_attention                         attn_backend.py:34-110
 *   bsgl_rope_kv_store          RotaryEmbedding.forward + split + set_kv_buffer qwen2.py:249-290
 *   bsgl_greedy_sample          Scheduler._sample_next_ids (greedy requests)   scheduler.py:509-546, sample.py:19-60
 *   bsgl_page_table_write_rows  prefill page-table rows                    scheduler.py:423-442
 *   bsgl_page_table_append      decode page-table cells                    scheduler.py:486-488
 *   bsgl_slot_ring_pop / _push  BaseTokenToKVPool.alloc / free             memory_pool.py:68-80
 *   bsgl_oproj_allreduce        o_proj + tp_all_reduce_sum (KV-head TP)    qwen2.py:166,195, tp_runtime.py:28-33
 *
 * Layout conventions (identical to the reference's tensors):
 *   q        [rows, n_q_heads, head_dim]            contiguous, bf16 or fp16
 *   k/v slab [S+1, n_kv_heads, head_dim]            contiguous, slot 0 is the pad slot
 *   out      [rows, n_q_heads*head_dim]             contiguous, q's dtype
 *   page table req_to_token [max_reqs, C] int32, req_pool_indices int64, seq_lens int64,
 *   out_cache_loc int32 (forward_batch.py:10-22 as probed in SURVEY.md §8a).
 *   q head h reads kv head h / (n_q_heads / n_kv_heads).  Softmax scale is passed explicitly;
 *   the reference's effective value is head_dim^-0.5 (it never forwards `scaling`,
 *   attn_backend.py:18,94-101).
 */
#ifndef BSGL_B200_H_
#define BSGL_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BSGL_VERSION 100       /* 0.1.0 */
#define BSGL_TILE_TOKENS 16    /* KV tokens per pipeline tile of the paged-attention kernel */
#define BSGL_MIN_TILES_PER_CTA 16  /* measured: B=1..4 decode 0.82 -> 0.59 ms per 28 layers vs 4 (fewer partials) */
#define BSGL_PLAN_HDR_INTS 4   /* plan_hdr = { n_tiles, n_ctas_eff, n_seqs, reserved } */
#define BSGL_STATIC_SCRATCH_BYTES (8 << 20)  /* built-in partials array of bsgl_paged_attention_static, per GPU */

enum bsgl_dtype { BSGL_BF16 = 0, BSGL_F16 = 1, BSGL_F32 = 2 };

enum bsgl_status {
  BSGL_OK = 0,
  BSGL_ERR_INVALID = 1,      /* bad argument (null pointer, unsupported shape/dtype, misalignment) */
  BSGL_ERR_CUDA = 2,         /* a CUDA runtime call or launch failed */
  BSGL_ERR_UNSUPPORTED = 3   /* valid request this build has no kernel for */
};

int bsgl_version(void);
/* Thread-local description of the last failure on the calling thread ("" if none). */
const char* bsgl_last_error(void);
/* Number of SMs on the current CUDA device (used to size the persistent grids). */
int bsgl_sm_count(int* out_sm_count);
/* KV heads one CTA of the paged-attention kernel handles (the "head group" width) for a pool
 * with n_kv_heads heads: the largest of {4,2,1} dividing n_kv_heads. */
int bsgl_heads_per_group(int n_kv_heads);
/* The value the split-KV plans are built with: BSGL_MIN_TILES_PER_CTA unless the environment variable of the same
 * name (read once per process, for tuning runs) overrides it. */
int bsgl_min_tiles_per_cta(void);

/* K1.  slab[loc[t], h, :] = src[t, h, :] for K and V in one launch.  Source rows may be
 * strided views (the reference hands `v` as a view of the fused QKV output, qwen2.py:285-290):
 * strides are in ELEMENTS; the innermost dimension must be contiguous.  src_dtype may differ
 * from slab_dtype (round-to-nearest-even cast, memory_pool.py:163-165).  Bit-exact.
 * n_slots = rows of each slab (S + 1): a loc[t] outside [0, n_slots) traps the kernel, like the
 * device-side assert of the reference's index-put, instead of writing into a neighbouring layer. */
int bsgl_kv_store(void* k_slab, void* v_slab, int64_t n_slots, const int32_t* loc,
                  const void* k_src, const void* v_src,
                  int64_t n_tokens, int n_kv_heads, int head_dim,
                  int64_t k_row_stride, int64_t k_head_stride,
                  int64_t v_row_stride, int64_t v_head_stride,
                  int src_dtype, int slab_dtype, void* stream);

/* K4 (decode).  One q row per request.  Outputs (all device, caller-allocated):
 *   positions   int64 [B]      = seq_lens - 1                      forward_batch.py:72
 *   kv_indptr   int32 [B+1]    exclusive cumsum of seq_lens
 *   qo_indptr   int32 [B+1]    0..B
 *   row_idx_off int64 [B]      element offset of the row's token list inside req_to_token
 *   row_kv_len  int32 [B]      KV tokens the row attends to (= seq_lens)
 *   plan_hdr    int32 [4], tile_indptr int32 [B*n_head_groups+1], cta_first_seq int32 [n_ctas+1]
 *                              the split-KV plan for bsgl_paged_attention */
int bsgl_build_decode_metadata(int64_t page_table_stride, const int64_t* req_pool_indices,
                               const int64_t* seq_lens, int batch, int n_head_groups, int n_ctas,
                               int64_t* positions, int32_t* kv_indptr, int32_t* qo_indptr,
                               int64_t* row_idx_off, int32_t* row_kv_len,
                               int32_t* plan_hdr, int32_t* tile_indptr, int32_t* cta_first_seq,
                               void* stream);

/* K4 (extend / prefill).  n_tokens = sum(seq_lens - prefix_lens) is known on the host (it is
 * q.shape[0]).  Outputs:
 *   extended_lens int64 [B]    = seq_lens - prefix_lens            forward_batch.py:36-38
 *   positions     int64 [T]    prefix..prefix+ext-1 per request    forward_batch.py:40-45
 *   qo_indptr     int32 [B+1], kv_indptr int32 [B+1]
 *   row_idx_off   int64 [T], row_kv_len int32 [T]  per q token: its request's page-table row
 *                              and prefix + local_row + 1 (mask col <= prefix+row, attn_backend.py:83-85)
 *   plan_hdr / tile_indptr [T*n_head_groups+1] / cta_first_seq [n_ctas+1]: row-wise plan
 *                              (may be NULL when the row-wise kernel will not be used).
 *   heavy_order   int32 [B]    request indices by descending seq_len (ties by index): the dispatch
 *                              order of bsgl_extend_attention's CTAs (may be NULL). */
int bsgl_build_extend_metadata(int64_t page_table_stride, const int64_t* req_pool_indices,
                               const int64_t* seq_lens, const int64_t* prefix_lens, int batch,
                               int64_t n_tokens, int n_head_groups, int n_ctas,
                               int64_t* extended_lens, int64_t* positions,
                               int32_t* qo_indptr, int32_t* kv_indptr,
                               int64_t* row_idx_off, int32_t* row_kv_len,
                               int32_t* plan_hdr, int32_t* tile_indptr, int32_t* cta_first_seq,
                               int32_t* heavy_order, void* stream);

/* kv_indices[kv_indptr[i] + p] = req_to_token[req_pool_indices[i], p] for p < seq_lens[i]. */
int bsgl_gather_kv_indices(const int32_t* req_to_token, int64_t page_table_stride,
                           const int64_t* req_pool_indices, const int32_t* kv_indptr, int batch,
                           int32_t* kv_indices, void* stream);

/* Number of floats the split-KV partial buffers need: part_o has 2*n_ctas slots of
 * heads_per_group*group_size*head_dim floats, part_lse 2*n_ctas slots of heads_per_group*group_size. */
int bsgl_partial_sizes(int n_ctas, int n_q_heads, int n_kv_heads, int head_dim,
                       int64_t* part_o_floats, int64_t* part_lse_floats);

/* K2/K2b.  Paged GQA attention of one q row per "row" over KV tokens
 * idx_base[row_idx_off[row] + 0 .. row_kv_len[row]).  Persistent grid of n_ctas CTAs; each CTA
 * owns an equal contiguous range of 16-token tiles of the flattened (row, head-group) space, so
 * long and ragged contexts split across CTAs.  Sequences that fall entirely inside one CTA are
 * written to `out` directly; the others leave normalised partials + base-2 LSE in
 * part_o/part_lse for bsgl_merge_partials. */
int bsgl_paged_attention(const void* q, const void* k_slab, const void* v_slab,
                         const int32_t* idx_base, const int64_t* row_idx_off,
                         const int32_t* row_kv_len, const int32_t* plan_hdr,
                         const int32_t* tile_indptr, const int32_t* cta_first_seq,
                         int n_rows, int n_q_heads, int n_kv_heads, int head_dim,
                         float sm_scale, int dtype, void* out,
                         float* part_o, float* part_lse, int n_ctas, void* stream);

/* K2c.  LSE merge of the partials of every sequence that was split; rows with no KV tokens
 * get zeros.  Must follow bsgl_paged_attention on the same stream with the same plan and n_ctas
 * (one CTA per plan boundary x 4 q heads: at most n_ctas-1 sequences are split). */
int bsgl_merge_partials(const float* part_o, const float* part_lse, const int32_t* plan_hdr,
                        const int32_t* tile_indptr, const int32_t* cta_first_seq, int n_ctas,
                        int n_rows, int n_q_heads, int n_kv_heads, int head_dim, int dtype, void* out,
                        void* stream);

/* K2 + K2c without caller scratch.  NO COUNTERPART IN THE REFERENCE (its attention path allocates freely through
 * torch); this exists for hosts of the library that cannot allocate device memory where they call it: a serving
 * worker whose HBM was divided up at start-up (weights, then a KV pool sized to what is left, memory_pool.py's
 * `size`) and for which a mid-serving cudaMalloc fails or synchronises the whole device under the overlap worker's
 * second stream; a pass-through C / cgo / JNI binding that owns no device allocator; code in a region that must
 * not allocate or synchronise at all.
 *
 * Same arguments and same results (bit for bit) as bsgl_paged_attention followed by bsgl_merge_partials on the same
 * stream, except that part_o / part_lse are not passed: the partials live in a BSGL_STATIC_SCRATCH_BYTES (8 MiB)
 * array that is part of the library's device image.  Things a caller must know:
 *   - every GPU whose context loads the library holds those 8 MiB until the process exits, used or not;
 *   - a request whose partials do not fit (bsgl_partial_sizes: (part_o + part_lse) floats * 4, part_o rounded up
 *     to 256 bytes) returns BSGL_ERR_INVALID with the needed and available byte counts in bsgl_last_error() and
 *     launches nothing.  148 CTAs need at most 4.8 MiB (32 q heads per group, head_dim 128);
 *   - the first call on each device resolves the array's device address (cudaGetSymbolAddress with that device
 *     current; this can make the driver load the module).  Make that first call before starting any stream
 *     capture on the device.  Later calls only launch the two kernels;
 *   - it calls no cudaMalloc / cudaFree / allocating library routine;
 *   - all calls on a device share the array and it keeps the previous call's contents.  Nothing in it needs to be
 *     zero on entry (the merge reads only what the same call's attention kernel wrote), so nothing is cleared.
 *     Calls whose kernels may run concurrently on one GPU (unordered streams, two host threads) race on it:
 *     keep them stream-ordered, or use bsgl_paged_attention with a buffer per stream.
 * bsgl_static_scratch_bytes() returns the capacity the loaded binary was built with. */
int bsgl_static_scratch_bytes(void);
int bsgl_paged_attention_static(const void* q, const void* k_slab, const void* v_slab,
                                const int32_t* idx_base, const int64_t* row_idx_off,
                                const int32_t* row_kv_len, const int32_t* plan_hdr,
                                const int32_t* tile_indptr, const int32_t* cta_first_seq,
                                int n_rows, int n_q_heads, int n_kv_heads, int head_dim,
                                float sm_scale, int dtype, void* out, int n_ctas, void* stream);

/* K3.  Ragged extend attention (radix-cache prefix + new tokens) on tcgen05/TMEM tensor cores.
 * q rows are packed in request order (qo_indptr); request i's row r sees KV columns
 * 0..prefix_lens[i]+r of req_to_token[req_pool_indices[i], :seq_lens[i]].  req_order (int32 [B], a
 * permutation of 0..B-1, NULL = identity) is the order in which requests are handed to the SMs;
 * K4's heavy_order (longest KV first) shortens the tail of ragged batches.  It never changes results.
 * One persistent CTA per SM walks the (request, KV head, 256-packed-row chunk) work items.  work_counter: int32 [2],
 * zero on entry and returned to zero by the launch (the CTAs draw items greedily from it; launches sharing one must be
 * stream-ordered), or NULL for a static dealing of the items. */
int bsgl_extend_attention(const void* q, const void* k_slab, const void* v_slab,
                          const int32_t* req_to_token, int64_t page_table_stride,
                          const int64_t* req_pool_indices, const int64_t* seq_lens,
                          const int64_t* prefix_lens, const int32_t* qo_indptr,
                          const int32_t* req_order, int batch, int64_t n_tokens, int n_q_heads, int n_kv_heads, int head_dim,
                          float sm_scale, int dtype, void* out, int32_t* work_counter, void* stream);

/* K5 (the step immediately before the path).  RoPE + q/k/v split of the fused
 * projection output + KV store in one launch.  Replaces RotaryEmbedding.forward/_apply_rope
 * (qwen2.py:249-268), the split (qwen2.py:285-290) and set_kv_buffer (memory_pool.py:152-171).
 *   qkv       [T, (Hq + 2*Hkv) * D] rows qkv_row_stride elements apart (the Linear's output)
 *   positions int64 [T], inv_freq float32 [D/2] (the module's own buffer, qwen2.py:244-245)
 *   q_out     [T, Hq, D] contiguous, rotated
 *   k_slab/v_slab[loc[t]] receive the rotated K row and the V row; all three NULL = rotate q only.
 *   n_slots   rows of each slab (S + 1); an out-of-range loc[t] traps (see bsgl_kv_store).
 * Rounding follows the reference: cos/sin rounded to the activation dtype, every product and sum
 * rounded to it (bit-exact against the reference run on the same device). */
int bsgl_rope_kv_store(const void* qkv, int64_t qkv_row_stride, const int64_t* positions,
                       const float* inv_freq, const int32_t* loc, void* k_slab, void* v_slab,
                       int64_t n_slots, void* q_out, int64_t n_tokens, int n_q_heads, int n_kv_heads, int head_dim,
                       int dtype, void* stream);

/* K6 (the step immediately after the path).  Batched greedy sampling on the
 * device: last-token gather (scheduler.py:513-521), repetition penalty over the distinct previous
 * ids (sample.py:19-36) and argmax (sample.py:59-60) for every request in one launch, no `.item()`.
 *   logits     [rows, vocab] bf16/fp16/fp32, rows row_stride elements apart
 *   qo_indptr  int32 [B+1] (K4): request i samples from row qo_indptr[i+1]-1; NULL = row i (decode)
 *   prev_ids   int64 [prev_indptr[B]] with prev_indptr int32 [B+1], penalties float32 [B]
 *              (all three NULL = no repetition penalty)
 *   next_ids   int64 [B]: first index of the row maximum (NaN counts as maximal), like torch.argmax */
int bsgl_greedy_sample(const void* logits, int64_t row_stride, int vocab, int dtype,
                       const int32_t* qo_indptr, int batch, const int64_t* prev_ids,
                       const int32_t* prev_indptr, const float* penalties, int64_t* next_ids,
                       void* stream);

/* K7 (device-side page-table update + slot allocator).  Integer, bit-exact.
 *
 * bsgl_page_table_write_rows: the prefill rows of one batch in one launch (scheduler.py:423-442:
 * zeros([B, C]) + per-request slice writes + whole-row index-put).  Row desc[0][i] of the table
 * becomes [prefix_flat[prefix_off .. +prefix) | new_slots[new_off .. +seq-prefix) | 0 ...] over
 * row_len cells.  desc is int64 [5][batch]: req_pool_idx, prefix_len, seq_len, new_off, prefix_off.
 * prefix_flat / new_slots may be NULL when no request reads them. */
int bsgl_page_table_write_rows(int32_t* req_to_token, int64_t page_table_stride, int64_t n_rows,
                               int row_len, const int64_t* desc, int batch,
                               const int32_t* prefix_flat, const int32_t* new_slots, void* stream);

/* bsgl_page_table_append: the decode step's B scalar writes (scheduler.py:486-488) in one launch:
 * req_to_token[req_pool_indices[i], seq_lens[i] - 1] = slots[i]; seq_lens already counts the new
 * token (scheduler.py:489-490). */
int bsgl_page_table_append(int32_t* req_to_token, int64_t page_table_stride, int64_t n_rows,
                           int row_len, const int64_t* req_pool_indices, const int64_t* seq_lens,
                           const int32_t* slots, int batch, void* stream);

/* The free-slot FIFO of BaseTokenToKVPool (memory_pool.py:46-102) as a ring in device memory.
 * Head, tail and count live on the host (all sizes are known there), so neither call syncs:
 *   pop : out[i] = ring[(head + i) mod capacity]            alloc, memory_pool.py:68-74
 *   push: ring[(tail + i) mod capacity] = vals[i*val_stride] free,  memory_pool.py:76-80
 * vals is int32 or int64 (vals_are_int64), e.g. a strided view of the page table. */
int bsgl_slot_ring_pop(const int32_t* ring, int64_t capacity, int64_t head, int64_t n, int32_t* out,
                       void* stream);
int bsgl_slot_ring_push(int32_t* ring, int64_t capacity, int64_t tail, int64_t n, const void* vals,
                        int64_t val_stride, int vals_are_int64, void* stream);

/* K8: o_proj GEMM with the KV-head-TP SUM all-reduce fused into its epilogue over
 * NVLink peer memory.  Replaces `o_proj(attn_output)` (qwen2.py:166,195) + `tp_all_reduce_sum`
 * (tp_runtime.py:28-33) with one kernel per rank.
 *
 * Communicator memory is the one thing this library allocates itself: it has to be exportable to the peer
 * processes (cudaMalloc + CUDA IPC), which torch's caching allocator does not promise.
 *   bsgl_comm_alloc  cudaMalloc + zero-fill + IPC export; handle is 64 bytes to hand to the peers
 *   bsgl_comm_open   map a peer's buffer into this process (peer access enabled lazily)
 *   bsgl_comm_close / bsgl_comm_free   the inverse of open / alloc
 *
 * bsgl_oproj_allreduce: out[T, hidden] = SUM_ranks( a_r[T, K_r] . w_r[hidden, K_r]^T ), identical bits on every rank.
 *   a            [n_tokens, k_local]  this rank's attention output, rows lda elements apart (bf16/fp16)
 *   w            [hidden, k_local]    this rank's column shard of o_proj.weight, rows ldw elements apart
 *   staging_ptrs [world]  every rank's staging buffer as mapped here: world * max_slots * 32 KiB
 *   out_ptrs     [world]  every rank's output buffer (out_bytes); this rank reads its own at out_ptrs[rank]
 *                         after the call (stream-ordered); valid until the rank's next call
 *   flag_ptrs    [world]  every rank's flag block, zero-initialised, bsgl_oproj_flag_bytes(max_slots, max_tiles)
 *                         bytes, laid out exactly as:
 *                           [0, W*4)            W = max_slots * 8 + max_tiles + 2 uint32 flag words; words W-2 and W-1
 *                                               of a rank's own block are the call counter (epoch) and a CTA counter,
 *                                               kept in device memory so that a captured CUDA graph numbers its
 *                                               calls like eager launches
 *                           [R, R + 256*32)     R = W*4 rounded UP to a multiple of 1024: one 4 x uint64 record of
 *                                               phase timestamps per CTA, written by EVERY call at R + 32*blockIdx.x
 *                                               (the grid never exceeds the SM count; 256 rows are reserved)
 *                         a block smaller than that is written out of bounds: size it with the query, not by hand
 *   mc_staging / mc_out   NVLS multicast views of the same staging / output regions (every rank's copy behind
 *                         one address), or both NULL: then partial tiles travel by peer stores (two-shot)
 *   timeout_s    a peer that never arrives traps the kernel after this many seconds
 * All ranks must make the same sequence of calls.  hidden % 128 == 0, k_local % 64 == 0, world <= 8.
 * Environment (read per call / once): BSGL_K8_GEMM=single|pair forces the GEMM form (128 x 128 single-CTA tiles or
 * 256 x 256 cta_group::2 pair blocks; default: pairs at world 1 when the tiles exceed one wave),
 * BSGL_K8_TILE_ORDER=grouped restores round 1's owner-by-owner visiting order.  Neither changes results. */
int bsgl_comm_alloc(int64_t bytes, void** dev_ptr, void* ipc_handle_64);
int bsgl_comm_open(const void* ipc_handle_64, void** dev_ptr);
int bsgl_comm_close(void* dev_ptr);
int bsgl_comm_free(void* dev_ptr);
int bsgl_oproj_flag_bytes(int max_slots, int max_tiles, int64_t* bytes);
int bsgl_oproj_allreduce(const void* a, int64_t lda, const void* w, int64_t ldw, int n_tokens, int hidden,
                         int k_local, int rank, int world, void* const* staging_ptrs,
                         void* const* out_ptrs, void* const* flag_ptrs, void* mc_staging, void* mc_out,
                         int max_slots, int max_tiles, int64_t out_bytes, double timeout_s, int dtype, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* BSGL_B200_H_ */
