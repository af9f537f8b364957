// K1  kv_store: scatter this step's K and V rows into the paged slabs, one launch.
// Replaces MHATokenToKVPool.set_kv_buffer (reference memory_pool.py:152-171): int32 loc -> int64
// widen (:159-160), optional cast to the pool dtype (:163-165), two index-puts (:170-171).
//
// HBM-bound byte work: every thread moves one 16-byte chunk of a token's K row and the same chunk of its V
// row (two vector loads, two vector stores; twice the loads when narrowing from fp32), one thread per chunk.
// Algorithmic bytes: 2 * (2 * T * Hkv * D * e) + 4 * T.
#include "common.cuh"

namespace bsgl {

template <typename SrcT, typename DstT>
struct Cvt8;

template <typename T>
struct Cvt8<T, T> {   // same dtype: 16-byte copy, bit-exact
  static __device__ __forceinline__ uint4 load(const T* p) { return *reinterpret_cast<const uint4*>(p); }
};

template <typename DstT>
struct Cvt8<float, DstT> {   // fp32 -> 16-bit, round-to-nearest-even (matches torch .to())
  static __device__ __forceinline__ uint4 load(const float* p) {
    float4 a = *reinterpret_cast<const float4*>(p);
    float4 b = *reinterpret_cast<const float4*>(p + 4);
    uint4 r;
    r.x = pack2<DstT>(a.x, a.y);
    r.y = pack2<DstT>(a.z, a.w);
    r.z = pack2<DstT>(b.x, b.y);
    r.w = pack2<DstT>(b.z, b.w);
    return r;
  }
};

template <>
struct Cvt8<__half, __nv_bfloat16> {
  static __device__ __forceinline__ uint4 load(const __half* p) {
    uint4 raw = *reinterpret_cast<const uint4*>(p);
    const __half2* h = reinterpret_cast<const __half2*>(&raw);
    uint4 r;
    uint32_t* o = &r.x;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float2 f = __half22float2(h[i]);
      o[i] = pack2<__nv_bfloat16>(f.x, f.y);
    }
    return r;
  }
};

template <>
struct Cvt8<__nv_bfloat16, __half> {
  static __device__ __forceinline__ uint4 load(const __nv_bfloat16* p) {
    uint4 raw = *reinterpret_cast<const uint4*>(p);
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw);
    uint4 r;
    uint32_t* o = &r.x;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float2 f = __bfloat1622float2(h[i]);
      o[i] = pack2<__half>(f.x, f.y);
    }
    return r;
  }
};

template <typename SrcT, typename DstT>
__global__ void __launch_bounds__(256) kv_store_kernel(DstT* __restrict__ k_slab, DstT* __restrict__ v_slab,
                                                       int64_t n_slots, const int32_t* __restrict__ loc,
                                                       const SrcT* __restrict__ k_src,
                                                       const SrcT* __restrict__ v_src, int64_t n_tokens,
                                                       int n_kv_heads, int head_dim, int64_t k_row_stride,
                                                       int64_t k_head_stride, int64_t v_row_stride,
                                                       int64_t v_head_stride) {
  // the attention kernel behind us is a programmatic dependent: let it launch and set up while the rows land
  // (it reads nothing before its griddepcontrol.wait, which returns once this grid has completed)
  asm volatile("griddepcontrol.launch_dependents;");
  const int chunks_per_head = head_dim >> 3;                 // 8 elements (16 B of 16-bit dst) per chunk
  const int chunks_per_row = n_kv_heads * chunks_per_head;   // per K (or V) row
  const int64_t slab_row = (int64_t)n_kv_heads * head_dim;
  // One thread = the same 16-byte chunk of a token's K row and V row: both source loads are issued before
  // the dependent slot load is consumed, and a CTA's 256 threads cover whole rows with coalesced 16-byte
  // accesses.  No loop: the grid covers every chunk (T x chunks_per_row <= 2^31 is checked by the launcher).
  const uint32_t i = blockIdx.x * 256u + threadIdx.x;
  const uint32_t t = i / (uint32_t)chunks_per_row;
  if ((int64_t)t >= n_tokens) return;
  const int c = (int)(i - t * (uint32_t)chunks_per_row);
  const int h = c / chunks_per_head;
  const int d = (c - h * chunks_per_head) << 3;
  const uint4 kq = Cvt8<SrcT, DstT>::load(k_src + (int64_t)t * k_row_stride + h * k_head_stride + d);
  const uint4 vq = Cvt8<SrcT, DstT>::load(v_src + (int64_t)t * v_row_stride + h * v_head_stride + d);
  const int64_t slot = (int64_t)loc[t];                      // int32 -> int64 widen
  if ((uint64_t)slot >= (uint64_t)n_slots) __trap();         // the reference's index-put device-asserts here
  const int64_t off = slot * slab_row + (int64_t)h * head_dim + d;
  *reinterpret_cast<uint4*>(k_slab + off) = kq;
  *reinterpret_cast<uint4*>(v_slab + off) = vq;
}

template <typename SrcT, typename DstT>
static int launch_kv_store(void* k_slab, void* v_slab, int64_t n_slots, const int32_t* loc, const void* k_src, const void* v_src,
                           int64_t n_tokens, int n_kv_heads, int head_dim, int64_t krs, int64_t khs, int64_t vrs,
                           int64_t vhs, cudaStream_t stream) {
  const int64_t total = n_tokens * n_kv_heads * (head_dim >> 3);   // (token, 16-byte chunk) pairs
  BSGL_REQUIRE(total < (1ll << 31), "bsgl_kv_store: %lld row chunks in one call (limit 2^31)", (long long)total);
  const int64_t blocks = (total + 255) / 256;
  kv_store_kernel<SrcT, DstT><<<(unsigned)blocks, 256, 0, stream>>>(
      (DstT*)k_slab, (DstT*)v_slab, n_slots, loc, (const SrcT*)k_src, (const SrcT*)v_src, n_tokens, n_kv_heads, head_dim, krs,
      khs, vrs, vhs);
  return check_launch("kv_store_kernel");
}

}  // namespace bsgl

extern "C" int bsgl_kv_store(void* k_slab, void* v_slab, int64_t n_slots, const int32_t* loc, const void* k_src, const void* v_src,
                             int64_t n_tokens, int n_kv_heads, int head_dim, int64_t k_row_stride,
                             int64_t k_head_stride, int64_t v_row_stride, int64_t v_head_stride, int src_dtype,
                             int slab_dtype, void* stream) {
  using namespace bsgl;
  BSGL_REQUIRE(n_tokens >= 0, "bsgl_kv_store: negative token count");
  if (n_tokens == 0) return BSGL_OK;
  BSGL_REQUIRE(k_slab && v_slab && loc && k_src && v_src, "bsgl_kv_store: null pointer");
  BSGL_REQUIRE(n_slots > 0, "bsgl_kv_store: n_slots must be the slab's row count (S + 1)");
  BSGL_REQUIRE(n_kv_heads > 0 && head_dim > 0 && head_dim % 8 == 0, "bsgl_kv_store: head_dim must be a multiple of 8");
  BSGL_REQUIRE(slab_dtype == BSGL_BF16 || slab_dtype == BSGL_F16,
               "bsgl_kv_store: slab dtype must be bf16 or fp16 (fp8 KV cache is unimplemented in the reference too, "
               "memory_pool.py:167-168)");
  const int src_elt = (src_dtype == BSGL_F32) ? 4 : 2;
  const int64_t src_align = 16 / src_elt;   // elements per 16 bytes
  BSGL_REQUIRE(((uintptr_t)k_src % 16 == 0) && ((uintptr_t)v_src % 16 == 0) && ((uintptr_t)k_slab % 16 == 0) &&
                   ((uintptr_t)v_slab % 16 == 0),
               "bsgl_kv_store: pointers must be 16-byte aligned");
  BSGL_REQUIRE(k_row_stride % src_align == 0 && k_head_stride % src_align == 0 && v_row_stride % src_align == 0 &&
                   v_head_stride % src_align == 0,
               "bsgl_kv_store: source strides must keep 16-byte alignment");
  cudaStream_t s = (cudaStream_t)stream;
#define BSGL_KVS(S, D) \
  return launch_kv_store<S, D>(k_slab, v_slab, n_slots, loc, k_src, v_src, n_tokens, n_kv_heads, head_dim, k_row_stride, \
                               k_head_stride, v_row_stride, v_head_stride, s)
  if (slab_dtype == BSGL_BF16) {
    if (src_dtype == BSGL_BF16) BSGL_KVS(__nv_bfloat16, __nv_bfloat16);
    if (src_dtype == BSGL_F16) BSGL_KVS(__half, __nv_bfloat16);
    if (src_dtype == BSGL_F32) BSGL_KVS(float, __nv_bfloat16);
  } else {
    if (src_dtype == BSGL_F16) BSGL_KVS(__half, __half);
    if (src_dtype == BSGL_BF16) BSGL_KVS(__nv_bfloat16, __half);
    if (src_dtype == BSGL_F32) BSGL_KVS(float, __half);
  }
#undef BSGL_KVS
  set_error("bsgl_kv_store: unsupported source dtype %d", src_dtype);
  return BSGL_ERR_INVALID;
}
