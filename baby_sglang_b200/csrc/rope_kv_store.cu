// K5  rope_kv_store: the step immediately before the attention path, fused into the KV store.
// One launch replaces, per layer:
//   * RotaryEmbedding.forward / _apply_rope        reference qwen2.py:249-268
//       freqs = outer(positions.float(), inv_freq); cos/sin in fp32, CAST TO THE ACTIVATION DTYPE,
//       then x1*cos - x2*sin and x1*sin + x2*cos with every product and sum rounded to that dtype
//       (two torch.cat round trips over q and k)
//   * the q/k/v split of the fused projection       reference qwen2.py:285-290 (v stays a strided view)
//   * MHATokenToKVPool.set_kv_buffer                reference memory_pool.py:152-171
// q is written rotated and contiguous ([T, Hq, D], what RoPE's cat hands to attention); rotated K and
// plain V go straight into the paged slabs at loc[t] -- they never exist as separate tensors.
//
// Bit-exactness: the kernel reproduces the reference's rounding sequence (cos/sin rounded to the
// 16-bit dtype, each product rounded, then the sum rounded), with cosf/sinf being the same libdevice
// routines torch.cos/torch.sin call for fp32 CUDA tensors, and inv_freq taken from the module's own
// buffer.  HBM-bound: bytes = T * ((Hq + 2 Hkv) * D * e  [read]  +  (Hq + 2 Hkv) * D * e  [write]).
#include "common.cuh"

namespace bsgl {

constexpr int kRopeThreads = 256;

template <typename T>
__device__ __forceinline__ float rnd(float x);   // round to T and back (one torch elementwise op)
template <>
__device__ __forceinline__ float rnd<__nv_bfloat16>(float x) {
  return __bfloat162float(__float2bfloat16_rn(x));
}
template <>
__device__ __forceinline__ float rnd<__half>(float x) {
  return __half2float(__float2half_rn(x));
}

template <typename T>
__device__ __forceinline__ float2 unpack2(uint32_t v);
template <>
__device__ __forceinline__ float2 unpack2<__nv_bfloat16>(uint32_t v) {
  return __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&v));
}
template <>
__device__ __forceinline__ float2 unpack2<__half>(uint32_t v) {
  return __half22float2(*reinterpret_cast<__half2*>(&v));
}

// One CTA per token.  Every thread first issues the loads of its 16-byte chunk pairs (x1 chunk c, x2 chunk c
// of a q or k head) and v chunks; cos/sin of the token (D/2 values each, already rounded to T) are computed
// once into shared memory while those are in flight; then the pairs are rotated and stored.
template <typename T>
__global__ void __launch_bounds__(kRopeThreads) rope_kv_store_kernel(
    const T* __restrict__ qkv, int64_t row_stride, const int64_t* __restrict__ positions,
    const float* __restrict__ inv_freq, const int32_t* __restrict__ loc, T* __restrict__ k_slab,
    T* __restrict__ v_slab, int64_t n_slots, T* __restrict__ q_out, int n_q_heads, int n_kv_heads, int head_dim) {
  extern __shared__ float cs[];                  // cos[D/2], sin[D/2]
  asm volatile("griddepcontrol.launch_dependents;");   // the attention kernel may launch and set up meanwhile
  const int half = head_dim >> 1;
  const int64_t t = blockIdx.x;
  const T* row = qkv + t * row_stride;
  const int cph = half >> 3;                     // chunk pairs per rotated head
  const int n_rot = (n_q_heads + n_kv_heads) * cph;
  const int n_v = n_kv_heads * (head_dim >> 3);
  const int n_work = n_rot + n_v;
  const int64_t slab_row = (int64_t)n_kv_heads * head_dim;
  // work item w < n_rot: chunk pair (x1 chunk, x2 chunk) of a q or k head; else a 16-byte chunk of the V row
  auto fetch = [&](int w, uint4& a, uint4& b) {
    if (w < n_rot) {
      const int hd = w / cph;                    // head among q heads then k heads
      const int c = (w - hd * cph) << 3;         // first element of the chunk inside each half
      const T* src = row + (int64_t)hd * head_dim + c;
      a = *reinterpret_cast<const uint4*>(src);          // x1[c..c+7]
      b = *reinterpret_cast<const uint4*>(src + half);   // x2[c..c+7]
    } else if (w < n_work) {
      a = *reinterpret_cast<const uint4*>(row + (int64_t)(n_q_heads + n_kv_heads) * head_dim + ((w - n_rot) << 3));
    }
  };
  // The first two work items of every thread (one whole 7B row) are in flight while cos/sin are evaluated.
  uint4 a0, b0, a1, b1;
  fetch(threadIdx.x, a0, b0);
  fetch(threadIdx.x + kRopeThreads, a1, b1);
  const int64_t slot = loc ? (int64_t)loc[t] : 0;
  const float pos = (float)positions[t];         // positions.float()
  for (int i = threadIdx.x; i < half; i += kRopeThreads) {
    const float f = pos * inv_freq[i];           // torch.outer in fp32
    cs[i] = rnd<T>(cosf(f));
    cs[half + i] = rnd<T>(sinf(f));
  }
  __syncthreads();
  if (loc && (uint64_t)slot >= (uint64_t)n_slots) __trap();   // the reference's index-put device-asserts here
  auto finish = [&](int w, const uint4& a, const uint4& b) {
    if (w < n_rot) {
      const int hd = w / cph;
      const int c = (w - hd * cph) << 3;
      const uint32_t* au = &a.x;
      const uint32_t* bu = &b.x;
      uint4 o1, o2;
      uint32_t* o1u = &o1.x;
      uint32_t* o2u = &o2.x;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float2 x1 = unpack2<T>(au[i]), x2 = unpack2<T>(bu[i]);
        const float c0 = cs[c + 2 * i], c1 = cs[c + 2 * i + 1];
        const float s0 = cs[half + c + 2 * i], s1 = cs[half + c + 2 * i + 1];
        // x1 * cos - x2 * sin ; x1 * sin + x2 * cos, each op rounded to T (qwen2.py:267-268)
        const float r0 = rnd<T>(x1.x * c0) - rnd<T>(x2.x * s0);
        const float r1 = rnd<T>(x1.y * c1) - rnd<T>(x2.y * s1);
        const float i0 = rnd<T>(x1.x * s0) + rnd<T>(x2.x * c0);
        const float i1 = rnd<T>(x1.y * s1) + rnd<T>(x2.y * c1);
        o1u[i] = pack2<T>(r0, r1);
        o2u[i] = pack2<T>(i0, i1);
      }
      T* dst;
      if (hd < n_q_heads) {
        dst = q_out + (t * n_q_heads + hd) * head_dim + c;
      } else {
        if (k_slab == nullptr) return;
        dst = k_slab + slot * slab_row + (int64_t)(hd - n_q_heads) * head_dim + c;
      }
      *reinterpret_cast<uint4*>(dst) = o1;
      *reinterpret_cast<uint4*>(dst + half) = o2;
    } else if (w < n_work && v_slab != nullptr) {
      *reinterpret_cast<uint4*>(v_slab + slot * slab_row + ((w - n_rot) << 3)) = a;
    }
  };
  finish(threadIdx.x, a0, b0);
  finish(threadIdx.x + kRopeThreads, a1, b1);
  for (int w = threadIdx.x + 2 * kRopeThreads; w < n_work; w += 2 * kRopeThreads) {   // wider rows than 512 items
    fetch(w, a0, b0);
    fetch(w + kRopeThreads, a1, b1);
    finish(w, a0, b0);
    finish(w + kRopeThreads, a1, b1);
  }
}

}  // namespace bsgl

extern "C" int bsgl_rope_kv_store(const void* qkv, int64_t qkv_row_stride, const int64_t* positions,
                                  const float* inv_freq, const int32_t* loc, void* k_slab, void* v_slab,
                                  int64_t n_slots, void* q_out, int64_t n_tokens, int n_q_heads, int n_kv_heads, int head_dim,
                                  int dtype, void* stream) {
  using namespace bsgl;
  BSGL_REQUIRE(n_tokens >= 0, "bsgl_rope_kv_store: negative token count");
  if (n_tokens == 0) return BSGL_OK;
  BSGL_REQUIRE(qkv && positions && inv_freq && q_out, "bsgl_rope_kv_store: null pointer");
  BSGL_REQUIRE((k_slab == nullptr) == (v_slab == nullptr) && (k_slab == nullptr) == (loc == nullptr),
               "bsgl_rope_kv_store: k_slab, v_slab and loc must be all set (store) or all NULL (rotate only)");
  BSGL_REQUIRE(n_q_heads > 0 && n_kv_heads > 0, "bsgl_rope_kv_store: bad head counts");
  BSGL_REQUIRE(loc == nullptr || n_slots > 0, "bsgl_rope_kv_store: n_slots must be the slab's row count (S + 1)");
  BSGL_REQUIRE(head_dim > 0 && head_dim % 16 == 0 && head_dim <= 1024,
               "bsgl_rope_kv_store: head_dim must be a multiple of 16 (two 16-byte half-row chunks)");
  BSGL_REQUIRE(dtype == BSGL_BF16 || dtype == BSGL_F16, "bsgl_rope_kv_store: dtype must be bf16 or fp16");
  BSGL_REQUIRE(qkv_row_stride % 8 == 0 && (uintptr_t)qkv % 16 == 0 && (uintptr_t)q_out % 16 == 0 &&
                   (uintptr_t)k_slab % 16 == 0 && (uintptr_t)v_slab % 16 == 0,
               "bsgl_rope_kv_store: pointers and the row stride must keep 16-byte alignment");
  BSGL_REQUIRE(n_tokens < (1ll << 31), "bsgl_rope_kv_store: too many tokens");
  cudaStream_t s = (cudaStream_t)stream;
  const size_t smem = (size_t)head_dim * sizeof(float);
  if (dtype == BSGL_BF16) {
    rope_kv_store_kernel<__nv_bfloat16><<<(unsigned)n_tokens, kRopeThreads, smem, s>>>(
        (const __nv_bfloat16*)qkv, qkv_row_stride, positions, inv_freq, loc, (__nv_bfloat16*)k_slab,
        (__nv_bfloat16*)v_slab, n_slots, (__nv_bfloat16*)q_out, n_q_heads, n_kv_heads, head_dim);
  } else {
    rope_kv_store_kernel<__half><<<(unsigned)n_tokens, kRopeThreads, smem, s>>>(
        (const __half*)qkv, qkv_row_stride, positions, inv_freq, loc, (__half*)k_slab, (__half*)v_slab, n_slots,
        (__half*)q_out, n_q_heads, n_kv_heads, head_dim);
  }
  return check_launch("rope_kv_store_kernel");
}
