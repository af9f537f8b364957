// K6  greedy_sample: the step immediately after the path, batched on the device.
// Replaces, for requests decoded greedily, the per-request Python loop of the reference
// (scheduler.py:509-546 `_sample_next_ids`, one `.item()` per request per step, the file carries a FIXME
// for it) and sample.py:49-60:
//   * prefill: last-token gather  logits[cumsum(extended_lens) - 1]           scheduler.py:513-521
//   * enforce_repetition_penalty_: for every DISTINCT previous token id,
//       logit = logit > 0 ? logit / penalty : logit * penalty  (in the logits' dtype)   sample.py:19-36
//   * torch.argmax over the vocabulary (first index among equal maxima; NaN counts as maximal)  sample.py:59-60
// One CTA per request streams its vocabulary row once (16-byte loads); previous ids are marked in a
// shared-memory bitmap so the penalty is applied on the fly; a block reduction picks (max, lowest index).
// HBM-bound: bytes = B * V * e  (+ 8 B per previous id).  Integer output: bit-exact.
#include "common.cuh"

namespace bsgl {

constexpr int kSampleThreads = 512;

template <typename T>
struct LogitIo;
template <>
struct LogitIo<float> {
  static constexpr int kVec = 4;
  static __device__ __forceinline__ void load(const float* p, float (&v)[8], int& n) {
    const float4 a = *reinterpret_cast<const float4*>(p);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w;
    n = 4;
  }
  static __device__ __forceinline__ float round(float x) { return x; }
};
template <>
struct LogitIo<__nv_bfloat16> {
  static constexpr int kVec = 8;
  static __device__ __forceinline__ void load(const __nv_bfloat16* p, float (&v)[8], int& n) {
    const uint4 a = *reinterpret_cast<const uint4*>(p);
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&a);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 f = __bfloat1622float2(h[i]);
      v[2 * i] = f.x; v[2 * i + 1] = f.y;
    }
    n = 8;
  }
  static __device__ __forceinline__ float round(float x) { return __bfloat162float(__float2bfloat16_rn(x)); }
};
template <>
struct LogitIo<__half> {
  static constexpr int kVec = 8;
  static __device__ __forceinline__ void load(const __half* p, float (&v)[8], int& n) {
    const uint4 a = *reinterpret_cast<const uint4*>(p);
    const __half2* h = reinterpret_cast<const __half2*>(&a);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 f = __half22float2(h[i]);
      v[2 * i] = f.x; v[2 * i + 1] = f.y;
    }
    n = 8;
  }
  static __device__ __forceinline__ float round(float x) { return __half2float(__float2half_rn(x)); }
};

// Total order used by the reduction: NaN above everything, then the value, then the LOWER index.
__device__ __forceinline__ bool better(float v, int i, float bv, int bi) {
  const bool vn = v != v, bn = bv != bv;
  if (vn != bn) return vn;
  if (!vn && v != bv) return v > bv;
  return i < bi;
}

template <typename T>
__global__ void __launch_bounds__(kSampleThreads) greedy_sample_kernel(
    const T* __restrict__ logits, int64_t row_stride, int vocab, const int32_t* __restrict__ qo_indptr,
    const int64_t* __restrict__ prev_ids, const int32_t* __restrict__ prev_indptr,
    const float* __restrict__ penalties, int64_t* __restrict__ next_ids) {
  extern __shared__ uint32_t bitmap[];                       // ceil(vocab / 32) words, then reduction scratch
  const int r = blockIdx.x;
  const int64_t row = qo_indptr ? (int64_t)qo_indptr[r + 1] - 1 : r;   // last token of the request
  const T* src = logits + row * row_stride;
  const int words = (vocab + 31) >> 5;
  float penalty = 1.f;
  int np = 0;
  if (prev_ids != nullptr) {
    penalty = penalties[r];
    np = prev_indptr[r + 1] - prev_indptr[r];
  }
  // sample.py:22-31: nothing happens for penalty <= 1 (or NaN) or an empty history
  const bool penalise = penalty > 1.f && np > 0;
  const float inv_penalty = 1.f / penalty;
  if (penalise) {
    for (int w = threadIdx.x; w < words; w += kSampleThreads) bitmap[w] = 0;
    __syncthreads();
    const int64_t* ids = prev_ids + prev_indptr[r];
    for (int i = threadIdx.x; i < np; i += kSampleThreads) {
      const int64_t t = ids[i];
      if (t >= 0 && t < vocab) atomicOr(&bitmap[t >> 5], 1u << (t & 31));
    }
    __syncthreads();
  }
  float best = -INFINITY;
  int best_i = 0x7fffffff;
  constexpr int VEC = LogitIo<T>::kVec;
  const int n_vec = vocab / VEC;
  auto consider = [&](float v, int idx) {
    if (penalise && (bitmap[idx >> 5] >> (idx & 31) & 1u))
      // sample.py:34-35, in the logits' dtype.  torch's CUDA true-division by a Python scalar multiplies by
      // the fp32 reciprocal (BinaryDivTrueKernel.cu), which is what the reference executes on a GPU.
      v = LogitIo<T>::round(v > 0.f ? v * inv_penalty : v * penalty);
    if (better(v, idx, best, best_i)) { best = v; best_i = idx; }
  };
  const bool aligned = ((uintptr_t)src % 16) == 0;
  if (aligned) {
    for (int c = threadIdx.x; c < n_vec; c += kSampleThreads) {
      float v[8];
      int n;
      LogitIo<T>::load(src + (int64_t)c * VEC, v, n);
#pragma unroll
      for (int i = 0; i < VEC; ++i) consider(v[i], c * VEC + i);
    }
    for (int idx = n_vec * VEC + threadIdx.x; idx < vocab; idx += kSampleThreads) consider((float)src[idx], idx);
  } else {
    for (int idx = threadIdx.x; idx < vocab; idx += kSampleThreads) consider((float)src[idx], idx);
  }
  // warp, then block reduction
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, best, o);
    const int oi = __shfl_xor_sync(0xffffffffu, best_i, o);
    if (better(ov, oi, best, best_i)) { best = ov; best_i = oi; }
  }
  __shared__ float wv[kSampleThreads / 32];
  __shared__ int wi[kSampleThreads / 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) { wv[warp] = best; wi[warp] = best_i; }
  __syncthreads();
  if (warp == 0) {
    best = lane < kSampleThreads / 32 ? wv[lane] : -INFINITY;
    best_i = lane < kSampleThreads / 32 ? wi[lane] : 0x7fffffff;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, best, o);
      const int oi = __shfl_xor_sync(0xffffffffu, best_i, o);
      if (better(ov, oi, best, best_i)) { best = ov; best_i = oi; }
    }
    if (lane == 0) next_ids[r] = best_i;
  }
}

}  // namespace bsgl

extern "C" int bsgl_greedy_sample(const void* logits, int64_t row_stride, int vocab, int dtype,
                                  const int32_t* qo_indptr, int batch, const int64_t* prev_ids,
                                  const int32_t* prev_indptr, const float* penalties, int64_t* next_ids,
                                  void* stream) {
  using namespace bsgl;
  BSGL_REQUIRE(batch >= 0, "bsgl_greedy_sample: negative batch");
  if (batch == 0) return BSGL_OK;
  BSGL_REQUIRE(logits && next_ids, "bsgl_greedy_sample: null pointer");
  BSGL_REQUIRE(vocab > 0 && vocab <= (1 << 20), "bsgl_greedy_sample: vocab %d out of range (1..2^20)", vocab);
  BSGL_REQUIRE((prev_ids == nullptr) == (prev_indptr == nullptr) && (prev_ids == nullptr) == (penalties == nullptr),
               "bsgl_greedy_sample: prev_ids, prev_indptr and penalties go together");
  BSGL_REQUIRE(dtype == BSGL_BF16 || dtype == BSGL_F16 || dtype == BSGL_F32, "bsgl_greedy_sample: bad dtype");
  const size_t smem = prev_ids ? (size_t)((vocab + 31) / 32) * 4 : 0;
  cudaStream_t s = (cudaStream_t)stream;
#define BSGL_GS(T)                                                                                          \
  do {                                                                                                      \
    static thread_local int configured_dev = -1;                                                            \
    int dev = 0;                                                                                            \
    BSGL_CUDA_TRY(cudaGetDevice(&dev));                                                                     \
    if (configured_dev != dev) {                                                                            \
      BSGL_CUDA_TRY(cudaFuncSetAttribute(greedy_sample_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                         (1 << 20) / 8));                                                   \
      configured_dev = dev;                                                                                 \
    }                                                                                                       \
    greedy_sample_kernel<T><<<(unsigned)batch, kSampleThreads, smem, s>>>(                                  \
        (const T*)logits, row_stride, vocab, qo_indptr, prev_ids, prev_indptr, penalties, next_ids);        \
    return check_launch("greedy_sample_kernel");                                                            \
  } while (0)
  if (dtype == BSGL_BF16) BSGL_GS(__nv_bfloat16);
  if (dtype == BSGL_F16) BSGL_GS(__half);
  BSGL_GS(float);
#undef BSGL_GS
}
