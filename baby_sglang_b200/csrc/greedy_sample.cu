// K6  greedy_sample: the step immediately after the path, batched on the device.
// Replaces, for requests decoded greedily, the per-request Python loop of the reference
// (scheduler.py:509-546 `_sample_next_ids`, one `.item()` per request per step, the file carries a FIXME
// for it) and sample.py:49-60:
//   * prefill: last-token gather  logits[cumsum(extended_lens) - 1]           scheduler.py:513-521
//   * enforce_repetition_penalty_: for every DISTINCT previous token id,
//       logit = logit > 0 ? logit / penalty : logit * penalty  (in the logits' dtype)   sample.py:19-36
//   * torch.argmax over the vocabulary (first index among equal maxima; NaN counts as maximal)  sample.py:59-60
// A thread-block cluster of up to 8 CTAs per request streams the vocabulary row once (16-byte loads, four in
// flight per thread); previous ids are marked in a shared-memory bitmap (each CTA its own slice) so the penalty
// is applied on the fly; warp -> block -> cluster (distributed shared memory) reduction picks (max, lowest index).
// HBM-bound: bytes = B * V * e  (+ 8 B per previous id).  Integer output: bit-exact.
#include <cooperative_groups.h>

#include "common.cuh"

namespace bsgl {

constexpr int kSampleThreads = 256;
constexpr int kSampleMaxSplits = 8;   // portable cluster size

// 16-byte streaming load: every logit is read exactly once
__device__ __forceinline__ uint4 ld_stream16(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}

template <typename T>
struct LogitIo;
template <>
struct LogitIo<float> {
  static constexpr int kVec = 4;
  static __device__ __forceinline__ void unpack(const uint4& a, float (&v)[8]) {
    v[0] = __uint_as_float(a.x); v[1] = __uint_as_float(a.y); v[2] = __uint_as_float(a.z); v[3] = __uint_as_float(a.w);
  }
  static __device__ __forceinline__ float round(float x) { return x; }
};
template <>
struct LogitIo<__nv_bfloat16> {
  static constexpr int kVec = 8;
  static __device__ __forceinline__ void unpack(const uint4& a, float (&v)[8]) {
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&a);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 f = __bfloat1622float2(h[i]);
      v[2 * i] = f.x; v[2 * i + 1] = f.y;
    }
  }
  static __device__ __forceinline__ float round(float x) { return __bfloat162float(__float2bfloat16_rn(x)); }
};
template <>
struct LogitIo<__half> {
  static constexpr int kVec = 8;
  static __device__ __forceinline__ void unpack(const uint4& a, float (&v)[8]) {
    const __half2* h = reinterpret_cast<const __half2*>(&a);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 f = __half22float2(h[i]);
      v[2 * i] = f.x; v[2 * i + 1] = f.y;
    }
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
  extern __shared__ uint32_t bitmap[];                       // this CTA's slice of the previous-id bitmap
  namespace cg = cooperative_groups;
  cg::cluster_group cluster = cg::this_cluster();
  const int splits = (int)gridDim.x;                         // == cluster size: the CTAs of one vocabulary row
  const int part = (int)blockIdx.x;                          // == rank in the cluster
  const int r = blockIdx.y;
  // Shared memory of another CTA may only be touched once that CTA is known to have started: every thread
  // arrives here and waits just before the remote stores below, so the barrier's latency hides behind the scan.
  cluster.barrier_arrive();
  const int64_t row = qo_indptr ? (int64_t)qo_indptr[r + 1] - 1 : r;   // last token of the request
  const T* src = logits + row * row_stride;
  constexpr int VEC = LogitIo<T>::kVec;
  // slice [e_lo, e_hi) of the row, a whole number of 16-byte vectors except for the row's tail
  const int n_vec = (vocab + VEC - 1) / VEC;
  const int per = (n_vec + splits - 1) / splits;
  const int e_lo = min(vocab, part * per * VEC), e_hi = min(vocab, (part + 1) * per * VEC);
  float penalty = 1.f;
  int np = 0;
  if (prev_ids != nullptr) {
    penalty = penalties[r];
    np = prev_indptr[r + 1] - prev_indptr[r];
  }
  // sample.py:22-31: nothing happens for penalty <= 1 (or NaN) or an empty history
  const bool penalise = penalty > 1.f && np > 0;
  const float inv_penalty = 1.f / penalty;
  const int w0 = e_lo >> 5, words = ((e_hi + 31) >> 5) - w0;
  if (penalise) {
    for (int w = threadIdx.x; w < words; w += kSampleThreads) bitmap[w] = 0;
    __syncthreads();
    const int64_t* ids = prev_ids + prev_indptr[r];
    for (int i = threadIdx.x; i < np; i += kSampleThreads) {
      const int64_t t = ids[i];
      if (t >= e_lo && t < e_hi) atomicOr(&bitmap[(int)(t >> 5) - w0], 1u << (t & 31));
    }
    __syncthreads();
  }
  float best = -INFINITY;
  int best_i = 0x7fffffff, nan_i = 0x7fffffff;
  auto consider = [&](float v, int idx) {
    if (penalise && (bitmap[(idx >> 5) - w0] >> (idx & 31) & 1u))
      // sample.py:34-35, in the logits' dtype.  torch's CUDA true-division by a Python scalar multiplies by
      // the fp32 reciprocal (BinaryDivTrueKernel.cu), which is what the reference executes on a GPU.
      v = LogitIo<T>::round(v > 0.f ? v * inv_penalty : v * penalty);
    // A thread meets its indices in increasing order, so inside a thread the total order of `better` reduces to
    // "strictly greater keeps the first of equals" plus "the first NaN" (tracked apart, NaN beats every value):
    // 5 instructions per element instead of the full comparison.
    if (v > best) { best = v; best_i = idx; }
    if (v != v) nan_i = min(nan_i, idx);
  };
  const bool aligned = ((uintptr_t)src % 16) == 0;
  if (aligned) {
    // four independent 16-byte loads per thread in flight before any of them is compared
    constexpr int U = 4;
    const int c_lo = e_lo / VEC, c_hi = e_hi / VEC;          // whole vectors of the slice
    for (int c0 = c_lo + threadIdx.x; c0 < c_hi; c0 += U * kSampleThreads) {
      uint4 raw[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int c = c0 + u * kSampleThreads;
        raw[u] = c < c_hi ? ld_stream16(src + (int64_t)c * VEC) : make_uint4(0, 0, 0, 0);
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int c = c0 + u * kSampleThreads;
        if (c < c_hi) {
          float v[8];
          LogitIo<T>::unpack(raw[u], v);
#pragma unroll
          for (int i = 0; i < VEC; ++i) consider(v[i], c * VEC + i);
        }
      }
    }
    for (int idx = c_hi * VEC + threadIdx.x; idx < e_hi; idx += kSampleThreads) consider((float)src[idx], idx);
  } else {
    for (int idx = e_lo + threadIdx.x; idx < e_hi; idx += kSampleThreads) consider((float)src[idx], idx);
  }
  if (best_i == 0x7fffffff) {
    // nothing exceeded -inf: the thread's candidate is its FIRST element (a row of -inf answers index 0)
    const int c_first = e_lo / VEC + (int)threadIdx.x;
    int f = aligned ? (c_first < e_hi / VEC ? c_first * VEC : (e_hi / VEC) * VEC + (int)threadIdx.x)
                    : e_lo + (int)threadIdx.x;
    if (f < e_hi) best_i = f;
  }
  if (nan_i != 0x7fffffff) { best = __int_as_float(0x7fc00000); best_i = nan_i; }
  // warp, then block, then cluster reduction (a total order: the result does not depend on the split)
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, best, o);
    const int oi = __shfl_xor_sync(0xffffffffu, best_i, o);
    if (better(ov, oi, best, best_i)) { best = ov; best_i = oi; }
  }
  __shared__ float wv[kSampleThreads / 32];
  __shared__ int wi[kSampleThreads / 32];
  __shared__ float cv[kSampleMaxSplits];                     // rank 0's copy collects one candidate per CTA
  __shared__ int ci[kSampleMaxSplits];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) { wv[warp] = best; wi[warp] = best_i; }
  cluster.barrier_wait();                                    // all CTAs of the row are running (arrived at entry)
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
    if (lane == 0) {
      cluster.map_shared_rank(cv, 0)[part] = best;           // distributed shared memory of the row's first CTA
      cluster.map_shared_rank(ci, 0)[part] = best_i;
    }
  }
  cluster.sync();
  if (part == 0 && threadIdx.x == 0) {
    best = cv[0];
    best_i = ci[0];
    for (int s = 1; s < splits; ++s)
      if (better(cv[s], ci[s], best, best_i)) { best = cv[s]; best_i = ci[s]; }
    next_ids[r] = best_i;
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
  // A vocabulary row is split over a thread-block cluster so that the grid fills the machine (8 CTAs of 256
  // threads per SM) even at small batch; the cluster's candidates meet in the first CTA's shared memory.
  int sms = 148, dev = 0;
  BSGL_CUDA_TRY(cudaGetDevice(&dev));
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  int splits = sms * 8 / batch;
  splits = splits < 1 ? 1 : (splits > kSampleMaxSplits ? kSampleMaxSplits : splits);
  const int vec = dtype == BSGL_F32 ? 4 : 8;
  const int per = ((vocab + vec - 1) / vec + splits - 1) / splits * vec;   // elements per slice
  const size_t smem = prev_ids ? (size_t)(per / 32 + 2) * 4 : 0;
  BSGL_REQUIRE(batch <= 65535, "bsgl_greedy_sample: batch %d > 65535", batch);
  cudaStream_t s = (cudaStream_t)stream;
#define BSGL_GS(T)                                                                                          \
  do {                                                                                                      \
    static thread_local int configured_dev = -1;                                                            \
    if (configured_dev != dev) {                                                                            \
      BSGL_CUDA_TRY(cudaFuncSetAttribute(greedy_sample_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                         (1 << 20) / 8 + 8));                                               \
      configured_dev = dev;                                                                                 \
    }                                                                                                       \
    cudaLaunchConfig_t cfg = {};                                                                            \
    cfg.gridDim = dim3((unsigned)splits, (unsigned)batch);                                                  \
    cfg.blockDim = dim3(kSampleThreads);                                                                    \
    cfg.dynamicSmemBytes = smem;                                                                            \
    cfg.stream = s;                                                                                         \
    cudaLaunchAttribute attr[1];                                                                            \
    attr[0].id = cudaLaunchAttributeClusterDimension;                                                       \
    attr[0].val.clusterDim.x = (unsigned)splits;                                                            \
    attr[0].val.clusterDim.y = 1;                                                                           \
    attr[0].val.clusterDim.z = 1;                                                                           \
    cfg.attrs = attr;                                                                                       \
    cfg.numAttrs = 1;                                                                                       \
    BSGL_CUDA_TRY(cudaLaunchKernelEx(&cfg, greedy_sample_kernel<T>, (const T*)logits, row_stride, vocab,    \
                                     qo_indptr, prev_ids, prev_indptr, penalties, next_ids));               \
    return check_launch("greedy_sample_kernel");                                                            \
  } while (0)
  if (dtype == BSGL_BF16) BSGL_GS(__nv_bfloat16);
  if (dtype == BSGL_F16) BSGL_GS(__half);
  BSGL_GS(float);
#undef BSGL_GS
}
