// K8  o_proj GEMM with the tensor-parallel SUM all-reduce fused into its epilogue (sm_100a, NVLink peer memory).
//
// Replaces, under KV-head tensor parallelism: `o_proj(attn_output)`
// (qwen2.py:166,195; bias-free, weight [hidden, Hq*D]) on this rank's column shard followed by
// `tp_all_reduce_sum` (tp_runtime.py:28-33, dist.all_reduce over NCCL) -- two kernels, a 2 x [T, hidden]
// HBM round trip and an NCCL launch per layer -- by ONE kernel per rank:
//
//   phase 1  partial[T, hidden] = attn_out[T, K_r] . W_r^T on tcgen05 tensor cores: TMA (2-D tensor maps,
//            128B swizzle) feeds a 6-stage shared-memory ring, one thread issues tcgen05.mma M128 N128 K16
//            into a double-buffered TMEM accumulator, and the epilogue warps push each finished 128x128
//            tile as bf16 straight from registers into the OWNER rank's staging buffer over NVLink
//            (peer-mapped memory, 512-byte coalesced bursts), then publish a per-(tile, source) flag with a
//            system-scope release.  Every rank visits the tiles in the same order: the partials of a tile reach
//            its owner together, the 148 tiles in flight per rank spread over all owners and links.
//   phase 2  the owner of a tile (tile % world) waits for its `world` partials, adds them in fp32 in rank
//            order (the result is bit-identical everywhere because only the owner computes it), rounds once
//            to bf16 and stores the final tile into EVERY rank's output buffer, then raises that tile's
//            done-flag on every rank.
//   phase 3  CTA 0 waits until every tile of the local output has been delivered.
//
// Two exchange paths, chosen by the communicator:
//   * NVLS (multicast mapping of the communicator buffers through the NVSwitch): phase 1 keeps each partial tile
//     in the rank's OWN staging (no wire traffic) and only raises a flag at the owner; the owner then reads the
//     tile with `multimem.ld_reduce` -- the switch sums the `world` copies in fp32 on the way -- and broadcasts
//     the result with `multimem.st`.  Per rank and direction that is ~T x hidden x 2 B on the wire (10.5 MB at
//     C5) instead of 2 (t-1)/t x that twice over (18.4 MB): one-shot instead of two-shot.  All ranks visit the
//     tiles in the same order, so a tile completes everywhere at about the same time and the reduction trails
//     the GEMM by a tile.
//   * peer stores (no multicast support): phases 1/2 as written above, two-shot.
//
// The transfer of tile i overlaps the MMAs of tile i+1 (TMEM double buffer) and the whole exchange overlaps
// the GEMM tail; no NCCL call, no separate reduction pass over HBM.  Flags carry the call's epoch, a counter in
// DEVICE memory that the last CTA of every launch advances (so a CUDA-graph replay numbers its calls like eager
// launches do; every rank makes the same sequence of calls).  Buffers are single: a rank cannot finish call n
// before every owner has read all of call n's partials and delivered all of its tiles, and the consumer of call
// n's output is stream-ordered before the rank's own contribution to call n+1, which every write of call n+1
// into this rank's buffers depends on.
//
// Bound: NVLink/NVSwitch bandwidth.
#include <cuda.h>

#include <cstdlib>
#include <mutex>
#include <unordered_map>

#include "tcgen05.cuh"

namespace bsgl {

constexpr int kOpBM = 128, kOpBN = 128, kOpBK = 64;
constexpr int kOpStages = 6;
constexpr int kOpStageBytes = (kOpBM + kOpBN) * kOpBK * 2;   // 32 KB: A tile + B tile, bf16
constexpr int kOpThreads = 320;                              // warp 0 TMA, warp 1 MMA, warps 2-5 epilogue, 6-9 owner reduce
constexpr int kOpTileBytes = kOpBM * kOpBN * 2;              // one bf16 partial tile in staging
constexpr int kOpMaxWorld = 8;
constexpr int kOpSmemBytes = kOpStages * kOpStageBytes + 1024 /*alignment slack*/ + 256 /*barriers*/;

struct OprojComm {
  uint8_t* staging[kOpMaxWorld];    // per rank: [world * max_slots] partial tiles (chunk-major 16-byte layout)
  uint8_t* out[kOpMaxWorld];        // per rank: [max_tokens * hidden] bf16
  uint32_t* flags[kOpMaxWorld];     // per rank: arrive[max_slots][8], done[max_tiles], {epoch, finished CTAs}, then a
                                    // 4 x u64 phase-timestamp record per CTA (globaltimer; read by comm.trace())
  uint8_t* mc_staging;              // multicast (NVLS) views of the same two regions; nullptr = peer-store path
  uint8_t* mc_out;
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint64_t globaltimer_ns() {
  uint64_t t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
// Spin until *p >= want (epochs are monotonic).  A peer that never shows up traps this rank after `timeout_ns`
// (communicator setting, default 30 s: first-call lazy peer mapping, a host-side hiccup on one rank or a profiler
// can delay a peer by seconds) instead of hanging the GPU.
__device__ __forceinline__ void wait_flag_ge(const uint32_t* p, uint32_t want, uint64_t timeout_ns) {
  if (ld_acquire_sys(p) >= want) return;
  const uint64_t t0 = globaltimer_ns();
  while (ld_acquire_sys(p) < want) {
    __nanosleep(64);
    if (globaltimer_ns() - t0 > timeout_ns) __trap();
  }
}

// NVLS: one 16-byte load that returns the element-wise SUM over every rank's copy (fp32 accumulation inside the
// switch), and one 16-byte store replicated into every rank's copy.
template <typename T>
__device__ __forceinline__ uint4 multimem_ld_reduce_add(const void* mc_ptr) {
  uint4 r;
  if constexpr (std::is_same<T, __nv_bfloat16>::value)
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(mc_ptr) : "memory");
  else
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.f16x2 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(mc_ptr) : "memory");
  return r;
}
__device__ __forceinline__ void multimem_st(void* mc_ptr, uint4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc_ptr), "r"(v.x), "r"(v.y), "r"(v.z),
               "r"(v.w) : "memory");
}

__device__ __forceinline__ void tma_load_2d(uint32_t dst_smem, const CUtensorMap* tmap, int c0, int c1, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
          dst_smem),
      "l"(reinterpret_cast<uint64_t>(tmap)), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}
// CTA-pair form: the bytes complete on the LEADER's barrier (the barrier at the same offset in the even CTA of the
// pair: bit 24 of the shared-window address selects the CTA), whichever CTA of the pair issues the load.
__device__ __forceinline__ void tma_load_2d_pair(uint32_t dst_smem, const CUtensorMap* tmap, int c0, int c1, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
          dst_smem),
      "l"(reinterpret_cast<uint64_t>(tmap)), "r"(bar & 0xFEFFFFFFu), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* tmap) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(tmap)) : "memory");
}

// Local visiting order: position i -> global tile id.  Owner groups (rank+1, rank+2, ..., rank) in turn;
// owner o holds tiles o, o + world, o + 2 world, ...
__device__ __forceinline__ int tile_at(int i, int n_tiles, int rank, int world) {
  for (int g = 0; g < world; ++g) {
    const int o = (rank + 1 + g) % world;
    const int cnt = (n_tiles - o + world - 1) / world;
    if (i < cnt) return o + world * i;
    i -= cnt;
  }
  return -1;
}


// Phase-2 worker: sum the `world` partial copies of one owned 128x128 tile (fp32, rank order), round once and
// store the result into every rank's output.  IPB items (16-byte chunks) per thread are loaded for all sources
// before the first add, so 16 loads per thread are in flight whatever the world size (W bounds `world`).
template <typename T, int W>
__device__ __forceinline__ void reduce_owned_tile(const OprojComm& comm, const uint8_t* part0, int64_t src_stride,
                                                  int world, int etid, int m0, int n0, int M, int N,
                                                  int64_t out_base) {
  constexpr int IPB = 16 / W;
#pragma unroll 1
  for (int it0 = 0; it0 < 16; it0 += IPB) {
    uint4 v[IPB][W];
#pragma unroll
    for (int u = 0; u < IPB; ++u) {
      const int item = (it0 + u) * 128 + etid;
      const int c = item & 15, r = item >> 4;                // 16 lanes cover one 256-byte output row segment
#pragma unroll
      for (int s = 0; s < W; ++s)   // unconditional (a source past `world` re-reads source 0 and is not summed):
        v[u][s] = __ldcg(reinterpret_cast<const uint4*>(part0 + (s < world ? s : 0) * src_stride) + c * 128 + r);
    }
#pragma unroll
    for (int u = 0; u < IPB; ++u) {
      const int item = (it0 + u) * 128 + etid;
      const int c = item & 15, r = item >> 4;
      float sum[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int s = 0; s < W; ++s) {                          // fixed order: independent of arrival order
        if (s < world) {
          const uint32_t w4[4] = {v[u][s].x, v[u][s].y, v[u][s].z, v[u][s].w};
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            float lo, hi;
            if constexpr (std::is_same<T, __nv_bfloat16>::value) {
              lo = __uint_as_float(w4[e] << 16);
              hi = __uint_as_float(w4[e] & 0xffff0000u);
            } else {
              const __half2 h = *reinterpret_cast<const __half2*>(&w4[e]);
              lo = __low2float(h);
              hi = __high2float(h);
            }
            sum[2 * e] += lo;
            sum[2 * e + 1] += hi;
          }
        }
      }
      if (m0 + r < M) {
        uint4 pk;
        pk.x = pack2<T>(sum[0], sum[1]);
        pk.y = pack2<T>(sum[2], sum[3]);
        pk.z = pack2<T>(sum[4], sum[5]);
        pk.w = pack2<T>(sum[6], sum[7]);
        const int64_t off = out_base + ((int64_t)(m0 + r) * N + n0 + c * 8) * 2;
#pragma unroll
        for (int p = 0; p < W; ++p)
          if (p < world) *reinterpret_cast<uint4*>(comm.out[p] + off) = pk;
      }
    }
  }
}

// Phase-2 worker, NVLS path: the switch sums the `world` copies of the owned tile on the way in and replicates the
// result on the way out.  16 independent 16-byte reductions per thread are issued before the first store.
template <typename T>
__device__ __forceinline__ void reduce_owned_tile_nvls(const OprojComm& comm, int tile, int etid, int m0, int n0,
                                                       int M, int N) {
  const uint4* src = reinterpret_cast<const uint4*>(comm.mc_staging + (int64_t)tile * kOpTileBytes);
  uint4 v[16];
#pragma unroll
  for (int u = 0; u < 16; ++u) {
    const int item = u * 128 + etid;
    const int c = item & 15, r = item >> 4;
    v[u] = multimem_ld_reduce_add<T>(src + c * 128 + r);
  }
#pragma unroll
  for (int u = 0; u < 16; ++u) {
    const int item = u * 128 + etid;
    const int c = item & 15, r = item >> 4;
    if (m0 + r < M) multimem_st(comm.mc_out + ((int64_t)(m0 + r) * N + n0 + c * 8) * 2, v[u]);
  }
}

// PAIR: two CTAs of a (2,1,1) cluster compute one 256 x 256 output block with tcgen05.mma.cta_group::2 (M 256,
// N 256): each CTA loads its own 128 rows of A and HALF of the weight tile (128 of the 256 rows) per k-block -- half
// the operand bytes through L2 per flop of the 128 x 128 single-CTA tiles, and the per-instruction cost of the MMA
// (43 clk + N / 2) is amortised over N = 256.  Each CTA still owns plain 128 x 128 output tiles (two per
// block, side by side), so the exchange protocol is unchanged.
template <typename T, bool PAIR>
__global__ void __launch_bounds__(kOpThreads, 1)
oproj_allreduce_kernel(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_b,
                       const __grid_constant__ OprojComm comm, int M, int N, int K, int rank, int world, int max_slots,
                       int max_tiles, uint64_t timeout_ns, int grouped_order) {
  extern __shared__ uint8_t smem_dyn[];
  const uint32_t raw = smem_u32(smem_dyn);
  const uint32_t base = (raw + 1023u) & ~1023u;          // SWIZZLE_128B tiles need 1024-byte alignment
  uint8_t* smem = smem_dyn + (base - raw);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_nt = N / kOpBN, n_mt = (M + kOpBM - 1) / kOpBM, n_tiles = n_mt * n_nt, n_kb = K / kOpBK;
  const bool nvls = comm.mc_staging != nullptr;
  uint32_t* ctl = comm.flags[rank] + max_slots * kOpMaxWorld + max_tiles;   // {epoch of the last finished call, CTAs done}
  const uint32_t epoch = *reinterpret_cast<volatile uint32_t*>(ctl) + 1u;   // advanced by the last CTA of each launch
  // phase timestamps of this CTA: {entry, last partial tile published, last owned tile delivered, exit}
  uint64_t* trace = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(comm.flags[rank]) +
                                                ((((size_t)max_slots * kOpMaxWorld + max_tiles + 2) * 4 + 1023) & ~(size_t)1023)) +
                    4 * blockIdx.x;
  if (threadIdx.x == 0) {
    trace[0] = globaltimer_ns();
    trace[1] = trace[2] = 0;
  }

  // PAIR work unit q = (row pair qm, column pair qn): CTA `crank` of the pair owns rows of tile row 2 qm + crank and
  // the two tile columns 2 qn, 2 qn + 1.
  const uint32_t crank = PAIR ? tc::cluster_ctarank() : 0u;
  const int n_qn = n_nt / 2, n_q = ((n_mt + 1) / 2) * n_qn;
  const int unit0 = PAIR ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
  const int unit_step = PAIR ? (int)(gridDim.x >> 1) : (int)gridDim.x;
  const int n_units = PAIR ? n_q : n_tiles;
  constexpr int kAccCols = PAIR ? 2 * kOpBN : kOpBN;                 // TMEM columns of one accumulator buffer

  const uint32_t bar0 = base + kOpStages * kOpStageBytes;
  const uint32_t full0 = bar0, empty0 = bar0 + 8 * kOpStages, accf0 = empty0 + 8 * kOpStages, acce0 = accf0 + 16;
  const uint32_t peerf0 = acce0 + 16;                                // leader: the peer CTA's stage has landed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + kOpStages * kOpStageBytes + 8 * (3 * kOpStages + 4));
  if (threadIdx.x == 0) {
    for (int s = 0; s < kOpStages; ++s) {
      mbar_init(full0 + 8 * s, 1);      // producer's arrive.expect_tx; TMA completes the bytes
      mbar_init(empty0 + 8 * s, 1);     // tcgen05.commit (multicast to both CTAs of a pair)
      mbar_init(peerf0 + 8 * s, 1);     // PAIR: remote arrive by the peer's relay warp
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(accf0 + 8 * b, 1);      // tcgen05.commit after a tile's last MMA
      // the epilogue threads, once the accumulator is in registers (PAIR: one arrive per epilogue warp of BOTH CTAs)
      mbar_init(acce0 + 8 * b, PAIR ? 8 : 128);
    }
    fence_mbar_init();
  }
  if constexpr (PAIR) tc::cluster_sync();                            // the peer's barriers exist before anyone signals them
  if (warp == 1) {
    if constexpr (PAIR) {
      tc::tmem_alloc2(smem_u32(tmem_slot), 512);
      tc::tmem_relinquish2();
    } else {
      tc::tmem_alloc(smem_u32(tmem_slot), 256);
      tc::tmem_relinquish();
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = *tmem_slot;

  if (warp == 0) {
    // ======================================================== TMA producer
    if (lane == 0) {
      tma_prefetch_desc(&tm_a);
      tma_prefetch_desc(&tm_b);
      uint32_t it = 0;
      for (int i = unit0; i < n_units; i += unit_step) {
        int m0, n0;
        if constexpr (PAIR) {
          m0 = (2 * (i / n_qn) + (int)crank) * kOpBM;             // this CTA's 128 rows of the 256-row block
          n0 = (2 * (i % n_qn)) * kOpBN + (int)crank * kOpBN;     // this CTA's half of the 256 weight rows
        } else {
          const int tile = (nvls || !grouped_order) ? i : tile_at(i, n_tiles, rank, world);
          m0 = (tile / n_nt) * kOpBM;
          n0 = (tile % n_nt) * kOpBN;
        }
        for (int kb = 0; kb < n_kb; ++kb, ++it) {
          const uint32_t s = it % kOpStages, ph = (it / kOpStages) & 1u;
          mbar_wait_guarded(empty0 + 8 * s, ph ^ 1u);
          const uint32_t a_dst = base + s * kOpStageBytes;
          if constexpr (PAIR) {
            // both CTAs' bytes complete on the leader's barrier, which expects the two stages together
            if (crank == 0) mbar_arrive_expect_tx(full0 + 8 * s, 2 * kOpStageBytes);
            tma_load_2d_pair(a_dst, &tm_a, kb * kOpBK, m0, full0 + 8 * s);               // rows past M arrive as zeros
            tma_load_2d_pair(a_dst + kOpBM * kOpBK * 2, &tm_b, kb * kOpBK, n0, full0 + 8 * s);
          } else {
            mbar_arrive_expect_tx(full0 + 8 * s, kOpStageBytes);
            tma_load_2d(a_dst, &tm_a, kb * kOpBK, m0, full0 + 8 * s);                    // rows past M arrive as zeros
            tma_load_2d(a_dst + kOpBM * kOpBK * 2, &tm_b, kb * kOpBK, n0, full0 + 8 * s);
          }
        }
      }
    }
  } else if (warp == 1) {
    // ======================================================== MMA issuer (warp-uniform loop, one elected lane)
    constexpr uint32_t idesc = PAIR ? tc::instr_desc<T>(2 * kOpBM, 2 * kOpBN, 0, 0) : tc::instr_desc<T>(kOpBM, kOpBN, 0, 0);
    uint32_t it = 0, tcount = 0;
    if (PAIR && crank != 0) {
      // the peer CTA of a pair issues nothing: the leader's MMAs read both CTAs' shared memory and write both TMEMs
    } else {
      for (int i = unit0; i < n_units; i += unit_step, ++tcount) {
        const uint32_t buf = tcount & 1u;
        mbar_wait_guarded(acce0 + 8 * buf, ((tcount >> 1) & 1u) ^ 1u);
        tc::fence_after_sync();
        for (int kb = 0; kb < n_kb; ++kb, ++it) {
          const uint32_t s = it % kOpStages, ph = (it / kOpStages) & 1u;
          mbar_wait_guarded(full0 + 8 * s, ph);
          tc::fence_after_sync();
          const uint32_t a = base + s * kOpStageBytes, b = a + kOpBM * kOpBK * 2;
          if (tc::elect_one()) {
#pragma unroll
            for (int kk = 0; kk < kOpBK / 16; ++kk) {
              if constexpr (PAIR)
                tc::mma_ss2(tmem + buf * kAccCols, tc::smem_desc_sw128(a + kk * 32, 16, 1024),
                            tc::smem_desc_sw128(b + kk * 32, 16, 1024), idesc, (kb | kk) != 0 ? 1u : 0u);
              else
                tc::mma_ss(tmem + buf * kAccCols, tc::smem_desc_sw128(a + kk * 32, 16, 1024),
                           tc::smem_desc_sw128(b + kk * 32, 16, 1024), idesc, (kb | kk) != 0 ? 1u : 0u);
            }
            if constexpr (PAIR) {
              tc::commit2(empty0 + 8 * s, 0b11);                       // both CTAs' stages are free again
              if (kb == n_kb - 1) tc::commit2(accf0 + 8 * buf, 0b11);  // both CTAs' accumulators are complete
            } else {
              tc::commit(empty0 + 8 * s);
              if (kb == n_kb - 1) tc::commit(accf0 + 8 * buf);
            }
          }
          __syncwarp();
        }
      }
    }
  } else if (warp < 6) {
    // ======================================================== epilogue: push partial tiles to their owners
    const int q = warp & 3;                                   // TMEM lane quadrant of this warp
    const int row = q * 32 + lane;
    const int etid = (warp - 2) * 32 + lane;                  // 0..127
    const uint32_t lane_base = (uint32_t)(q * 32) << 16;
    uint32_t tcount = 0;
    const uint32_t leader_acce0 = PAIR ? tc::mapa(acce0, 0) : 0u;
    constexpr int kHalves = PAIR ? 2 : 1;                      // 128 x 128 output tiles per accumulator buffer
    for (int i = unit0; i < n_units; i += unit_step, ++tcount) {
      const uint32_t buf = tcount & 1u;
      mbar_wait_guarded(accf0 + 8 * buf, (tcount >> 1) & 1u);
      tc::fence_after_sync();
      uint32_t acc[4][32];
      auto tile_of = [&](int half) {
        if constexpr (PAIR) return (2 * (i / n_qn) + (int)crank) * n_nt + 2 * (i % n_qn) + half;
        return (nvls || !grouped_order) ? i : tile_at(i, n_tiles, rank, world);
      };
      // odd number of tile rows: the peer CTA's rows of the last row pair do not exist
      const bool valid = !PAIR || (2 * (i / n_qn) + (int)crank) < n_mt;
      auto load_half = [&](int half) {
#pragma unroll
        for (int c = 0; c < 4; ++c) tc::ld32(tmem + lane_base + buf * kAccCols + half * kOpBN + c * 32, acc[c]);
        tc::wait_ld();
      };
      auto release_acc = [&]() {                               // the MMA warp may start the unit after next
        tc::fence_before_sync();
        if constexpr (PAIR) {
          __syncwarp();
          if (lane == 0) tc::mbar_arrive_cluster(leader_acce0 + 8 * buf);
        } else {
          mbar_arrive(acce0 + 8 * buf);
        }
      };
      // staging layout: 16-byte chunk c of row r at (c * 128 + r) * 16 -> a warp store is one 512-byte burst
      // NVLS: the partial stays in this rank's own staging (the owner's multimem.ld_reduce pulls it through the
      // switch); peer-store path: straight into the owner's staging over NVLink
      auto store_tile = [&](int tile) {
        const int owner = tile % world, slot = tile / world;
        uint4* dst = reinterpret_cast<uint4*>(
            nvls ? comm.staging[rank] + (int64_t)tile * kOpTileBytes
                 : comm.staging[owner] + ((int64_t)rank * max_slots + slot) * kOpTileBytes);
#pragma unroll
        for (int c = 0; c < 16; ++c) {
          const uint32_t* v = &acc[c >> 2][(c & 3) * 8];
          uint4 pk;
          pk.x = pack2<T>(__uint_as_float(v[0]), __uint_as_float(v[1]));
          pk.y = pack2<T>(__uint_as_float(v[2]), __uint_as_float(v[3]));
          pk.z = pack2<T>(__uint_as_float(v[4]), __uint_as_float(v[5]));
          pk.w = pack2<T>(__uint_as_float(v[6]), __uint_as_float(v[7]));
          dst[c * 128 + row] = pk;
        }
      };
      auto publish_tile = [&](int tile) {
        named_bar_sync(1, 128);    // bar.sync orders the 128 threads' stores before thread 0's release (cumulative)
        if (etid == 0) st_release_sys(comm.flags[tile % world] + (tile / world) * kOpMaxWorld + rank, epoch);
      };
      load_half(0);
      if constexpr (PAIR) {
        // the accumulator is handed back before the first tile's (slow, system-scope) release: the leader's next
        // block waits for BOTH CTAs' epilogues
        if (valid) store_tile(tile_of(0));
        load_half(1);
        release_acc();
        if (valid) {
          publish_tile(tile_of(0));
          store_tile(tile_of(1));
          publish_tile(tile_of(1));
        }
      } else {
        release_acc();
        store_tile(tile_of(0));
        publish_tile(tile_of(0));
      }
    }
    if (etid == 0) trace[1] = globaltimer_ns();

    // ---- phase 3: the local output is complete once every owner delivered every tile
    if (blockIdx.x == 0) {
      const uint32_t* my_flags = comm.flags[rank];
      const int done_off = max_slots * kOpMaxWorld;
      for (int tile = etid; tile < n_tiles; tile += 128) wait_flag_ge(my_flags + done_off + tile, epoch, timeout_ns);
    }
  } else {
    // ======================================================== owner reduce (phase 2), concurrent with phase 1:
    // owned tiles complete in increasing j (every rank visits an owner's tiles in the same order, its own last),
    // so these warps trail the GEMM by about one tile instead of starting after it
    const int etid = (warp - 6) * 32 + lane;                  // 0..127
    const int n_owned = (n_tiles - rank + world - 1) / world;
    uint32_t* my_flags = comm.flags[rank];
    const int done_off = max_slots * kOpMaxWorld;
    for (int j = blockIdx.x; j < n_owned; j += gridDim.x) {
      const int tile = rank + world * j;
      const int m0 = (tile / n_nt) * kOpBM, n0 = (tile % n_nt) * kOpBN;
      if (etid < world) wait_flag_ge(my_flags + j * kOpMaxWorld + etid, epoch, timeout_ns);
      named_bar_sync(2, 128);
      const uint8_t* part0 = comm.staging[rank] + (int64_t)j * kOpTileBytes;
      const int64_t src_stride = (int64_t)max_slots * kOpTileBytes;
      const int64_t out_base = 0;
      if (nvls)
        reduce_owned_tile_nvls<T>(comm, tile, etid, m0, n0, M, N);
      else if (world <= 2)
        reduce_owned_tile<T, 2>(comm, part0, src_stride, world, etid, m0, n0, M, N, out_base);
      else if (world <= 4)
        reduce_owned_tile<T, 4>(comm, part0, src_stride, world, etid, m0, n0, M, N, out_base);
      else
        reduce_owned_tile<T, 8>(comm, part0, src_stride, world, etid, m0, n0, M, N, out_base);
      named_bar_sync(2, 128);      // CTA-scope order of everyone's stores before the cumulative sys-scope release
      if (etid < world) st_release_sys(comm.flags[etid] + done_off + tile, epoch);
    }
    if (etid == 0) trace[2] = globaltimer_ns();

  }

  tc::fence_before_sync();
  __syncthreads();
  if constexpr (PAIR) tc::cluster_sync();   // the peer no longer signals this CTA's barriers nor reads its shared memory
  if (warp == 1) {
    tc::fence_after_sync();
    if constexpr (PAIR)
      tc::tmem_dealloc2(tmem, 512);
    else
      tc::tmem_dealloc(tmem, 256);
  }
  // the last CTA to get here closes the call: every CTA has read `epoch` by now, none will read it again
  if (threadIdx.x == 0) {
    trace[3] = globaltimer_ns();
    __threadfence();
    if (atomicAdd(ctl + 1, 1u) == gridDim.x - 1) {
      ctl[1] = 0u;
      *reinterpret_cast<volatile uint32_t*>(ctl) = epoch;
      __threadfence();
    }
  }
}

}  // namespace bsgl

// ------------------------------------------------------------------------------------ host side
namespace bsgl {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      p = nullptr;
    return reinterpret_cast<EncodeTiledFn>(p);
  }();
  return fn;
}

// [rows, K] row-major 16-bit matrix, rows `ld` elements apart -> 2-D map with a {64 x 128} box, 128B swizzle.
// Encoding is a driver call (~1-2 us); the weight's map never changes and the activation buffer of a layer is
// usually the same allocation step after step, so maps are cached by (pointer, shape, pitch, dtype).
struct MapKey {
  const void* ptr;
  int64_t rows, K, ld;
  int dtype;
  bool operator==(const MapKey& o) const {
    return ptr == o.ptr && rows == o.rows && K == o.K && ld == o.ld && dtype == o.dtype;
  }
};
struct MapKeyHash {
  size_t operator()(const MapKey& k) const {
    size_t h = std::hash<const void*>()(k.ptr);
    for (int64_t v : {k.rows, k.K, k.ld, (int64_t)k.dtype}) h = h * 1315423911u + std::hash<int64_t>()(v);
    return h;
  }
};

static int make_map(CUtensorMap* m, const void* ptr, int64_t rows, int64_t K, int64_t ld, int dtype) {
  static std::mutex mu;
  static std::unordered_map<MapKey, CUtensorMap, MapKeyHash> cache;
  const MapKey key{ptr, rows, K, ld, dtype};
  {
    std::lock_guard<std::mutex> lock(mu);
    auto it = cache.find(key);
    if (it != cache.end()) {
      *m = it->second;
      return BSGL_OK;
    }
  }
  EncodeTiledFn fn = encode_tiled_fn();
  BSGL_REQUIRE(fn != nullptr, "cuTensorMapEncodeTiled is not available from this driver");
  const cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)rows};
  const cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
  const cuuint32_t box[2] = {(cuuint32_t)kOpBK, (cuuint32_t)kOpBM};
  const cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(m, dtype == BSGL_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2,
                  const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  BSGL_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  std::lock_guard<std::mutex> lock(mu);
  if (cache.size() >= 1024) cache.clear();
  cache.emplace(key, *m);
  return BSGL_OK;
}

template <typename T, bool PAIR>
static int launch_oproj_as(const CUtensorMap& ta, const CUtensorMap& tb, const OprojComm& comm, int M, int N, int K,
                           int rank, int world, int max_slots, int max_tiles, uint64_t timeout_ns, int grid, int grouped,
                           cudaStream_t stream) {
  auto kern = oproj_allreduce_kernel<T, PAIR>;
  static thread_local int configured_dev = -1;
  int dev = 0;
  BSGL_CUDA_TRY(cudaGetDevice(&dev));
  if (configured_dev != dev) {
    BSGL_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kOpSmemBytes));
    configured_dev = dev;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)grid);
  cfg.blockDim = dim3(kOpThreads);
  cfg.dynamicSmemBytes = kOpSmemBytes;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = PAIR ? 2 : 1;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  BSGL_CUDA_TRY(cudaLaunchKernelEx(&cfg, kern, ta, tb, comm, M, N, K, rank, world, max_slots, max_tiles, timeout_ns,
                                   grouped));
  return check_launch("oproj_allreduce_kernel");
}

template <typename T>
static int launch_oproj(const CUtensorMap& ta, const CUtensorMap& tb, const OprojComm& comm, int M, int N, int K,
                        int rank, int world, int max_slots, int max_tiles, uint64_t timeout_ns, cudaStream_t stream) {
  int dev = 0, sms = 0;
  BSGL_CUDA_TRY(cudaGetDevice(&dev));
  BSGL_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  if (sms > 256) sms = 256;   // the flag block reserves 256 per-CTA timestamp records (bsgl_oproj_flag_bytes)
  const int n_mt = (M + kOpBM - 1) / kOpBM, n_nt = N / kOpBN, n_tiles = n_mt * n_nt;
  // Visiting order of the peer-store path.  "global" (default): every rank walks the tiles in the same order, so
  // the `world` partials of a tile reach its owner at about the same time and the owner's reduction trails the
  // GEMM tile by tile (the 148 tiles in flight cover all owners evenly: no hot link).  "grouped": owner by owner
  // starting at rank + 1 (round 1's order; an owner's tiles then all complete in the last 1/world of the GEMM).
  static const int grouped = [] {
    const char* e = getenv("BSGL_K8_TILE_ORDER");
    return (e && e[0] == 'g' && e[1] == 'r') ? 1 : 0;
  }();
  // CTA pairs (cta_group::2, 256 x 256 blocks): measured 67.5 vs 82.0 us at world 1 on 1024 x 5120 x 5120 (795 vs
  // 655 TFLOP/s), but 69.4 vs 62.0 us at TP2 on the same output: 80 blocks on 74 pairs finish in two coarse rounds, so
  // all partial tiles are published in two bursts, where the 320 single-CTA tiles spread their peer stores under the
  // GEMM.  Chosen automatically only without an exchange (world 1) and when the single-CTA tiles would need more
  // than one wave (below that they spread over twice as many SMs).  BSGL_K8_GEMM=single|pair overrides.
  const char* gm = getenv("BSGL_K8_GEMM");                    // read per call: tests and benches flip it in-process
  const int gemm_mode = !gm ? 0 : (gm[0] == 's' ? 1 : (gm[0] == 'p' ? 2 : 0));
  const bool pair_ok = n_mt >= 2 && n_nt % 2 == 0 && sms >= 2;
  const bool pair = pair_ok && (gemm_mode == 2 || (gemm_mode == 0 && world == 1 && n_tiles > sms));
  // every CTA must be resident: phase 2 waits on tiles other CTAs of this grid produce (one CTA per SM: 197 KB of
  // shared memory each; a persistent kernel of another stream holding SMs would stall it until the timeout)
  if (pair) {
    // how many CTA pairs the device can hold at once (a pair needs both SMs of a TPC), asked once per device
    static thread_local int pairs_dev = -1, pairs_max = 0;
    if (pairs_dev != dev) {
      auto kern = oproj_allreduce_kernel<T, true>;
      BSGL_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kOpSmemBytes));
      cudaLaunchConfig_t cfg = {};
      cfg.gridDim = dim3((unsigned)(2 * (sms / 2)));
      cfg.blockDim = dim3(kOpThreads);
      cfg.dynamicSmemBytes = kOpSmemBytes;
      cudaLaunchAttribute attr[1];
      attr[0].id = cudaLaunchAttributeClusterDimension;
      attr[0].val.clusterDim.x = 2;
      attr[0].val.clusterDim.y = 1;
      attr[0].val.clusterDim.z = 1;
      cfg.attrs = attr;
      cfg.numAttrs = 1;
      int n = 0;
      BSGL_CUDA_TRY(cudaOccupancyMaxActiveClusters(&n, kern, &cfg));
      pairs_max = n;
      pairs_dev = dev;
    }
    const int n_q = ((n_mt + 1) / 2) * (n_nt / 2);
    const int cap = pairs_max < sms / 2 ? pairs_max : sms / 2;
    BSGL_REQUIRE(cap >= 1, "bsgl_oproj_allreduce: the device cannot hold one CTA pair of this kernel");
    const int pairs = n_q < cap ? n_q : cap;
    return launch_oproj_as<T, true>(ta, tb, comm, M, N, K, rank, world, max_slots, max_tiles, timeout_ns, 2 * pairs,
                                    grouped, stream);
  }
  const int grid = n_tiles < sms ? n_tiles : sms;
  return launch_oproj_as<T, false>(ta, tb, comm, M, N, K, rank, world, max_slots, max_tiles, timeout_ns, grid, grouped,
                                   stream);
}

}  // namespace bsgl

extern "C" int bsgl_comm_alloc(int64_t bytes, void** dev_ptr, void* ipc_handle_64) {
  using namespace bsgl;
  BSGL_REQUIRE(bytes > 0 && dev_ptr && ipc_handle_64, "bsgl_comm_alloc: bad arguments");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  void* p = nullptr;
  BSGL_CUDA_TRY(cudaMalloc(&p, (size_t)bytes));
  BSGL_CUDA_TRY(cudaMemset(p, 0, (size_t)bytes));
  BSGL_CUDA_TRY(cudaDeviceSynchronize());
  cudaIpcMemHandle_t h;
  BSGL_CUDA_TRY(cudaIpcGetMemHandle(&h, p));
  memcpy(ipc_handle_64, &h, 64);
  *dev_ptr = p;
  return BSGL_OK;
}

extern "C" int bsgl_comm_open(const void* ipc_handle_64, void** dev_ptr) {
  using namespace bsgl;
  BSGL_REQUIRE(ipc_handle_64 && dev_ptr, "bsgl_comm_open: null pointer");
  cudaIpcMemHandle_t h;
  memcpy(&h, ipc_handle_64, 64);
  BSGL_CUDA_TRY(cudaIpcOpenMemHandle(dev_ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return BSGL_OK;
}

extern "C" int bsgl_comm_close(void* dev_ptr) {
  using namespace bsgl;
  if (dev_ptr) BSGL_CUDA_TRY(cudaIpcCloseMemHandle(dev_ptr));
  return BSGL_OK;
}

extern "C" int bsgl_comm_free(void* dev_ptr) {
  using namespace bsgl;
  if (dev_ptr) BSGL_CUDA_TRY(cudaFree(dev_ptr));
  return BSGL_OK;
}

extern "C" int bsgl_oproj_flag_bytes(int max_slots, int max_tiles, int64_t* bytes) {
  using namespace bsgl;
  BSGL_REQUIRE(max_slots > 0 && max_tiles > 0 && bytes != nullptr, "bsgl_oproj_flag_bytes: bad arguments");
  // must match the kernel's `trace` address: flag words rounded up to 1024 bytes, then 256 CTA records of 32 bytes
  const int64_t words = (int64_t)max_slots * kOpMaxWorld + max_tiles + 2;
  *bytes = ((words * 4 + 1023) & ~(int64_t)1023) + 256 * 32;
  return BSGL_OK;
}

extern "C" int bsgl_oproj_allreduce(const void* a, int64_t lda, const void* w, int64_t ldw, int n_tokens, int hidden,
                                    int k_local, int rank, int world, void* const* staging_ptrs,
                                    void* const* out_ptrs, void* const* flag_ptrs, void* mc_staging, void* mc_out,
                                    int max_slots, int max_tiles, int64_t out_bytes, double timeout_s, int dtype,
                                    void* stream) {
  using namespace bsgl;
  BSGL_REQUIRE(n_tokens >= 0, "bsgl_oproj_allreduce: negative token count");
  BSGL_REQUIRE(a && w && staging_ptrs && out_ptrs && flag_ptrs, "bsgl_oproj_allreduce: null pointer");
  BSGL_REQUIRE(world >= 1 && world <= kOpMaxWorld && rank >= 0 && rank < world,
               "bsgl_oproj_allreduce: rank %d / world %d out of range (max %d)", rank, world, kOpMaxWorld);
  BSGL_REQUIRE(hidden > 0 && hidden % kOpBN == 0, "bsgl_oproj_allreduce: hidden %d is not a multiple of %d", hidden,
               kOpBN);
  BSGL_REQUIRE(k_local > 0 && k_local % kOpBK == 0, "bsgl_oproj_allreduce: K %d is not a multiple of %d", k_local,
               kOpBK);
  BSGL_REQUIRE(dtype == BSGL_BF16 || dtype == BSGL_F16, "bsgl_oproj_allreduce: dtype must be bf16 or fp16");
  BSGL_REQUIRE(lda >= k_local && ldw >= k_local && lda % 8 == 0 && ldw % 8 == 0 && (uintptr_t)a % 16 == 0 &&
                   (uintptr_t)w % 16 == 0,
               "bsgl_oproj_allreduce: operands must be 16-byte aligned with row strides that are multiples of 8");
  BSGL_REQUIRE((mc_staging == nullptr) == (mc_out == nullptr),
               "bsgl_oproj_allreduce: the multicast views of staging and output go together");
  BSGL_REQUIRE(timeout_s > 0.0, "bsgl_oproj_allreduce: timeout must be positive");
  if (n_tokens == 0) return BSGL_OK;
  const int n_tiles = ((n_tokens + kOpBM - 1) / kOpBM) * (hidden / kOpBN);
  BSGL_REQUIRE(n_tiles <= max_tiles && (n_tiles + world - 1) / world <= max_slots,
               "bsgl_oproj_allreduce: %d tiles exceed the communicator (max_tiles %d, max_slots %d)", n_tiles,
               max_tiles, max_slots);
  BSGL_REQUIRE((int64_t)n_tokens * hidden * 2 <= out_bytes, "bsgl_oproj_allreduce: output buffer too small");
  OprojComm comm;
  for (int i = 0; i < kOpMaxWorld; ++i) {
    const int j = i < world ? i : 0;
    BSGL_REQUIRE(staging_ptrs[j] && out_ptrs[j] && flag_ptrs[j], "bsgl_oproj_allreduce: null peer pointer");
    comm.staging[i] = (uint8_t*)staging_ptrs[j];
    comm.out[i] = (uint8_t*)out_ptrs[j];
    comm.flags[i] = (uint32_t*)flag_ptrs[j];
  }
  comm.mc_staging = (uint8_t*)mc_staging;
  comm.mc_out = (uint8_t*)mc_out;
  CUtensorMap ta, tb;
  int rc = make_map(&ta, a, n_tokens, k_local, lda, dtype);
  if (rc != BSGL_OK) return rc;
  rc = make_map(&tb, w, hidden, k_local, ldw, dtype);
  if (rc != BSGL_OK) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  const uint64_t timeout_ns = (uint64_t)(timeout_s * 1e9);
  if (dtype == BSGL_BF16)
    return launch_oproj<__nv_bfloat16>(ta, tb, comm, n_tokens, hidden, k_local, rank, world, max_slots, max_tiles,
                                       timeout_ns, s);
  return launch_oproj<__half>(ta, tb, comm, n_tokens, hidden, k_local, rank, world, max_slots, max_tiles, timeout_ns,
                              s);
}
