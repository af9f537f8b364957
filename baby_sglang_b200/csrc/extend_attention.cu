// K3  ragged extend attention on tcgen05 tensor cores with TMEM accumulators (sm_100a).
//
// Replaces the reference's per-request prefill loop (attn_backend.py:34-110): page-table slice
// (:54-57), K/V gather of prefix + new tokens (:58-59), GQA repeat_interleave (:61-66), dense
// additive mask `col > prefix + row` (:74-92), SDPA (:94-101), cat (:110).
//
// PERSISTENT: one CTA per SM walks the work items (dispatch order, dealt boustrophedon over the CTAs); TMEM, the
// mbarriers (phases carried across items) and the K/V ring live for the whole launch, and the producers gather the
// next item's first K/V tiles and its Q tiles while the current item drains -- the ~9 us of launch + allocation +
// pipeline fill a one-item CTA paid are gone (prefix 1024: +8 %, no prefix: +17 %; long prefixes, which run at the
// board's power cap, +-1 %).
// Work item = (request, KV head, q chunk).  The G q heads that share a KV head are packed into the
// M dimension: packed row m = token * G + head, 256 packed rows per chunk = two 128-row tiles
// (A and B) that share one stream of K/V tiles, so every gathered KV row feeds 2 x 128 MMA rows.
//
//   producers (3 warps)  gather Q once, then K_j / V_j tiles of 128 KV tokens through the page
//                        table with 16-byte cp.async into 128B-swizzled UMMA operand layout
//                        (invalid tokens are zero-filled); K and V tiles alternate through one ring
//                        of 5 x 32 KB slots; completion is published per slot with
//                        cp.async.mbarrier.arrive, the MMA thread orders the generic and async
//                        proxies (fence.proxy.async) after waiting and before issuing.
//   MMA warp (1 thread)  S_X = Q_X . K_j^T   (tcgen05.mma SS, M128 N128 K16 x D/16, fp32 in TMEM)
//                        O_X += P_X . V_j    (tcgen05.mma TS: P read from TMEM, V MN-major from smem)
//                        issue order S_A S_B | PV_A S_A' PV_B S_B' | ... so one tile's softmax
//                        overlaps the other tile's MMAs; tcgen05.commit signals the mbarriers.
//   softmax (2 x 4 warps) one thread per row: tcgen05.ld the S row, mask `col <= prefix + token`,
//                        online softmax in base 2 with lazy rescaling of O (only when the running
//                        max grows by > 2^8).  The first half's exponentials are evaluated under the
//                        running max while the tile max is still being reduced (redone only when the
//                        max grew past the threshold), so the max pass is off the critical path; math is
//                        packed fp32x2 (FFMA2/FADD2); one pair in four of the second half goes through a
//                        polynomial on the FMA pipe instead of the SFU.  P is written back over S as
//                        packed 16-bit pairs (tcgen05.st) and handed over in two halves; final O / l -> out.
//                        Padding rows (beyond the chunk's tokens) are never masked and never stored.
//
// TMEM columns: S_A [0,128) S_B [128,256) O_A [256,256+D) O_B [256+D,256+2D); P_X aliases S_X.
// Tensor-bound: flops = 4 * Hq * D * sum_i ext_i * (prefix_i + (ext_i + 1) / 2).
#include <cstdlib>

#include "tcgen05.cuh"

namespace bsgl {

// Development aid (scripts/ext_trace.cu): per-role clock64 timestamps of one CTA.  Compiled out of the library.
#ifdef BSGL_EXT_TRACE
__device__ long long* g_ext_trace = nullptr;   // [role 10][step 128][event 8]; roles: softmax warps 0-7, MMA 8, producer 9
__device__ int g_ext_trace_cta = 0;
#define BSGL_TR_INIT() long long* const tr_base = (blockIdx.x == g_ext_trace_cta) ? g_ext_trace : nullptr
#define BSGL_TR(role, step, ev)                                                              \
  do {                                                                                       \
    if (tr_base && (step) < 128) tr_base[((role) * 128 + (step)) * 8 + (ev)] = clock64();    \
  } while (0)
#else
#define BSGL_TR_INIT() do {} while (0)
#define BSGL_TR(role, step, ev) do {} while (0)
#endif

constexpr int kExtSoftmaxWarps = 8;
constexpr int kExtMmaWarp = 8;
constexpr int kExtProducerWarps = 3;
constexpr int kExtProducerThreads = kExtProducerWarps * 32;
constexpr int kExtThreads = (kExtSoftmaxWarps + 1 + kExtProducerWarps) * 32;   // 384
constexpr int kExtRows = 256;        // packed q rows per chunk (2 MMA tiles)
constexpr int kExtKvTile = 128;      // KV tokens per tile
constexpr float kRescaleThreshold = 8.f;   // log2 units
constexpr int kExtPParts = 2;        // P is handed to the MMA warp in this many column groups per tile

template <int D>
struct ExtCfg {
  static constexpr int NH = D / 64;                      // 128-byte column blocks per row
  static constexpr int HALF_BYTES = 128 * 128;           // [128 rows x 128 B]
  static constexpr int TILE_BYTES = NH * HALF_BYTES;     // one 128-row operand tile
  // K and V tiles alternate through ONE ring (K_0 V_0 K_1 V_1 ...): tile u = 2j (K_j) or 2j+1 (V_j)
  // lives in slot u % RING.  An odd ring size keeps 2.5 KV steps in flight next to the 2 q tiles.
  static constexpr int RING = D == 128 ? 5 : 8;
  static constexpr int Q_OFF = 0;
  static constexpr int RING_OFF = 2 * TILE_BYTES;
  static constexpr int BAR_OFF = RING_OFF + RING * TILE_BYTES;
  // q_full, q_free, full/empty per slot, s_full[2], p_full[parts][2], o_full[2]
  static constexpr int N_BARS = 2 + 2 * RING + 2 + 2 * kExtPParts + 2;
  static constexpr int MISC_OFF = BAR_OFF + N_BARS * 8;
  static constexpr int SMEM_BYTES = MISC_OFF + 64;
  static_assert(SMEM_BYTES <= 227 * 1024, "extend kernel shared memory over budget");
  static constexpr int TMEM_COLS = 512;
  static constexpr int O_COL = 256;
  static constexpr int CPR = D / 8;                      // 16-byte chunks per row
  static constexpr int RSTEP = kExtProducerThreads / CPR;
  static constexpr int NR = (128 + RSTEP - 1) / RSTEP;   // rows per producer thread per 128-row tile
};

template <typename T, int D, int EXPMODE>
__global__ void __launch_bounds__(kExtThreads, 1)
extend_attention_kernel(const T* __restrict__ q, const T* __restrict__ k_slab, const T* __restrict__ v_slab,
                        const int32_t* __restrict__ req_to_token, int64_t pt_stride,
                        const int64_t* __restrict__ req_pool_indices, const int64_t* __restrict__ prefix_lens,
                        const int32_t* __restrict__ qo_indptr, const int32_t* __restrict__ req_order, int batch,
                        int n_q_heads, int n_kv_heads, float scale_log2, T* __restrict__ out, int total_vb,
                        int32_t* __restrict__ work_counter) {
  using C = ExtCfg<D>;
  extern __shared__ __align__(1024) uint8_t smem_raw[];   // SWIZZLE_128B operand tiles need 1024-byte alignment
  const uint32_t smem_base = smem_u32(smem_raw);
  if (smem_base & 1023u) __trap();
  uint8_t* smem = smem_raw;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  BSGL_TR_INIT();
  const int G = n_q_heads / n_kv_heads;
  const int TQ = kExtRows / G;                // tokens per chunk

  // ---- PERSISTENT: one CTA per SM walks the work items vb = blockIdx.x, blockIdx.x + gridDim.x, ...  (vb % Hkv =
  // KV head, fastest, so the heads of one chunk run together; vb / Hkv = position in dispatch order).  Every warp
  // resolves an item by itself (warp scan over the chunk counts ceil(ext_i / TQ) in dispatch order -- req_order:
  // longest KV first, and a request's longest-KV chunk first), so the roles stay decoupled: barrier phases, the
  // K/V ring and TMEM carry across items, the producers prefetch the next item's tiles while this one drains,
  // and nothing is re-allocated or re-initialised per item.
  struct Work {
    int found, req, h, q_begin, prefix, t0, n_tok, kv_end, nkv;
  };
  // Items come in decreasing cost (dispatch order).  With a work counter the CTAs draw them greedily (the next item
  // goes to whichever CTA is free first, like the hardware's block scheduler did for one-item CTAs): producer warp 0
  // is the dealer -- it runs ahead of the other roles anyway -- and publishes every draw in a small shared-memory
  // FIFO that the other warps read; no role ever waits for a slower one to read it (the dealer is at most the depth of
  // the K/V ring ahead, the FIFO holds 8 draws).  Without a counter the rounds are dealt boustrophedon: CTA b takes
  // position b of the even rounds and position G - 1 - b of the odd ones (plain round-robin gave CTA 0 the dearest
  // item of every round: 7 % slower on the ragged C3 mix).
  volatile int* deal_fifo = reinterpret_cast<volatile int*>(smem + C::MISC_OFF + 16);   // [8]
  volatile int* deal_pub = reinterpret_cast<volatile int*>(smem + C::MISC_OFF + 48);    // draws published so far
  const bool is_dealer = warp == kExtMmaWarp + 1;
  auto vb_of = [&](int round) -> int {
    if (work_counter == nullptr) {
      const int g = (int)gridDim.x, b = (int)blockIdx.x;
      return round * g + ((round & 1) ? g - 1 - b : b);
    }
    int vb = 0;
    if (is_dealer) {
      if (lane == 0) {
        vb = atomicAdd(work_counter, 1);
        deal_fifo[round & 7] = vb;
        __threadfence_block();
        *deal_pub = round + 1;
      }
      vb = __shfl_sync(0xffffffffu, vb, 0);
    } else {
      if (lane == 0) {
        while (*deal_pub <= round) __nanosleep(64);
        __threadfence_block();
        vb = deal_fifo[round & 7];
      }
      vb = __shfl_sync(0xffffffffu, vb, 0);
    }
    return vb;
  };
  auto find_work = [&](int vb) {
    Work w;
    w.found = 0;
    w.h = vb % n_kv_heads;
    const int target = vb / n_kv_heads;
    int base = 0, req = 0, chunk = 0;
    for (int i0 = 0; i0 < batch && !w.found; i0 += 32) {
      const int pos = i0 + lane;
      const int i = pos < batch ? (req_order ? req_order[pos] : pos) : 0;
      const int cnt = pos < batch ? (qo_indptr[i + 1] - qo_indptr[i] + TQ - 1) / TQ : 0;
      int incl = cnt;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int n = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += n;
      }
      const int lo = base + incl - cnt;
      const bool hit = target >= lo && target < lo + cnt;
      const uint32_t ballot = __ballot_sync(0xffffffffu, hit);
      if (ballot) {
        const int src = __ffs(ballot) - 1;
        req = __shfl_sync(0xffffffffu, i, src);
        chunk = __shfl_sync(0xffffffffu, cnt - 1 - (target - lo), src);   // a request's longest-KV chunk goes first
        w.found = 1;
      }
      base += __shfl_sync(0xffffffffu, incl, 31);
    }
    if (w.found) {
      w.req = req;
      w.q_begin = qo_indptr[req];
      const int ext = qo_indptr[req + 1] - w.q_begin;
      w.prefix = (int)prefix_lens[req];
      w.t0 = chunk * TQ;                                   // first token of the chunk inside the extend
      w.n_tok = min(TQ, ext - w.t0);
      w.kv_end = w.prefix + w.t0 + w.n_tok;                // KV tokens the last row of the chunk sees
      w.nkv = (w.kv_end + kExtKvTile - 1) / kExtKvTile;
    }
    return w;
  };

  // ---- barriers, TMEM
  const uint32_t bar0 = smem_base + C::BAR_OFF;
  constexpr int RING = C::RING;
  const uint32_t q_full = bar0;
  const uint32_t full0 = bar0 + 8, empty0 = full0 + 8 * RING, s_full = empty0 + 8 * RING, p_full = s_full + 16;
  const uint32_t q_free = p_full + 16 * kExtPParts;     // the item's last S MMA has read the Q tiles
  // O_X holds the item's last PV.  Its own barrier (one completion per item): on s_full it would be a second
  // completion between two waits of the softmax warps, which a parity wait cannot tell from none.
  const uint32_t o_full = q_free + 8;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + C::MISC_OFF);
  if (threadIdx.x == 0) {
    *reinterpret_cast<volatile int*>(smem + C::MISC_OFF + 48) = 0;     // no work item dealt yet
    mbar_init(q_full, kExtProducerThreads);       // one cp.async-completion arrive per producer thread
    mbar_init(q_free, 1);                         // tcgen05.commit
    mbar_init(o_full, 1);                         // tcgen05.commit
    mbar_init(o_full + 8, 1);
    for (int s = 0; s < RING; ++s) {
      mbar_init(full0 + 8 * s, kExtProducerThreads);
      mbar_init(empty0 + 8 * s, 1);               // tcgen05.commit
    }
    for (int x = 0; x < 2; ++x) {
      mbar_init(s_full + 8 * x, 1);
      for (int c = 0; c < kExtPParts; ++c) mbar_init(p_full + 16 * c + 8 * x, 128);   // P column group c of tile x
    }
    fence_mbar_init();
  }
  if (warp == kExtMmaWarp) {
    tc::tmem_alloc(smem_u32(tmem_slot), C::TMEM_COLS);
    tc::tmem_relinquish();
  }
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = *tmem_slot;

  const uint32_t q_smem = smem_base + C::Q_OFF, ring_smem = smem_base + C::RING_OFF;

  if (warp < kExtSoftmaxWarps) {
    // ============================================================ softmax: one thread per row
    tc::setmaxnreg_inc<208>();   // 2 x 128 x 208 + 128 x 88 = 384 x 168 registers
    const int x = warp >> 2;                               // tile A / B
    const int row = (warp & 3) * 32 + lane;                // TMEM lane
    const int m = x * 128 + row;                           // packed row in the chunk
    const int tok = m / G;
    const int head = m - tok * G;
    const uint32_t lane_base = (uint32_t)((warp & 3) * 32) << 16;
    const uint32_t s_addr = tmem + lane_base + x * 128;
    const uint32_t o_addr = tmem + lane_base + C::O_COL + x * D;
    uint32_t sf_par = 0;                                   // parity of the s_full[x] completions consumed by earlier items
   for (int round = 0;; ++round) {
    const int vb = vb_of(round);
    if (vb >= total_vb) break;
    const Work W = find_work(vb);
    if (!W.found) break;
    const int nkv = W.nkv;
    // last visible KV column.  Padding rows (zero-filled Q, never stored) see everything: a row that
    // masked every tile would drag its whole warp through the mask code on every tile.
    const int limit = (tok < W.n_tok && m < TQ * G) ? W.prefix + W.t0 + tok : 0x3fffffff;
    float m_used = -INFINITY, l = 0.f;

    for (int j = 0; j < nkv; ++j) {
      mbar_wait_guarded(s_full + 8 * x, (sf_par + j) & 1);
      tc::fence_after_sync();
      if (lane == 0) BSGL_TR(warp, j, 0);
      uint32_t s[4][32];
#pragma unroll
      for (int c = 0; c < 4; ++c) tc::ld32(s_addr + c * 32, s[c]);
      tc::wait_ld();
      if (lane == 0) BSGL_TR(warp, j, 1);
      const int nvalid = limit - j * kExtKvTile + 1;       // visible columns of this tile
      if (nvalid < kExtKvTile) {
#pragma unroll
        for (int c = 0; c < 4; ++c)
#pragma unroll
          for (int i = 0; i < 32; ++i)
            if (c * 32 + i >= nvalid) s[c][i] = 0xff800000u;   // -inf
      }
      // Exponentials of P column group c under the reference max -neg_m (base 2), packed to 16-bit pairs.
      constexpr int NP = 64 / kExtPParts;                 // packed P columns per group
      auto exp_group = [&](int c, float neg_m, uint32_t (&p)[NP], uint64_t (&sums)[2], float (&mxs)[4], bool with_max) {
        const uint64_t scale2 = pack_f32x2(scale_log2, scale_log2), negm2 = pack_f32x2(neg_m, neg_m);
        // Software pipeline: the exponentials of batch b+1 are issued before batch b is summed and packed,
        // so 2 * BP SFU results are in flight and their latency is covered (the consumer of an ex2 placed
        // right behind it would stall the warp for the SFU latency on every pair).
        constexpr int BP = 8, NB = NP / BP;
        float e0[2][BP], e1[2][BP];
        auto issue = [&](int bt, int buf) {
#pragma unroll
          for (int k = 0; k < BP; ++k) {
            const int i = bt * BP + k;
            const int e = c * 2 * NP + 2 * i;          // adjacent columns = adjacent registers = one FFMA2
            const uint64_t x2 = ffma2(pack_f32x2(__uint_as_float(s[e >> 5][e & 31]), __uint_as_float(s[e >> 5][(e & 31) + 1])),
                                      scale2, negm2);
            // EXPMODE: share of the exponential PAIRS evaluated on the FMA pipe instead of the SFU
            //   0 none, 1 one in four, 2 two in four, 3 three in four; 4/5/6 = timing experiments only
            // (only in the later column groups: the first group runs with all 128 S values live and has no
            // registers to spare for the polynomial's temporaries)
            const bool poly = c > 0 && ((EXPMODE == 1 && (i & 3) == 3) || (EXPMODE == 2 && (i & 1) == 1) ||
                                        (EXPMODE == 3 && (i & 3) != 0));
            if (EXPMODE == 4 || EXPMODE == 6) {
              unpack_f32x2(x2, e0[buf][k], e1[buf][k]);
            } else if (poly) {
              poly_exp2_x2(x2, e0[buf][k], e1[buf][k]);
            } else {
              float x0, x1;
              unpack_f32x2(x2, x0, x1);
              e0[buf][k] = fast_exp2_ordered(x0);
              e1[buf][k] = fast_exp2_ordered(x1);
            }
          }
        };
        auto retire = [&](int bt, int buf) {
#pragma unroll
          for (int k = 0; k < BP; ++k) {
            const int i = bt * BP + k;
            sums[i & 1] = fadd2_ordered(sums[i & 1], pack_f32x2(e0[buf][k], e1[buf][k]));
            p[i] = pack2<T>(e0[buf][k], e1[buf][k]);
            if (with_max) {
              // the tile's row max rides along in the issue slots the SFU-bound exponentials leave free:
              // 128 / NP columns per pair
              constexpr int MPI = 128 / NP;
#pragma unroll
              for (int t = 0; t < MPI; ++t) {
                const int col = i * MPI + t;
                mxs[t & 3] = fmaxf(mxs[t & 3], __uint_as_float(s[col >> 5][col & 31]));
              }
            }
          }
        };
        issue(0, 0);
#pragma unroll
        for (int bt = 0; bt < NB; ++bt) {
          if (bt + 1 < NB) issue(bt + 1, (bt + 1) & 1);
          retire(bt, bt & 1);
        }
      };
      // Group 0 is evaluated speculatively under the running reference max while the tile max is still
      // being reduced; only if the max grew past the lazy-rescale threshold (first tile, rare later) is it
      // redone after the rescale.  Results are identical to max-then-exp; the max pass leaves the critical path.
      float neg_m = (m_used == -INFINITY) ? 0.f : -m_used;
      uint64_t sums[2] = {0ull, 0ull};                    // two packed fp32 pairs (+0.0f bit patterns)
      float mxs[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
      uint32_t p[NP];
      exp_group(0, neg_m, p, sums, mxs, true);
      const float mx = fmaxf(fmaxf(mxs[0], mxs[1]), fmaxf(mxs[2], mxs[3]));
      const float m_new = fmaxf(m_used, mx * scale_log2);
      const bool grow = (m_new - m_used) > kRescaleThreshold;   // false for NaN (-inf - -inf)
      if (lane == 0) BSGL_TR(warp, j, 2);
      if (__any_sync(0xffffffffu, grow)) {
        const float alpha = (m_new == -INFINITY) ? 1.f : fast_exp2(m_used - m_new);
        l *= alpha;
        m_used = m_new;
        if (j > 0) {   // PV_X(j-1) completed before S_X(j) was produced (in-order tensor pipe)
#pragma unroll
          for (int c = 0; c < D / 32; ++c) {
            uint32_t o[32];
            tc::ld32(o_addr + c * 32, o);
            tc::wait_ld();
#pragma unroll
            for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
            tc::st32(o_addr + c * 32, o);
          }
        }
        neg_m = (m_used == -INFINITY) ? 0.f : -m_used;
        sums[0] = sums[1] = 0ull;
        exp_group(0, neg_m, p, sums, mxs, false);
      }
#pragma unroll
      for (int c = 0; c < kExtPParts; ++c) {
        if (c > 0) exp_group(c, neg_m, p, sums, mxs, false);
        tc::st_cols(s_addr + c * NP, p);
        // hand each column group of P to the MMA warp as soon as it is stored: PV_X of the earlier
        // groups runs under the later groups' exponentials
        tc::wait_st();
        tc::fence_before_sync();
        mbar_arrive(p_full + 16 * c + 8 * x);
        if (lane == 0) BSGL_TR(warp, j, 3 + c);
      }
      {
        float a0, a1;
        unpack_f32x2(fadd2(sums[0], sums[1]), a0, a1);
        l += a0 + a1;
      }
    }
    // ---- epilogue: O / l -> out[token, head, :]
    mbar_wait_guarded(o_full + 8 * x, round & 1);          // one item per round
    sf_par ^= (uint32_t)nkv & 1u;
    tc::fence_after_sync();
    // the item is resolved once more here, so that its address fields are not live registers across the KV loop
    // (the softmax warpgroups run at their 208-register ceiling)
    int vb_again = vb;
    asm volatile("" : "+r"(vb_again));
    const Work E = find_work(vb_again);
    const bool valid = limit != 0x3fffffff;
    const float inv = (valid && l > 0.f) ? 1.f / l : 0.f;
    T* orow = out + ((int64_t)(E.q_begin + E.t0 + tok) * n_q_heads + (int64_t)E.h * G + head) * D;
#pragma unroll
    for (int c = 0; c < D / 32; ++c) {
      uint32_t o[32];
      tc::ld32(o_addr + c * 32, o);
      tc::wait_ld();
      if (valid) {
#pragma unroll
        for (int i = 0; i < 32; i += 8) {
          uint4 pk;
          pk.x = pack2<T>(__uint_as_float(o[i]) * inv, __uint_as_float(o[i + 1]) * inv);
          pk.y = pack2<T>(__uint_as_float(o[i + 2]) * inv, __uint_as_float(o[i + 3]) * inv);
          pk.z = pack2<T>(__uint_as_float(o[i + 4]) * inv, __uint_as_float(o[i + 5]) * inv);
          pk.w = pack2<T>(__uint_as_float(o[i + 6]) * inv, __uint_as_float(o[i + 7]) * inv);
          *reinterpret_cast<uint4*>(orow + c * 32 + i) = pk;
        }
      }
    }
    // O_X is read: the next item's first PV_X (which overwrites it) is issued only after this thread has published
    // that item's first P group, i.e. after this point in program order
    tc::fence_before_sync();
   }   // next work item
  } else if (warp == kExtMmaWarp) {
    // ============================================================ MMA issuer
    // The whole warp runs the loop with uniform control flow (descriptors, addresses and counters
    // stay in uniform registers, which is where UTCHMMA takes its operands from); one elected lane
    // issues the tcgen05 instructions.
    tc::setmaxnreg_dec<88>();
    {
      constexpr uint32_t idesc_s = tc::instr_desc<T>(128, kExtKvTile, 0, 0);
      constexpr uint32_t idesc_pv = tc::instr_desc<T>(128, D, 0, 1);
      const uint32_t tm = __shfl_sync(0xffffffffu, tmem, 0);
      auto wait_tile = [&](int u) {
        mbar_wait_guarded(full0 + 8 * (u % RING), (u / RING) & 1);
        fence_proxy_async_smem();
        tc::fence_after_sync();
      };
      auto issue_s = [&](int x, int u, bool release_k) {
        const uint32_t slot = u % RING;
        const uint32_t qa = q_smem + x * C::TILE_BYTES, kb = ring_smem + slot * C::TILE_BYTES;
        if (tc::elect_one()) {
#pragma unroll
          for (int kk = 0; kk < D / 16; ++kk) {
            const uint32_t off = (kk >> 2) * C::HALF_BYTES + (kk & 3) * 32;
            tc::mma_ss(tm + x * 128, tc::smem_desc_sw128(qa + off, 16, 1024),
                       tc::smem_desc_sw128(kb + off, 16, 1024), idesc_s, kk > 0 ? 1u : 0u);
          }
          tc::commit(s_full + 8 * x);
          if (release_k) tc::commit(empty0 + 8 * slot);
        }
        __syncwarp();
      };
      auto issue_pv = [&](int x, int u, int hlf, bool acc, bool release_v, bool last) {
        const uint32_t slot = u % RING;
        const uint32_t vb = ring_smem + slot * C::TILE_BYTES;
        if (tc::elect_one()) {
#pragma unroll
          for (int k4 = 0; k4 < kExtKvTile / 16 / kExtPParts; ++k4) {
            const int kk = hlf * (kExtKvTile / 16 / kExtPParts) + k4;
            tc::mma_ts(tm + C::O_COL + x * D, tm + x * 128 + kk * 8,
                       tc::smem_desc_sw128(vb + kk * 2048, C::HALF_BYTES, 1024), idesc_pv, (acc || kk > 0) ? 1u : 0u);
          }
          if (hlf == kExtPParts - 1 && release_v) tc::commit(empty0 + 8 * slot);
          if (hlf == kExtPParts - 1 && last) tc::commit(o_full + 8 * x);     // O_X complete
        }
        __syncwarp();
      };
      auto commit_q_free = [&]() {                     // the item's last S MMAs are issued: Q may be overwritten
        if (tc::elect_one()) tc::commit(q_free);
        __syncwarp();
      };
      uint32_t items_done = 0, u0 = 0, step0 = 0;      // items, ring tiles and softmax steps of earlier items
      for (int round = 0;; ++round) {
        const int vb = vb_of(round);
        if (vb >= total_vb) break;
        const Work W = find_work(vb);
        if (!W.found) break;
        const int nkv = W.nkv;
        mbar_wait_guarded(q_full, items_done & 1);
        wait_tile(u0);
        // S_X(0) overwrites the previous item's last P_X: the tensor pipe runs in issue order, PV_X(last) is ahead
        issue_s(0, u0, false);
        issue_s(1, u0, true);
        if (nkv == 1) commit_q_free();
        for (int j = 0; j < nkv; ++j) {
          const int uv = u0 + 2 * j + 1, uk = u0 + 2 * j + 2;     // V_j, K_{j+1}
          wait_tile(uv);
          if (lane == 0) BSGL_TR(8, j, 0);
#pragma unroll
          for (int x = 0; x < 2; ++x) {
#pragma unroll
            for (int hlf = 0; hlf < kExtPParts; ++hlf) {
              // (j = 0: the softmax threads of tile x publish this group only after their epilogue of the previous
              // item has read O_X, which this PV overwrites)
              mbar_wait_guarded(p_full + 16 * hlf + 8 * x, (step0 + j) & 1);
              tc::fence_after_sync();
              if (lane == 0) BSGL_TR(8, j, 1 + 3 * x + hlf);
              issue_pv(x, uv, hlf, j > 0, x == 1, j + 1 >= nkv);
            }
            if (j + 1 < nkv) {
              if (x == 0) wait_tile(uk);
              issue_s(x, uk, x == 1);
              if (x == 1 && j + 2 == nkv) commit_q_free();
              if (lane == 0) BSGL_TR(8, j, 3 + 3 * x);
            }
          }
        }
        u0 += 2 * nkv;
        step0 += nkv;
        ++items_done;
      }
    }
  } else {
    // ============================================================ producers: gather Q, K_j, V_j
    tc::setmaxnreg_dec<88>();
    const int p = threadIdx.x - (kExtMmaWarp + 1) * 32;
    const int c16 = p % C::CPR;                            // 16-byte chunk of the row
    const int r0 = p / C::CPR;
    const uint32_t col_off = (c16 >> 3) * C::HALF_BYTES;
    const int cc = c16 & 7;
    auto dst_of = [&](int r) { return col_off + (uint32_t)r * 128u + (uint32_t)((cc ^ (r & 7)) << 4); };

    const int64_t slab_row = (int64_t)n_kv_heads * D;
    uint32_t items_done = 0, u0 = 0;                       // items and ring tiles of earlier items
    for (int round = 0;; ++round) {
      const int vb = vb_of(round);
      if (vb >= total_vb) break;
      const Work W = find_work(vb);
      if (!W.found) break;
      const int nkv = W.nkv, kv_end = W.kv_end;
      const int32_t* idxp = req_to_token + req_pool_indices[W.req] * pt_stride;
      const T* kb = k_slab + (int64_t)W.h * D + c16 * 8;
      const T* vb_ = v_slab + (int64_t)W.h * D + c16 * 8;
      // K_j then V_j (same rows) into the next two ring slots
      auto load_pair = [&](int j) {
        // page-table entries of this tile (-1 = past the chunk's last visible token: zero-fill);
        // their latency hides behind the wait for a free slot
        int idx[C::NR];
#pragma unroll
        for (int i = 0; i < C::NR; ++i) {
          const int r = r0 + i * C::RSTEP;
          const int t = j * kExtKvTile + r;
          idx[i] = (r < kExtKvTile && t < kv_end) ? __ldg(idxp + t) : -1;
        }
#pragma unroll
        for (int kv = 0; kv < 2; ++kv) {
          const uint32_t u = u0 + 2 * j + kv;
          const uint32_t slot = u % RING;
          mbar_wait_guarded(empty0 + 8 * slot, ((u / RING) & 1) ^ 1);
          if (p == 0) BSGL_TR(9, j, 2 * kv);
          const T* base = kv == 0 ? kb : vb_;
          const uint32_t dst = ring_smem + slot * C::TILE_BYTES;
#pragma unroll
          for (int i = 0; i < C::NR; ++i) {
            const int r = r0 + i * C::RSTEP;
            if (r < kExtKvTile && !((EXPMODE == 5 || EXPMODE == 6) && u >= RING)) {   // 5/6: timing experiment, no gather
              const bool ok = idx[i] >= 0;
              cp_async16_zfill(dst + dst_of(r), base + (ok ? idx[i] : 0) * slab_row, ok ? 16u : 0u);
            }
          }
          cp_async_mbar_arrive(full0 + 8 * slot);    // arrives when this thread's copies have landed
          if (p == 0) BSGL_TR(9, j, 2 * kv + 1);
        }
      };
      // The item's first K/V tiles go into whatever ring slots the previous item has already released, BEFORE
      // the wait for its last S MMA to let go of the Q tiles: the pipeline fill of this item hides behind the
      // drain of the previous one.
      load_pair(0);
      if (items_done > 0) mbar_wait_guarded(q_free, (items_done - 1) & 1);
      // Q: packed row m = token * G + head; the G heads of a token are contiguous in q.
      for (int mrow = r0; mrow < kExtRows; mrow += C::RSTEP) {
        const int tok = mrow / G;
        const int head = mrow - tok * G;
        const bool ok = tok < W.n_tok && mrow < TQ * G;
        const T* src = q + ((int64_t)(W.q_begin + W.t0 + (ok ? tok : 0)) * n_q_heads + (int64_t)W.h * G + (ok ? head : 0)) * D +
                       c16 * 8;
        cp_async16_zfill(q_smem + (mrow >> 7) * C::TILE_BYTES + dst_of(mrow & 127), src, ok ? 16u : 0u);
      }
      cp_async_mbar_arrive(q_full);
      for (int j = 1; j < nkv; ++j) load_pair(j);
      u0 += 2 * nkv;
      ++items_done;
    }
    cp_async_wait_all();                             // nothing of this CTA is in flight at exit
  }

  // ---- teardown: every TMEM read is done before the allocation is released
  __syncthreads();
  if (work_counter != nullptr && threadIdx.x == 0) {
    // the last CTA to get here returns the counter pair to zero for the next launch (every CTA has stopped drawing)
    __threadfence();
    if (atomicAdd(work_counter + 1, 1) == (int)gridDim.x - 1) {
      work_counter[0] = 0;
      work_counter[1] = 0;
      __threadfence();
    }
  }
  if (warp == kExtMmaWarp) {
    tc::fence_after_sync();
    tc::tmem_dealloc(tmem, C::TMEM_COLS);
  }
}

}  // namespace bsgl

// ------------------------------------------------------------------------------------ launcher
namespace bsgl {

template <typename T, int D, int EXPMODE>
static int launch_extend(const void* q, const void* k_slab, const void* v_slab, const int32_t* req_to_token,
                         int64_t pt_stride, const int64_t* req_pool_indices, const int64_t* prefix_lens,
                         const int32_t* qo_indptr, const int32_t* req_order, int batch, int64_t n_tokens,
                         int n_q_heads, int n_kv_heads, float sm_scale, void* out, int32_t* work_counter,
                         cudaStream_t stream) {
  using C = ExtCfg<D>;
  auto kern = extend_attention_kernel<T, D, EXPMODE>;
  static thread_local int configured_dev = -1;
  int dev = 0;
  BSGL_CUDA_TRY(cudaGetDevice(&dev));
  if (configured_dev != dev) {
    BSGL_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM_BYTES));
    configured_dev = dev;
  }
  const int G = n_q_heads / n_kv_heads;
  const int TQ = kExtRows / G;
  // sum_i ceil(ext_i / TQ) <= floor(T / TQ) + batch work items per KV head; one persistent CTA per SM walks them
  // (items past the real count resolve to "not found" and end a CTA's walk)
  const int64_t max_chunks = n_tokens / TQ + batch;
  BSGL_REQUIRE(max_chunks * n_kv_heads < (1ll << 31), "bsgl_extend_attention: too many q chunks");
  const int total_vb = (int)(max_chunks * n_kv_heads);
  static thread_local int sms_dev = -1, sms = 0;
  if (sms_dev != dev) {
    BSGL_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    sms_dev = dev;
  }
  // BSGL_EXT_PERSIST=0: one CTA per work item (the loop then ends after its first round) -- experiments only
  const char* pe = getenv("BSGL_EXT_PERSIST");
  const bool persist = !(pe && pe[0] == '0');
  dim3 grid((unsigned)((total_vb < sms || !persist) ? total_vb : sms));
  const float scale_log2 = sm_scale * 1.4426950408889634f;
  kern<<<grid, kExtThreads, C::SMEM_BYTES, stream>>>((const T*)q, (const T*)k_slab, (const T*)v_slab, req_to_token,
                                                     pt_stride, req_pool_indices, prefix_lens, qo_indptr, req_order,
                                                     batch, n_q_heads, n_kv_heads, scale_log2, (T*)out, total_vb,
                                                     work_counter);
  return check_launch("extend_attention_kernel");
}

}  // namespace bsgl

extern "C" int bsgl_extend_attention(const void* q, const void* k_slab, const void* v_slab,
                                     const int32_t* req_to_token, int64_t page_table_stride,
                                     const int64_t* req_pool_indices, const int64_t* seq_lens,
                                     const int64_t* prefix_lens, const int32_t* qo_indptr,
                                     const int32_t* req_order, int batch, int64_t n_tokens, int n_q_heads, int n_kv_heads, int head_dim, float sm_scale,
                                     int dtype, void* out, int32_t* work_counter, void* stream) {
  using namespace bsgl;
  (void)seq_lens;   // seq = prefix + ext; ext comes from qo_indptr
  BSGL_REQUIRE(batch >= 0 && n_tokens >= 0, "bsgl_extend_attention: negative sizes");
  if (batch == 0 || n_tokens == 0) return BSGL_OK;
  BSGL_REQUIRE(q && k_slab && v_slab && req_to_token && req_pool_indices && prefix_lens && qo_indptr && out,
               "bsgl_extend_attention: null pointer");
  BSGL_REQUIRE(n_kv_heads > 0 && n_q_heads % n_kv_heads == 0, "bsgl_extend_attention: n_q_heads %% n_kv_heads != 0");
  const int G = n_q_heads / n_kv_heads;
  BSGL_REQUIRE(G <= 128, "bsgl_extend_attention: GQA group size %d > 128 is not supported", G);
  BSGL_REQUIRE(n_kv_heads <= 65535, "bsgl_extend_attention: too many KV heads");
  BSGL_REQUIRE(head_dim == 64 || head_dim == 128, "bsgl_extend_attention: head_dim %d not in {64,128}", head_dim);
  BSGL_REQUIRE(dtype == BSGL_BF16 || dtype == BSGL_F16, "bsgl_extend_attention: dtype must be bf16 or fp16");
  BSGL_REQUIRE(((uintptr_t)q % 16 == 0) && ((uintptr_t)k_slab % 16 == 0) && ((uintptr_t)v_slab % 16 == 0) &&
                   ((uintptr_t)out % 16 == 0),
               "bsgl_extend_attention: pointers must be 16-byte aligned");
  cudaStream_t s = (cudaStream_t)stream;
#define BSGL_EXT1(T, D, M)                                                                                          \
  return launch_extend<T, D, M>(q, k_slab, v_slab, req_to_token, page_table_stride, req_pool_indices, prefix_lens, \
                                qo_indptr, req_order, batch, n_tokens, n_q_heads, n_kv_heads, sm_scale, out, \
                                work_counter, s)
#ifdef BSGL_EXT_EXPERIMENTS
  // scripts/ builds only (-DBSGL_EXT_EXPERIMENTS): the other exponential splits and the timing experiments that
  // skip the exponentials (4, 6) or the K/V gather (5, 6) -- the latter give WRONG results by design, which is why
  // none of this is compiled into the shipped library.
  static const int expmode = [] {
    const char* e = getenv("BSGL_EXT_EXPMODE");
    return e ? atoi(e) : 1;
  }();
#define BSGL_EXT(T, D)                    \
  do {                                    \
    if (expmode == 0) BSGL_EXT1(T, D, 0); \
    if (expmode == 2) BSGL_EXT1(T, D, 2); \
    if (expmode == 3) BSGL_EXT1(T, D, 3); \
    if (expmode == 4) BSGL_EXT1(T, D, 4); \
    if (expmode == 5) BSGL_EXT1(T, D, 5); \
    if (expmode == 6) BSGL_EXT1(T, D, 6); \
    BSGL_EXT1(T, D, 1);                   \
  } while (0)
#else
  // EXPMODE 1 (one exponential pair in four on the FMA pipe) is the measured best and the only one shipped
#define BSGL_EXT(T, D) BSGL_EXT1(T, D, 1)
#endif
  if (dtype == BSGL_BF16) {
    if (head_dim == 128) BSGL_EXT(__nv_bfloat16, 128);
    BSGL_EXT(__nv_bfloat16, 64);
  }
  if (head_dim == 128) BSGL_EXT(__half, 128);
  BSGL_EXT(__half, 64);
#undef BSGL_EXT
#undef BSGL_EXT1
}
