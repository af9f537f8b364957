// K2 / K2b  paged GQA attention, one q row per sequence, split-KV persistent kernel (sm_100a).
//
// Replaces the reference's per-request decode loop (attn_backend.py:112-149): page-table slice
// (:122-125), K/V gather (:126-128), GQA repeat_interleave (:130-135), SDPA (:141-143), cat (:149).
//
// Shape of the work.  A "sequence" is (q row, KV-head group): HPB adjacent KV heads of one row and
// the HPB*G q heads reading them.  Its KV tokens are cut into 16-token tiles; the flattened tile
// space of all sequences is cut into n_ctas equal contiguous ranges (plan built by K4), so ragged
// and long contexts split across CTAs with at most two partial sequences per CTA.
//
// One CTA per SM, persistent:
//   * 4 producer warps: read the page-table entries (prefetched one tile ahead) and gather the
//     K/V rows with 16-byte cp.async.cg (LDGSTS): consecutive lanes take consecutive 16 B chunks of
//     one token row (HPB*D*2 contiguous bytes = every KV head of the group), so each warp
//     instruction is one 512 B burst.  Rows land in a ring of shared-memory stages with a +16 B
//     row pitch (conflict-free ldmatrix); completion is signalled on a per-stage mbarrier with
//     cp.async.mbarrier.arrive.  (A first version used one TMA bulk copy per row: measured
//     ~44 ns per copy per SM regardless of size, i.e. request-rate bound at 1 KiB rows.)
//   * 8 consumer warps: warp (h, sub) owns KV head h of the group and every NSUB-th tile.  Per
//     tile: S^T[16 tok x 8 heads] = K[16 x D] . Q^T (mma.sync m16n8k16, the G<=8 q heads of the
//     KV head sit in the N=8 dimension so nothing is padded), online softmax with warp shuffles,
//     P^T moved to the B-operand layout with movmatrix, O^T[D x 8] += V^T . P^T (V^T fragments come
//     from ldmatrix.trans).  fp32 accumulation throughout.
//   * per item: the NSUB partial states of each head are merged through shared memory and either
//     the final row (bf16/fp16) or a normalised fp32 partial + base-2 LSE is written.
//
// HBM-bound: algorithmic bytes per launch = sum(kv_len) * Hkv * D * 2 * e (+ page table + q/out).
#include "common.cuh"

namespace bsgl {

constexpr int kTile = BSGL_TILE_TOKENS;   // 16 KV tokens per tile
constexpr int kConsumerWarps = 8;
constexpr int kProducerWarps = 4;
constexpr int kProducerThreads = kProducerWarps * 32;
constexpr int kThreads = (kConsumerWarps + kProducerWarps) * 32;
constexpr int kMaxItems = 64;
constexpr int kSmemBudget = 227 * 1024;

struct Item {
  long long idx_off;   // element offset of the sequence's token list in idx_base
  int row;             // q / out row
  int hg;              // head group
  int tok_begin;       // first KV token of this item
  int tok_end;         // one past the last KV token
  int part_slot;       // -1: write the final row; else partial slot
  int valid;
};

template <int D, int HPB>
struct Cfg {
  static constexpr int NSUB = kConsumerWarps / HPB;
  static constexpr int ROW_BYTES = HPB * D * 2;
  static constexpr int ROW_PITCH = ROW_BYTES + 16;
  static constexpr int KV_BYTES = kTile * ROW_PITCH;     // K (or V) part of a stage
  static constexpr int STAGE_BYTES = 2 * KV_BYTES;
  static constexpr int CPR = ROW_BYTES / 16;               // 16-byte chunks per row
  static constexpr int RSTEP = kProducerThreads / CPR;      // rows covered by one pass of the producers
  static constexpr int NR = kTile / RSTEP;                  // rows per producer thread per tile
  static constexpr int PD_RAW = 16 / NR;                    // page-table prefetch distance in tiles
  static constexpr int PD = PD_RAW < 2 ? 2 : (PD_RAW > 8 ? 8 : PD_RAW);
  // Narrow rows (one KV head per group, 256 B): a tile is only 8 KB, i.e. ~360 clk of HBM time per SM, less than
  // the dependent-issue latency of one producer iteration (~84 instructions per warp per tile, measured with
  // ncu).  There each producer WARP takes every 4th tile on its own (WPT): 4 tiles are being issued at once and
  // the per-tile bookkeeping is paid by one warp instead of four.  (Tried for 512 B rows too: 4 % slower there.)
  static constexpr bool WPT = ROW_BYTES <= 256;
  static constexpr int RPI = 512 / ROW_BYTES;               // WPT: rows covered by one warp instruction
  static constexpr int NI = RPI > 0 ? kTile / RPI : 0;      // WPT: instructions per K (or V) tile
  static constexpr int FULL_ARRIVALS = WPT ? 32 : kProducerThreads;
  static_assert(CPR <= kProducerThreads && kProducerThreads % CPR == 0 && kTile % RSTEP == 0 && NR >= 1,
                "row width does not map onto the producer threads");
  static constexpr int DP = D + 4;                        // scratch row pitch in floats
  static constexpr int SCR_FLOATS = 8 * DP;               // per consumer warp
  static constexpr int SCR_BYTES = kConsumerWarps * SCR_FLOATS * 4;
  static constexpr int ML_BYTES = kConsumerWarps * 8 * 2 * 4;
  static constexpr int ITEM_BYTES = kMaxItems * (int)sizeof(Item);
  static constexpr int FIXED_BYTES = SCR_BYTES + ML_BYTES + ITEM_BYTES + 1024 /*barriers + slack*/;
  static constexpr int NSTAGES_RAW = (kSmemBudget - FIXED_BYTES) / STAGE_BYTES;
  static constexpr int NSTAGES_CAP = NSTAGES_RAW > 24 ? 24 : NSTAGES_RAW;
  // WPT: a stage must belong to the same producer warp in every round (NSTAGES % 4 == 0).  Otherwise the warp
  // owning round r+1 of a stage can run two rounds ahead of the warp owning round r, and its parity wait on the
  // empty barrier passes on the stale phase (seen as a hang at full load, never in the small parity cases).
  static constexpr int NSTAGES = WPT ? (NSTAGES_CAP / kProducerWarps) * kProducerWarps : NSTAGES_CAP;
  static constexpr int SMEM_BYTES = NSTAGES * STAGE_BYTES + FIXED_BYTES;
  static_assert(NSTAGES >= 3, "not enough shared memory for the KV ring");
};

template <typename T, int D, int HPB>
__global__ void __launch_bounds__(kThreads, 1)
paged_attention_kernel(const T* __restrict__ q, const T* __restrict__ k_slab, const T* __restrict__ v_slab,
                       const int32_t* __restrict__ idx_base, const int64_t* __restrict__ row_idx_off,
                       const int32_t* __restrict__ row_kv_len, const int32_t* __restrict__ plan_hdr,
                       const int32_t* __restrict__ tile_indptr, const int32_t* __restrict__ cta_first_seq,
                       int n_q_heads, int n_kv_heads, float scale_log2, T* __restrict__ out,
                       float* __restrict__ part_o, float* __restrict__ part_lse) {
  using C = Cfg<D, HPB>;
  constexpr int NST = C::NSTAGES;
  constexpr int NSUB = C::NSUB;
  constexpr int KS = D / 16;   // k-steps of QK^T == m-tiles of PV

  extern __shared__ __align__(128) uint8_t smem[];
  uint8_t* stages = smem;
  float* scratch = reinterpret_cast<float*>(smem + NST * C::STAGE_BYTES);
  float* ml = scratch + kConsumerWarps * C::SCR_FLOATS;                 // [warp][8 heads][m, l]
  Item* items = reinterpret_cast<Item*>(ml + kConsumerWarps * 16);
  uint64_t* bars = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(items) + C::ITEM_BYTES);
  const uint32_t full_bar0 = smem_u32(bars);
  const uint32_t empty_bar0 = smem_u32(bars + NST);
  const uint32_t stages_u32 = smem_u32(stages);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int G = n_q_heads / n_kv_heads;
  const int nhg = n_kv_heads / HPB;
  const int64_t slab_row = (int64_t)n_kv_heads * D;   // elements per slab row

  if (threadIdx.x == 0) {
    for (int s = 0; s < NST; ++s) {
      mbar_init(full_bar0 + 8 * s, C::FULL_ARRIVALS);   // one cp.async-completion arrive per issuing thread
      mbar_init(empty_bar0 + 8 * s, HPB);    // one arrive per consumer warp of the owning sub-group
    }
    fence_mbar_init();
    fence_proxy_async_smem();
  }

  // Programmatic dependent launch, both ways.  This kernel is launched as a dependent of whatever precedes it
  // on the stream (K1 kv_store triggers at its start): launch, CTA placement and the barrier set-up above
  // overlap that kernel; nothing it may have written (KV rows, q, the plan) is read before this wait.  The
  // merge kernel behind us is in turn our dependent: let it launch and park at its own wait.
  asm volatile("griddepcontrol.wait;" ::: "memory");
  asm volatile("griddepcontrol.launch_dependents;");
  const int cta = blockIdx.x;
  const int n_tiles = plan_hdr[0];
  const int n_eff = plan_hdr[1];
  const int n_seqs = plan_hdr[2];
  // read with the header, not after the early exit: one round trip less on the kernel's start-up chain
  // (cta_first_seq has gridDim.x + 1 entries whatever n_eff is)
  const int seq_first = cta_first_seq[cta];
  int seq_last = cta_first_seq[cta + 1];
  if (cta >= n_eff || n_tiles == 0) return;

  const int tile_lo = cta_tile_begin(cta, n_tiles, n_eff);
  const int tile_hi = cta_tile_begin(cta + 1, n_tiles, n_eff);
  if (seq_last > n_seqs - 1) seq_last = n_seqs - 1;

  uint32_t stage_ctr = 0;   // running tile counter of this CTA; identical in every warp

  for (int seq0 = seq_first; seq0 <= seq_last; seq0 += kMaxItems) {
    __syncthreads();   // previous pass fully consumed (and barrier init visible on the first pass)
    const int n_items = min(kMaxItems, seq_last - seq0 + 1);
    if (threadIdx.x < n_items) {
      const int j = seq0 + threadIdx.x;
      const int tb = tile_indptr[j], te = tile_indptr[j + 1];
      const int lo = max(tb, tile_lo), hi = min(te, tile_hi);
      Item it;
      it.valid = lo < hi;
      it.row = j / nhg;
      it.hg = j - it.row * nhg;
      const int kvl = row_kv_len[it.row];
      it.idx_off = row_idx_off[it.row];
      it.tok_begin = (lo - tb) * kTile;
      it.tok_end = min((hi - tb) * kTile, kvl);
      const bool whole = (lo == tb) && (hi == te);
      it.part_slot = whole ? -1 : 2 * cta + (lo == tb ? 1 : 0);
      items[threadIdx.x] = it;
    }
    __syncthreads();

    if (warp >= kConsumerWarps) {
      // ================================================================ producer warps
      // Two cursors walk the same (item, tile) sequence: `pf` loads page-table entries PD tiles
      // ahead into a register ring, `is` waits for a free stage and issues the row gathers.  The
      // distance hides the dependent idx -> row-address latency even when stages are short
      // (narrow rows), and it carries across item boundaries.
      const int p = threadIdx.x - kConsumerWarps * 32;
      const int col = p % C::CPR;             // this thread's 16-byte column of every row it copies
      // first row, then every RSTEP-th; WPT: the tile token whose page-table entry this lane fetches
      const int r0 = C::WPT ? (lane & 15) : p / C::CPR;
      const uint32_t dst0 = stages_u32 + r0 * C::ROW_PITCH + col * 16;
      constexpr int PD = C::PD;

      struct Cursor {
        const int32_t* idxp;
        int ii, ti, ntile, tok0, last_tok, hg;
      };
      auto load_item = [&](Cursor& c) {
        while (c.ii < n_items && !items[c.ii].valid) ++c.ii;
        if (c.ii < n_items) {
          const Item& it = items[c.ii];
          c.ti = 0;
          c.ntile = (it.tok_end - it.tok_begin + kTile - 1) / kTile;
          c.tok0 = it.tok_begin + r0;
          c.last_tok = it.tok_end - 1;      // never read the page table past the sequence
          c.idxp = idx_base + it.idx_off;
          c.hg = it.hg;
        }
      };
      auto advance = [&](Cursor& c) {
        ++c.ti;
        c.tok0 += kTile;
        if (c.ti == c.ntile) {
          ++c.ii;
          load_item(c);
        }
      };
      if constexpr (C::WPT) {
        const int pw = warp - kConsumerWarps;              // this warp owns tiles with (tile counter % 4) == pw
        const int rsub = lane / C::CPR;                    // row inside one warp instruction
        constexpr int PDW = 4;                             // own tiles of page-table prefetch (= 16 tiles ahead)
        auto advance_n = [&](Cursor& c, int n) {
          while (n > 0 && c.ii < n_items) {
            const int left = c.ntile - c.ti;
            if (n < left) {
              c.ti += n;
              c.tok0 += n * kTile;
              n = 0;
            } else {
              n -= left;
              ++c.ii;
              load_item(c);
            }
          }
        };
        Cursor pf;
        pf.ii = 0;
        load_item(pf);
        const int skip = (pw - (int)stage_ctr) & 3;
        advance_n(pf, skip);
        Cursor is = pf;
        uint32_t sc = stage_ctr + skip;
        int ring[PDW];
#pragma unroll
        for (int u = 0; u < PDW; ++u) {
          ring[u] = 0;
          if (pf.ii < n_items) {
            ring[u] = __ldg(pf.idxp + min(pf.tok0, pf.last_tok));
            advance_n(pf, kProducerWarps);
          }
        }
        const uint32_t dstw = stages_u32 + rsub * C::ROW_PITCH + col * 16;
        while (is.ii < n_items) {
#pragma unroll
          for (int u = 0; u < PDW; ++u) {
            if (is.ii >= n_items) break;
            const int idx_tile = ring[u];
            if (pf.ii < n_items) {
              ring[u] = __ldg(pf.idxp + min(pf.tok0, pf.last_tok));
              advance_n(pf, kProducerWarps);
            }
            const int64_t head_off = (int64_t)is.hg * HPB * D + col * 8;
            const T* kb = k_slab + head_off;
            const T* vb = v_slab + head_off;
            const uint32_t st = sc % NST;
            const uint32_t ph = (sc / NST) & 1;
            mbar_wait(empty_bar0 + 8 * st, ph ^ 1);
            const uint32_t dst = dstw + st * C::STAGE_BYTES;
#pragma unroll
            for (int i = 0; i < C::NI; ++i) {
              const int idx = __shfl_sync(0xffffffffu, idx_tile, i * C::RPI + rsub);
              const int64_t off = (int64_t)idx * slab_row;
              cp_async16(dst + i * C::RPI * C::ROW_PITCH, kb + off);
              cp_async16(dst + i * C::RPI * C::ROW_PITCH + C::KV_BYTES, vb + off);
            }
            cp_async_mbar_arrive(full_bar0 + 8 * st);
            sc += kProducerWarps;
            advance_n(is, kProducerWarps);
          }
        }
        // every warp leaves the pass with the same tile counter
        int nt = 0;
        for (int i = lane; i < n_items; i += 32)
          if (items[i].valid) nt += (items[i].tok_end - items[i].tok_begin + kTile - 1) / kTile;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) nt += __shfl_xor_sync(0xffffffffu, nt, o);
        stage_ctr += nt;
      } else {
        Cursor pf;
        pf.ii = 0;
        load_item(pf);
        Cursor is = pf;
        int ring[PD][C::NR];
#pragma unroll
      for (int u = 0; u < PD; ++u) {
        if (pf.ii < n_items) {
#pragma unroll
          for (int i = 0; i < C::NR; ++i) ring[u][i] = __ldg(pf.idxp + min(pf.tok0 + i * C::RSTEP, pf.last_tok));
          advance(pf);
        }
      }
      while (is.ii < n_items) {
#pragma unroll
        for (int u = 0; u < PD; ++u) {
          if (is.ii >= n_items) break;
          int idx_cur[C::NR];
#pragma unroll
          for (int i = 0; i < C::NR; ++i) idx_cur[i] = ring[u][i];
          if (pf.ii < n_items) {
#pragma unroll
            for (int i = 0; i < C::NR; ++i) ring[u][i] = __ldg(pf.idxp + min(pf.tok0 + i * C::RSTEP, pf.last_tok));
            advance(pf);
          }
          const int64_t head_off = (int64_t)is.hg * HPB * D + col * 8;
          const T* kb = k_slab + head_off;
          const T* vb = v_slab + head_off;
          const uint32_t st = stage_ctr % NST;
          const uint32_t ph = (stage_ctr / NST) & 1;
          mbar_wait(empty_bar0 + 8 * st, ph ^ 1);
          const uint32_t dst = dst0 + st * C::STAGE_BYTES;
#pragma unroll
          for (int i = 0; i < C::NR; ++i) {
            const int64_t off = (int64_t)idx_cur[i] * slab_row;
            cp_async16(dst + i * C::RSTEP * C::ROW_PITCH, kb + off);
            cp_async16(dst + i * C::RSTEP * C::ROW_PITCH + C::KV_BYTES, vb + off);
          }
          cp_async_mbar_arrive(full_bar0 + 8 * st);
          ++stage_ctr;
          advance(is);
        }
      }
      }   // !WPT
    } else {
      // ================================================================ consumer warps
      const int h = warp % HPB;        // KV head inside the group
      const int sub = warp / HPB;      // token sub-group
      const int g = lane >> 2;         // MMA group id: S^T row (token), B-fragment column (q head)
      const int t4 = lane & 3;
      float* my_scr = scratch + warp * C::SCR_FLOATS;

      // ldmatrix lane addresses inside a stage
      const uint32_t k_lane_off =
          ((lane & 7) + ((lane >> 3) & 1) * 8) * C::ROW_PITCH + h * D * 2 + (lane >> 4) * 16;
      const uint32_t v_lane_off = C::KV_BYTES + ((lane & 7) + (lane >> 4) * 8) * C::ROW_PITCH + h * D * 2 +
                                  ((lane >> 3) & 1) * 16;

      for (int ii = 0; ii < n_items; ++ii) {
        const Item it = items[ii];
        if (!it.valid) continue;
        const int ntile = (it.tok_end - it.tok_begin + kTile - 1) / kTile;

        // Q^T B-fragments: column g = q head g of this KV head, rows = head-dim
        uint32_t qb[KS][2];
        {
          const int kvh = it.hg * HPB + h;
          const uint32_t* qrow =
              reinterpret_cast<const uint32_t*>(q + ((int64_t)it.row * n_q_heads + (int64_t)kvh * G + g) * D);
#pragma unroll
          for (int ks = 0; ks < KS; ++ks) {
            qb[ks][0] = g < G ? __ldg(qrow + ks * 8 + t4) : 0u;
            qb[ks][1] = g < G ? __ldg(qrow + ks * 8 + 4 + t4) : 0u;
          }
        }

        float o[KS][4];
#pragma unroll
        for (int mt = 0; mt < KS; ++mt) o[mt][0] = o[mt][1] = o[mt][2] = o[mt][3] = 0.f;
        float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;   // heads 2*t4, 2*t4+1

        for (int ti = 0; ti < ntile; ++ti) {
          const uint32_t sc = stage_ctr + ti;
          if ((int)(sc % NSUB) != sub) continue;
          const uint32_t st = sc % NST;
          const uint32_t ph = (sc / NST) & 1;
          const uint32_t sbase = stages_u32 + st * C::STAGE_BYTES;
          mbar_wait(full_bar0 + 8 * st, ph);

          // ---- S^T = K . Q^T  (two accumulators to shorten the dependent chain)
          float sa[4] = {0.f, 0.f, 0.f, 0.f}, sb[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
          for (int ks = 0; ks < KS; ks += 2) {
            uint32_t a0, a1, a2, a3;
            ldsm_x4(sbase + k_lane_off + ks * 32, a0, a1, a2, a3);
            mma16816<T>(sa, a0, a1, a2, a3, qb[ks][0], qb[ks][1]);
            ldsm_x4(sbase + k_lane_off + (ks + 1) * 32, a0, a1, a2, a3);
            mma16816<T>(sb, a0, a1, a2, a3, qb[ks + 1][0], qb[ks + 1][1]);
          }
          // sa[0]: (token g, head 2t4)  sa[1]: (token g, head 2t4+1)  sa[2]/[3]: token g+8
          float s0 = (sa[0] + sb[0]) * scale_log2, s1 = (sa[1] + sb[1]) * scale_log2;
          float s2 = (sa[2] + sb[2]) * scale_log2, s3 = (sa[3] + sb[3]) * scale_log2;
          const int valid = it.tok_end - (it.tok_begin + ti * kTile);   // >= 1
          if (valid < kTile) {
            if (g >= valid) { s0 = -INFINITY; s1 = -INFINITY; }
            if (g + 8 >= valid) { s2 = -INFINITY; s3 = -INFINITY; }
          }
          // ---- online softmax, base 2
          float mx0 = fmaxf(s0, s2), mx1 = fmaxf(s1, s3);
#pragma unroll
          for (int off = 4; off < 32; off <<= 1) {
            mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, off));
            mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, off));
          }
          const float mn0 = fmaxf(m0, mx0), mn1 = fmaxf(m1, mx1);   // finite: every tile has a valid token
          const float c0 = fast_exp2(m0 - mn0), c1 = fast_exp2(m1 - mn1);
          m0 = mn0;
          m1 = mn1;
          const float p0 = fast_exp2(s0 - mn0), p1 = fast_exp2(s1 - mn1);
          const float p2 = fast_exp2(s2 - mn0), p3 = fast_exp2(s3 - mn1);
          l0 = l0 * c0 + p0 + p2;
          l1 = l1 * c1 + p1 + p3;
          if (__any_sync(0xffffffffu, (c0 != 1.f) || (c1 != 1.f))) {
#pragma unroll
            for (int mt = 0; mt < KS; ++mt) {
              o[mt][0] *= c0; o[mt][1] *= c1; o[mt][2] *= c0; o[mt][3] *= c1;
            }
          }
          // ---- P^T as the B operand: transpose the two 8x8 (token x head) blocks
          const uint32_t pb0 = movmatrix_t(pack2<T>(p0, p1));
          const uint32_t pb1 = movmatrix_t(pack2<T>(p2, p3));
          // ---- O^T += V^T . P^T
#pragma unroll
          for (int mt = 0; mt < KS; ++mt) {
            uint32_t a0, a1, a2, a3;
            ldsm_x4_t(sbase + v_lane_off + mt * 32, a0, a1, a2, a3);
            mma16816<T>(o[mt], a0, a1, a2, a3, pb0, pb1);
          }
          __syncwarp();
          if (lane == 0) mbar_arrive(empty_bar0 + 8 * st);
        }
        stage_ctr += ntile;

        // ---- item epilogue: publish this warp's state, merge the NSUB states of each head
#pragma unroll
        for (int off = 4; off < 32; off <<= 1) {
          l0 += __shfl_xor_sync(0xffffffffu, l0, off);
          l1 += __shfl_xor_sync(0xffffffffu, l1, off);
        }
#pragma unroll
        for (int mt = 0; mt < KS; ++mt) {
          my_scr[(2 * t4) * C::DP + mt * 16 + g] = o[mt][0];
          my_scr[(2 * t4 + 1) * C::DP + mt * 16 + g] = o[mt][1];
          my_scr[(2 * t4) * C::DP + mt * 16 + g + 8] = o[mt][2];
          my_scr[(2 * t4 + 1) * C::DP + mt * 16 + g + 8] = o[mt][3];
        }
        if (g == 0) {
          ml[(warp * 8 + 2 * t4) * 2 + 0] = m0;
          ml[(warp * 8 + 2 * t4) * 2 + 1] = l0;
          ml[(warp * 8 + 2 * t4 + 1) * 2 + 0] = m1;
          ml[(warp * 8 + 2 * t4 + 1) * 2 + 1] = l1;
        }
        named_bar_sync(1, kConsumerWarps * 32);

        // one warp per (kv head, q head) pair per round; lane owns 4 consecutive head-dim values
        for (int pair = warp; pair < HPB * G; pair += kConsumerWarps) {
          const int ph_ = pair / G, pg = pair - ph_ * G;
          float mm = -INFINITY;
#pragma unroll
          for (int s = 0; s < NSUB; ++s) mm = fmaxf(mm, ml[((s * HPB + ph_) * 8 + pg) * 2]);
          float den = 0.f;
          float acc[4] = {0.f, 0.f, 0.f, 0.f};
          const int dd = lane * 4;            // D/4 <= 32 float4 per row
#pragma unroll
          for (int s = 0; s < NSUB; ++s) {
            const int w = s * HPB + ph_;
            const float wm = ml[(w * 8 + pg) * 2], wl = ml[(w * 8 + pg) * 2 + 1];
            const float wgt = (wm == -INFINITY) ? 0.f : fast_exp2(wm - mm);
            den += wl * wgt;
            if (dd < D) {
              const float4 v = *reinterpret_cast<const float4*>(scratch + w * C::SCR_FLOATS + pg * C::DP + dd);
              acc[0] += wgt * v.x; acc[1] += wgt * v.y; acc[2] += wgt * v.z; acc[3] += wgt * v.w;
            }
          }
          const float inv = 1.f / den;
          const int qh = (it.hg * HPB + ph_) * G + pg;
          if (dd < D) {
            const float x0 = acc[0] * inv, x1 = acc[1] * inv, x2 = acc[2] * inv, x3 = acc[3] * inv;
            if (it.part_slot < 0) {
              uint2 pk;
              pk.x = pack2<T>(x0, x1);
              pk.y = pack2<T>(x2, x3);
              *reinterpret_cast<uint2*>(out + ((int64_t)it.row * n_q_heads + qh) * D + dd) = pk;
            } else {
              *reinterpret_cast<float4*>(part_o + ((int64_t)it.part_slot * (HPB * G) + pair) * D + dd) =
                  make_float4(x0, x1, x2, x3);
            }
          }
          if (it.part_slot >= 0 && lane == 0)
            part_lse[(int64_t)it.part_slot * (HPB * G) + pair] = mm + log2f(den);
        }
        named_bar_sync(1, kConsumerWarps * 32);   // scratch may be overwritten by the next item
      }
    }
  }
}

// ------------------------------------------------------------------------------------ launcher
template <typename T, int D, int HPB>
static int launch_paged_attention(const void* q, const void* k_slab, const void* v_slab, const int32_t* idx_base,
                                  const int64_t* row_idx_off, const int32_t* row_kv_len, const int32_t* plan_hdr,
                                  const int32_t* tile_indptr, const int32_t* cta_first_seq, int n_q_heads,
                                  int n_kv_heads, float sm_scale, void* out, float* part_o, float* part_lse,
                                  int n_ctas, cudaStream_t stream) {
  using C = Cfg<D, HPB>;
  auto kern = paged_attention_kernel<T, D, HPB>;
  static thread_local int configured_dev = -1;
  int dev = 0;
  BSGL_CUDA_TRY(cudaGetDevice(&dev));
  if (configured_dev != dev) {
    BSGL_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM_BYTES));
    configured_dev = dev;
  }
  const float scale_log2 = sm_scale * 1.4426950408889634f;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)n_ctas);
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = C::SMEM_BYTES;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  BSGL_CUDA_TRY(cudaLaunchKernelEx(&cfg, kern, (const T*)q, (const T*)k_slab, (const T*)v_slab, idx_base, row_idx_off,
                                   row_kv_len, plan_hdr, tile_indptr, cta_first_seq, n_q_heads, n_kv_heads,
                                   scale_log2, (T*)out, part_o, part_lse));
  return check_launch("paged_attention_kernel");
}

}  // namespace bsgl

extern "C" int bsgl_paged_attention(const void* q, const void* k_slab, const void* v_slab, const int32_t* idx_base,
                                    const int64_t* row_idx_off, const int32_t* row_kv_len, const int32_t* plan_hdr,
                                    const int32_t* tile_indptr, const int32_t* cta_first_seq, int n_rows,
                                    int n_q_heads, int n_kv_heads, int head_dim, float sm_scale, int dtype, void* out,
                                    float* part_o, float* part_lse, int n_ctas, void* stream) {
  using namespace bsgl;
  BSGL_REQUIRE(n_rows >= 0, "bsgl_paged_attention: negative row count");
  if (n_rows == 0) return BSGL_OK;
  BSGL_REQUIRE(q && k_slab && v_slab && idx_base && row_idx_off && row_kv_len && plan_hdr && tile_indptr &&
                   cta_first_seq && out && part_o && part_lse,
               "bsgl_paged_attention: null pointer");
  BSGL_REQUIRE(n_kv_heads > 0 && n_q_heads % n_kv_heads == 0, "bsgl_paged_attention: n_q_heads %% n_kv_heads != 0");
  const int G = n_q_heads / n_kv_heads;
  BSGL_REQUIRE(G <= 8, "bsgl_paged_attention: GQA group size %d > 8 is not supported", G);
  BSGL_REQUIRE(head_dim == 64 || head_dim == 128, "bsgl_paged_attention: head_dim %d not in {64,128}", head_dim);
  BSGL_REQUIRE(dtype == BSGL_BF16 || dtype == BSGL_F16, "bsgl_paged_attention: dtype must be bf16 or fp16");
  BSGL_REQUIRE(n_ctas > 0, "bsgl_paged_attention: n_ctas must be positive");
  BSGL_REQUIRE(((uintptr_t)q % 16 == 0) && ((uintptr_t)k_slab % 16 == 0) && ((uintptr_t)v_slab % 16 == 0) &&
                   ((uintptr_t)out % 16 == 0) && ((uintptr_t)part_o % 16 == 0),
               "bsgl_paged_attention: pointers must be 16-byte aligned");
  const int hpb = bsgl_heads_per_group(n_kv_heads);
  cudaStream_t s = (cudaStream_t)stream;
#define BSGL_PA(T, D, H)                                                                                          \
  return launch_paged_attention<T, D, H>(q, k_slab, v_slab, idx_base, row_idx_off, row_kv_len, plan_hdr,          \
                                         tile_indptr, cta_first_seq, n_q_heads, n_kv_heads, sm_scale, out, part_o, \
                                         part_lse, n_ctas, s)
#define BSGL_PA_D(T, H)                \
  do {                                 \
    if (head_dim == 128) BSGL_PA(T, 128, H); \
    BSGL_PA(T, 64, H);                 \
  } while (0)
#define BSGL_PA_H(T)                   \
  do {                                 \
    if (hpb == 4) BSGL_PA_D(T, 4);     \
    if (hpb == 2) BSGL_PA_D(T, 2);     \
    BSGL_PA_D(T, 1);                   \
  } while (0)
  if (dtype == BSGL_BF16) BSGL_PA_H(__nv_bfloat16);
  BSGL_PA_H(__half);
#undef BSGL_PA_H
#undef BSGL_PA_D
#undef BSGL_PA
}
