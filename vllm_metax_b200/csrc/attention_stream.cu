// Streaming paged-attention decode kernel for B200 (sm_100a): head_size 128,
// block_size 16/32, bf16 queries, bf16 or fp8-e4m3 KV cache, any GQA group
// size up to 16.  Serves paged_attention_v1 and paged_attention_v2
// (reference: csrc/attention/attention_kernels.cuh:79-485).
//
// Design:
//  * persistent grid, one CTA per SM; the CTA serves ALL query heads of a kv
//    head, so every KV byte is read from HBM once (the reference re-reads it
//    once per query head).  Two schedules:
//      FLAT (paged_attention_v1): the 32-token tiles of all (seq, kv head)
//        pairs form one flat sequence cut into equal contiguous ranges, one
//        per CTA — perfect balance, <= 2 split pairs per CTA, and a pair that
//        lies inside one CTA is finished with no global round trip; split
//        pairs are merged by the CTA that delivers the last segment (arrival
//        counters), so the op is a single launch;
//      items (paged_attention_v2): work item = (512-token partition, seq, kv
//        head), partials in the reference's exp_sums/max_logits/tmp_out format;
//  * four producer warps (one issuing lane each, tiles dealt round-robin) walk
//    the block table and stream the contiguous K chunk [D/x][B][x] and V chunk
//    [D][B] of each (block, kv head) into a shared-memory ring with 1-D TMA
//    bulk copies (cp.async.bulk) completing on mbarriers; the ring keeps
//    128-192 KiB per SM in flight;
//  * 8 consumer warps each take whole 32-token tiles out of the ring and run
//    S^T = K.Q^T and O^T += V^T.P^T on mma.sync m16n8k16 with the roles
//    swapped so that the 8-wide N dimension is the GQA group (M = tokens /
//    head dims): no padding rows for G <= 8.  K/V fragments come straight out
//    of the cache layout with ldmatrix (a 16-byte unit of the K chunk is 8
//    dims of one token; a row of the V chunk is 16 tokens of one dim);
//  * online softmax in fp32 with exp2, running max per head kept in
//    registers, probabilities rounded to the MMA operand type (as the
//    reference oracle rounds them to the value dtype), 8x8 transposes by
//    movmatrix; the consumer warps' partial results are merged through shared
//    memory once per work item;
//  * fp8: e4m3 bytes are expanded to f16 in registers (exact), the MMA runs in
//    f16, k_scale is folded into the logit scale and v_scale into the output.
#include <cuda.h>

#include "stream_common.cuh"

namespace b200 {
namespace {

#ifdef B200_TIMELINE
// Developer build (-DB200_TIMELINE): per-CTA %globaltimer stamps of the last streaming launch, read back with
// b200_debug_timeline(): [cta][0] entry, [1] schedule ready, [2] first tile landed (warp 0), [3] last tile consumed,
// [4] merges done / exit.
__device__ unsigned long long g_timeline[256][8];
__device__ __forceinline__ unsigned long long gtime() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
#define TL(slot) do { if ((threadIdx.x & 31) == 0) atomicMax(&g_timeline[blockIdx.x][slot], gtime()); } while (0)
// first tile of the warp only (a stamp per tile would slow the consumers down)
#define TL_FIRST(slot) do { if ((threadIdx.x & 31) == 0 && !tl_seen) { tl_seen = true; atomicMin(&g_timeline[blockIdx.x][slot], gtime()); } } while (0)
#else
#define TL(slot) do { } while (0)
#define TL_FIRST(slot) do { } while (0)
#endif

// ---- the kernel ------------------------------------------------------------------
// PW ("per-warp items", item schedule only): every consumer warp owns WHOLE work items (512-token partitions) and is
// fed by its own producer lane through its own slice of the ring, so an item ends without any cross-warp merge —
// the warp normalises and writes its partial itself.  With fp8 caches the consumers are the busy side and the
// CTA-wide merge per 512-token item (barrier, 32 KiB through shared memory, barrier) cost 0.88 -> 0.66 of peak.
template <int B, bool FP8, int NT, bool FLAT, int NP, bool FLASH = false, bool PW = false>
__global__ void __launch_bounds__((kConsumerWarps + NP) * kWarpSize, 1)
paged_attention_stream_kernel(const __grid_constant__ StreamParams p) {
  using C = Cfg<B, FP8>;
  static_assert(NP == 4 || NP == 8, "whole warpgroups of producers");
  static_assert(!PW || (!FLAT && !FLASH && NT == 1 && NP == kConsumerWarps), "per-warp items: item schedule, one head tile, one producer per consumer");
  static_assert(!FLASH || (!FP8 && B == 16), "flash layout: bf16 cache; the block size is a runtime value");
  // bytes of one ring stage (32 tokens of K and V)
  constexpr int kStage = FLASH ? kFlashStageBytes : C::kStageBytes;
  constexpr bool F16 = FP8;  // MMA operand type: f16 for fp8 caches, bf16 otherwise
  constexpr int NH = 8 * NT;  // padded heads per kv head
  extern __shared__ uint8_t smem_raw[];
  // ring base: 128-byte aligned; 1024 for the flash layout (the 128-byte TMA swizzle pattern repeats every 1 KiB)
  constexpr uintptr_t kAlign = FLASH ? 1024 : 128;
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + kAlign - 1) & ~(kAlign - 1));
  const int nst = p.nstages;
  uint8_t* ring = smem;
  float* merge_o = reinterpret_cast<float*>(ring + (size_t)nst * kStage);
  float* merge_m = merge_o + kConsumerWarps * NH * kMergeStride;
  float* merge_l = merge_m + kConsumerWarps * NH;
  uint64_t* bars = reinterpret_cast<uint64_t*>(merge_l + kConsumerWarps * NH);
  Sched* sched = reinterpret_cast<Sched*>(bars + 2 * nst);
  int* merge_flag = reinterpret_cast<int*>(sched + 1);
  const uint32_t ring_u32 = smem_u32(ring);
  const uint32_t full_u32 = smem_u32(bars);
  const uint32_t empty_u32 = full_u32 + 8u * nst;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
#ifdef B200_TIMELINE
  if (tid == 0) {
    g_timeline[blockIdx.x][0] = gtime();
    g_timeline[blockIdx.x][2] = ~0ull;
    g_timeline[blockIdx.x][3] = 0; g_timeline[blockIdx.x][4] = 0;
    unsigned smid;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    g_timeline[blockIdx.x][5] = smid;  // tools/timeline_probe.py --per-sm
  }
#endif
  if (tid == 32) {  // not warp 0: that one computes the schedule meanwhile
    for (int i = 0; i < nst; ++i) {
      mbar_init(full_u32 + 8u * i, 1);
      mbar_init(empty_u32 + 8u * i, 1);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if constexpr (FLAT) {
    if (warp == 0) compute_sched(p, sched, lane);
  }
  __syncthreads();
#ifdef B200_TIMELINE
  if (tid == 0) g_timeline[blockIdx.x][1] = gtime();
#endif

  if (warp >= kConsumerWarps) {
    // ===================== producer warps ========================================
    // Producer warp pw owns the tiles gt == pw (mod kProducerWarps) of the CTA's tile stream; one
    // lane walks the block table and issues the bulk copies (a single issuing thread costs ~1000
    // cycles per tile, so four of them share the stream).
    if (lane != 0) {
      // The idle lanes of the last producer warp give empty sequences (seq_len 0: CUDA-graph padding) their result:
      // a zero row, as the reference kernels leave (acc = 0 times 1/(0 + 1e-6), csrc/attention/attention_kernels.cuh:331),
      // and lse = -inf.  Such sequences own no tile, so no consumer ever writes them.
      if (warp == kConsumerWarps + NP - 1) {
        for (int s = (int)blockIdx.x; s < p.S; s += (int)gridDim.x) {
          if (__ldg(p.seq_lens + s) > 0) continue;
          uint4* row = reinterpret_cast<uint4*>(p.out + (int64_t)s * p.H * kD);
          for (int i = lane - 1; i < p.H * (kD * 2 / 16); i += 31) row[i] = make_uint4(0u, 0u, 0u, 0u);
          if (p.lse != nullptr)
            for (int h = lane - 1; h < p.H; h += 31) p.lse[(int64_t)h * p.S + s] = -INFINITY;
        }
      }
      return;
    }
    if constexpr (FLASH) {
      producer_loop_flash<FLAT, NP>(p, &p.tm_k, &p.tm_v, sched, warp - kConsumerWarps, ring_u32, full_u32, empty_u32, nst);
      return;
    }
    if constexpr (PW) {
      producer_loop_pw<B, FP8>(p, warp - kConsumerWarps, ring_u32, full_u32, empty_u32, nst / kConsumerWarps);
      return;
    }
    producer_loop<B, FP8, FLAT, NP>(p, sched, warp - kConsumerWarps, ring_u32, full_u32, empty_u32, nst);
    return;
  }

  // ===================== consumer warps ============================================
#ifdef B200_TIMELINE
  bool tl_seen = false;
#endif
  const int g = lane >> 2, c = lane & 3;
  // ldmatrix lane addressing (byte offsets inside a K / V chunk)
  const int mi = lane >> 3, r8 = lane & 7;
  uint32_t k_lane_off, v_lane_off;
  [[maybe_unused]] uint32_t k_swz = 0, v_swz = 0;  // flash layout: swizzle term of this lane's 16-byte chunk
  if constexpr (FLASH) {
    // a box is 16 token rows of 128 B (64 dims), 16-byte chunk c of row r stored at chunk c ^ (r & 7) (128B swizzle).
    // K (A = K, rows = tokens): matrices (tok 0-7, dims 0-7) (tok 8-15, dims 0-7) (tok 0-7, dims 8-15) (tok 8-15, dims 8-15):
    // row = (mi&1)*8 + r8, chunk = 2*kk + (mi>>1)  ->  byte = row*128 + ((kk*32) ^ k_swz)
    k_lane_off = (uint32_t)(((mi & 1) * 8 + r8) * 128);
    k_swz = (uint32_t)(((mi >> 1) ^ r8) << 4);
    // V (A = V^T, rows = dims) through ldmatrix.trans: stored matrices (tok 0-7, dims 0-7) (tok 0-7, dims 8-15)
    // (tok 8-15, dims 0-7) (tok 8-15, dims 8-15) come out as (dim g | tok 2c, 2c+1) fragments a0..a3
    v_lane_off = (uint32_t)(((mi >> 1) * 8 + r8) * 128);
    v_swz = (uint32_t)(((mi & 1) ^ r8) << 4);
  } else if constexpr (!FP8) {
    // K: matrices (tok 0-7,g0) (tok 8-15,g0) (tok 0-7,g1) (tok 8-15,g1); unit = 16 B = 8 dims of a token
    k_lane_off = (uint32_t)(((mi >> 1) * B + (mi & 1) * 8 + r8) * 16);
    // V: matrices (dim 0-7,tok 0-7) (dim 8-15,tok 0-7) (dim 0-7,tok 8-15) (dim 8-15,tok 8-15)
    v_lane_off = (uint32_t)((((mi & 1) * 8 + r8) * B * 2) + (mi >> 1) * 16);
  } else {
    // K: unit = 16 B = 16 dims of a token; rows are permuted tokens (see tok_of_row)
    const int tok = (r8 >> 1) * 4 + (r8 & 1) + 2 * (mi & 1);
    k_lane_off = (uint32_t)(((mi >> 1) * B + tok) * 16);
    // V: row = dim (B bytes, 16 tokens per 16 B); four matrices = 32 consecutive dims
    v_lane_off = (uint32_t)((mi * 8 + r8) * B);
  }
  // token (inside its 16-token half) of accumulator rows g and g+8
  const int row_tok = FP8 ? ((g >> 1) * 4 + (g & 1)) : g;
  constexpr int kRowHiDelta = FP8 ? 2 : 8;
  // tokens (inside a 16-token half) held by V fragment registers a0/a1 (lo) and a2/a3 (hi)
  const int vtok_lo = FP8 ? 4 * c : 2 * c;
  const int vtok_hi = FP8 ? 4 * c + 2 : 2 * c + 8;

  const float ksc = FP8 ? __ldg(p.k_scale) : 1.f;
  const float vsc = FP8 ? __ldg(p.v_scale) : 1.f;

  uint32_t tile_base = 0;  // tiles issued to the ring before this segment
  // ring stage / phase of this warp's next tile (PW: the warp's own slice [warp * spw, (warp + 1) * spw) of the ring)
  [[maybe_unused]] const uint32_t spw = (uint32_t)nst / kConsumerWarps;
  [[maybe_unused]] uint32_t lstage = 0;
  uint32_t stage = PW ? (uint32_t)warp * spw : (uint32_t)warp, parity = 0;
  WorkIter<FLAT> wi;
  wi.init(p, sched);
  if constexpr (PW) wi.init_warp_items(warp);  // items k == warp (mod 8) of the CTA's list, as the warp's producer lane
  auto next_item = [&](Seg& sg) -> bool { return wi.next(p, sg); };
  // ---- Q fragments (B operand: k = dims, n = heads) of a segment --------------------------------------
  // Loaded in two steps so that the global-memory latency of the NEXT segment's query overlaps the merge of the
  // current one (with 512-token items the exposed wait was ~1/3 of the per-item overhead): q_issue only issues the
  // loads (into qf for bf16, raw words for fp8), q_finish converts and derives the per-head scalars.
  uint32_t qf[NT][8][2];
  [[maybe_unused]] uint2 qraw[NT][8];
  float sc2[NT][2];   // logit multiplier (log2 units) of heads 2c, 2c+1
  float al2[NT][2];   // alibi slope * log2(e)
  auto q_issue = [&](const Seg& sg) {
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      const int hl = nt * 8 + g;  // head inside the group owned by this lane's B column
      const bool hv = hl < p.G;
      const __nv_bfloat16* qrow = p.q + (int64_t)sg.s * p.q_stride + (int64_t)(sg.kvh * p.G + hl) * kD;
#pragma unroll
      for (int ks = 0; ks < 8; ++ks) {
        if constexpr (!FP8) {
          qf[nt][ks][0] = hv ? __ldg(reinterpret_cast<const uint32_t*>(qrow + 16 * ks + 2 * c)) : 0u;
          qf[nt][ks][1] = hv ? __ldg(reinterpret_cast<const uint32_t*>(qrow + 16 * ks + 8 + 2 * c)) : 0u;
        } else {
          qraw[nt][ks] = hv ? __ldg(reinterpret_cast<const uint2*>(qrow + 16 * ks + 4 * c)) : make_uint2(0u, 0u);
        }
      }
    }
  };
  auto q_finish = [&](const Seg& sg) {
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      float qscale_inv = 1.f;
      if constexpr (FP8) {
        // bf16 -> f16 with a per-head power-of-two prescale so that no value
        // overflows f16; the prescale is undone in the logit multiplier.
        float qv[8][4];
        float amax = 0.f;
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
          const uint2 raw = qraw[nt][ks];
          const __nv_bfloat162 lo = *reinterpret_cast<const __nv_bfloat162*>(&raw.x);
          const __nv_bfloat162 hi = *reinterpret_cast<const __nv_bfloat162*>(&raw.y);
          qv[ks][0] = __low2float(lo); qv[ks][1] = __high2float(lo);
          qv[ks][2] = __low2float(hi); qv[ks][3] = __high2float(hi);
#pragma unroll
          for (int e = 0; e < 4; ++e) amax = fmaxf(amax, fabsf(qv[ks][e]));
        }
        amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 1));
        amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 2));
        float qscale = 1.f;
        if (amax > 16384.f) {  // 2^14; keep |q| * qscale <= 2^14
          int ex;
          frexpf(amax, &ex);          // amax = f * 2^ex, f in [0.5, 1)
          qscale = ldexpf(1.f, 14 - ex);
          qscale_inv = ldexpf(1.f, ex - 14);
        }
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
          qf[nt][ks][0] = pack2<true>(qv[ks][0] * qscale, qv[ks][1] * qscale);
          qf[nt][ks][1] = pack2<true>(qv[ks][2] * qscale, qv[ks][3] * qscale);
        }
      }
      // per-column (head 2c / 2c+1) scalars live in the lanes with g = 2c, 2c+1
      const float mine = p.scale * kLog2e * ksc * qscale_inv;
      sc2[nt][0] = __shfl_sync(0xffffffffu, mine, (2 * c) * 4);
      sc2[nt][1] = __shfl_sync(0xffffffffu, mine, (2 * c + 1) * 4);
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const int hc = nt * 8 + 2 * c + e;
        al2[nt][e] = (p.alibi != nullptr && hc < p.G) ? __ldg(p.alibi + sg.kvh * p.G + hc) * kLog2e : 0.f;
      }
    }
  };

  // Prefetching costs registers across the merge: worth it (and spill-free) for the item schedule with one head
  // tile, where a CTA walks dozens of short items; the flat schedule has <= 3 segments per CTA.
  constexpr bool kPrefetchQ = !FLAT && NT == 1;
  Seg it, nx;
  int item = -1;
  [[maybe_unused]] int seen_cnt = 0;  // flat schedule, thread 0: arrival counter of the segment's pair as the segment began
  bool has = next_item(nx);
  if (kPrefetchQ && has) q_issue(nx);
  while (has) {
    it = nx;
    ++item;
    const int ntiles = (it.t1 - it.t0 + kTileTokens - 1) / kTileTokens;
    if constexpr (!kPrefetchQ) q_issue(it);
    if constexpr (FLAT) {
      // A piece of a split pair: have ALL the other pieces been delivered already?  Then this CTA is the one that
      // merges, and it can do so straight from its own result — no partial written, no fence, no counter round trip
      // (the case of nearly every CTA's LAST segment: the rest of that pair was the next CTA's FIRST segment, long
      // done).  The counter is read now — the load is in flight during the tile loop, its value is looked at when the
      // segment ends — so the answer can only turn out pessimistic.
      if (tid == 0 && !(it.starts && it.ends))
        asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(seen_cnt) : "l"(p.counters + it.s * p.Hkv + it.kvh) : "memory");
    }
    q_finish(it);
    const bool use_alibi = p.alibi != nullptr;

    // ---- running state ----------------------------------------------------------------
    float oacc[NT][8][4];
    float mrun[NT][2], lrun[NT][2];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      mrun[nt][0] = mrun[nt][1] = -INFINITY;
      lrun[nt][0] = lrun[nt][1] = 0.f;
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int e = 0; e < 4; ++e) oacc[nt][i][e] = 0.f;
    }

    // Tile gt of the CTA's tile stream goes to warp gt % 8 and ring stage gt % nst; nst is a multiple
    // of 8, so every use of a stage is consumed by the same warp, in order.  (A warp that skipped a
    // phase of a stage could pass a parity wait one round early: mbarrier parity only tells odd from
    // even.)
    // one 32-token tile at shared-memory address sb, first token tok0: S^T = K.Q^T, online softmax, O^T += V^T.P^T
    auto do_tile = [&](const uint32_t sb, const int tok0) {
      const int nvalid = min(kTileTokens, it.t1 - tok0);
      const bool partial = nvalid < kTileTokens;

      // ---- S^T = K . Q^T : rows = tokens, cols = heads --------------------------------
      float sacc[2][NT][4];
#pragma unroll
      for (int j = 0; j < 2; ++j)
#pragma unroll
        for (int nt = 0; nt < NT; ++nt)
#pragma unroll
          for (int e = 0; e < 4; ++e) sacc[j][nt][e] = 0.f;
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        // half j: block j (B = 16) or tokens 16j.. of the block (B = 32)
        const uint32_t kb = FLASH ? sb + j * 2 * kFlashBoxBytes + k_lane_off
                                  : sb + (B == 16 ? j * C::kBlkBytes : j * 16 * 16) + k_lane_off;
        if constexpr (FLASH) {
#pragma unroll
          for (int ks = 0; ks < 8; ++ks) {
            uint32_t a[4];
            ldmatrix_x4(a, kb + (ks >> 2) * kFlashBoxBytes + (((ks & 3) * 32) ^ k_swz));
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) mma16816<F16>(sacc[j][nt], a, qf[nt][ks]);
          }
        } else if constexpr (!FP8) {
#pragma unroll
          for (int ks = 0; ks < 8; ++ks) {
            uint32_t a[4];
            ldmatrix_x4(a, kb + ks * 2 * B * 16);
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) mma16816<F16>(sacc[j][nt], a, qf[nt][ks]);
          }
        } else {
#pragma unroll
          for (int kk = 0; kk < 4; ++kk) {
            uint32_t r[4];
            ldmatrix_x4(r, kb + kk * 2 * B * 16);
            uint32_t a0[4], a1[4];
            cvt_e4m3x4_f16x2x2(r[0], a0[0], a0[2]);
            cvt_e4m3x4_f16x2x2(r[1], a0[1], a0[3]);
            cvt_e4m3x4_f16x2x2(r[2], a1[0], a1[2]);
            cvt_e4m3x4_f16x2x2(r[3], a1[1], a1[3]);
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) {
              mma16816<F16>(sacc[j][nt], a0, qf[nt][2 * kk]);
              mma16816<F16>(sacc[j][nt], a1, qf[nt][2 * kk + 1]);
            }
          }
        }
      }

      // ---- online softmax ------------------------------------------------------------------
      // accumulator element e of half j: token = 16j + row_tok + (e>>1)*kRowHiDelta, head = 2c + (e&1)
      uint32_t pf[2][NT][2];
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
        float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
        for (int j = 0; j < 2; ++j)
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const int tt = 16 * j + row_tok + (e >> 1) * kRowHiDelta;  // token inside the tile
            float v = sacc[j][nt][e] * sc2[nt][e & 1];
            if (use_alibi) v = fmaf(al2[nt][e & 1], (float)(tok0 + tt - it.L + 1), v);
            if (partial && tt >= nvalid) v = -INFINITY;
            sacc[j][nt][e] = v;
            mx[e & 1] = fmaxf(mx[e & 1], v);
          }
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          mx[e] = fmaxf(mx[e], __shfl_xor_sync(0xffffffffu, mx[e], 4));
          mx[e] = fmaxf(mx[e], __shfl_xor_sync(0xffffffffu, mx[e], 8));
          mx[e] = fmaxf(mx[e], __shfl_xor_sync(0xffffffffu, mx[e], 16));
        }
        float mnew[2], corr[2];
        bool changed = false;
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          mnew[e] = fmaxf(mrun[nt][e], mx[e]);
          changed |= mnew[e] > mrun[nt][e];
        }
        if (__any_sync(0xffffffffu, changed)) {
#pragma unroll
          for (int e = 0; e < 2; ++e) {
            corr[e] = fast_exp2(mrun[nt][e] - mnew[e]);  // first tile: exp2(-inf) = 0
            mrun[nt][e] = mnew[e];
            lrun[nt][e] *= corr[e];
          }
#pragma unroll
          for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int e = 0; e < 4; ++e) oacc[nt][i][e] *= corr[e & 1];
        }
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          float pv[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            pv[e] = fast_exp2(sacc[j][nt][e] - mrun[nt][e & 1]);
            lrun[nt][e & 1] += pv[e];
          }
          // rows g (tokens lo) and g+8 (tokens hi) -> transposed B fragments
          pf[j][nt][0] = movmatrix_trans(pack2<F16>(pv[0], pv[1]));
          pf[j][nt][1] = movmatrix_trans(pack2<F16>(pv[2], pv[3]));
        }
      }

      // ---- O^T += V^T . P^T : rows = head dims, cols = heads -------------------------------
      // Two copies of the loop (full tile / last, partial tile of a sequence): with a runtime predicate
      // ptxas emits the 64 masking LOP3s as predicated instructions that occupy issue slots on every tile.
      auto pv = [&](auto partial_tag) {
        constexpr bool kPartial = decltype(partial_tag)::value;
        uint32_t vmask[2][2];
        if constexpr (kPartial) {
#pragma unroll
          for (int j = 0; j < 2; ++j) {
            const int tl = 16 * j + vtok_lo, th = 16 * j + vtok_hi;
            vmask[j][0] = (tl < nvalid ? 0x0000ffffu : 0u) | (tl + 1 < nvalid ? 0xffff0000u : 0u);
            vmask[j][1] = (th < nvalid ? 0x0000ffffu : 0u) | (th + 1 < nvalid ? 0xffff0000u : 0u);
          }
        }
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const uint32_t vb = FLASH ? sb + kFlashVOffset + j * 2 * kFlashBoxBytes + v_lane_off
                                    : sb + C::kVOffset + (B == 16 ? j * C::kBlkBytes : j * 16 * C::kElemBytes) + v_lane_off;
          if constexpr (FLASH) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              uint32_t a[4];
              ldmatrix_x4_trans(a, vb + (i >> 2) * kFlashBoxBytes + (((i & 3) * 32) ^ v_swz));
              if constexpr (kPartial) {  // rows past seq_len: stale cache (or a half that was not loaded)
                a[0] &= vmask[j][0]; a[1] &= vmask[j][0];
                a[2] &= vmask[j][1]; a[3] &= vmask[j][1];
              }
#pragma unroll
              for (int nt = 0; nt < NT; ++nt) mma16816<F16>(oacc[nt][i], a, pf[j][nt]);
            }
          } else if constexpr (!FP8) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              uint32_t a[4];
              ldmatrix_x4(a, vb + i * 16 * B * 2);
              if constexpr (kPartial) {  // stale cache past seq_len may hold NaN/Inf: zero it
                a[0] &= vmask[j][0]; a[1] &= vmask[j][0];
                a[2] &= vmask[j][1]; a[3] &= vmask[j][1];
              }
#pragma unroll
              for (int nt = 0; nt < NT; ++nt) mma16816<F16>(oacc[nt][i], a, pf[j][nt]);
            }
          } else {
#pragma unroll
            for (int ii = 0; ii < 4; ++ii) {
              uint32_t r[4];
              ldmatrix_x4(r, vb + ii * 32 * B);
              uint32_t a0[4], a1[4];
              cvt_e4m3x4_f16x2x2(r[0], a0[0], a0[2]);
              cvt_e4m3x4_f16x2x2(r[1], a0[1], a0[3]);
              cvt_e4m3x4_f16x2x2(r[2], a1[0], a1[2]);
              cvt_e4m3x4_f16x2x2(r[3], a1[1], a1[3]);
              if constexpr (kPartial) {
                a0[0] &= vmask[j][0]; a0[1] &= vmask[j][0]; a0[2] &= vmask[j][1]; a0[3] &= vmask[j][1];
                a1[0] &= vmask[j][0]; a1[1] &= vmask[j][0]; a1[2] &= vmask[j][1]; a1[3] &= vmask[j][1];
              }
#pragma unroll
              for (int nt = 0; nt < NT; ++nt) {
                mma16816<F16>(oacc[nt][2 * ii], a0, pf[j][nt]);
                mma16816<F16>(oacc[nt][2 * ii + 1], a1, pf[j][nt]);
              }
            }
          }
        }
      };
      if (partial) pv(std::true_type{}); else pv(std::false_type{});
    };

    if constexpr (PW) {
      // the warp walks ALL tiles of its item through its own ring slice
      for (int ti = 0; ti < ntiles; ++ti) {
        MBAR_WAIT(full_u32 + 8u * stage, parity, 2, item, ti);
        TL_FIRST(2);
        do_tile(ring_u32 + stage * kStage, it.t0 + ti * kTileTokens);
        __threadfence_block();  // see the note in the shared-ring loop below
        __syncwarp();
        if (lane == 0) mbar_arrive(empty_u32 + 8u * stage);
        if (++lstage == spw) {
          lstage = 0;
          parity ^= 1u;
        }
        stage = (uint32_t)warp * spw + lstage;
      }
    }
    for (int ti = PW ? ntiles : (int)((warp - tile_base) & (kConsumerWarps - 1)); ti < ntiles; ti += kConsumerWarps) {
      if constexpr (FLAT) {
        // the arrival counter once more, one tile of this warp (~8 tiles of the CTA) before the segment ends: a piece
        // whose partner was still running when the segment began (pairs longer than a CTA's share) may have it now
        if (tid == 0 && ti + kConsumerWarps >= ntiles && !(it.starts && it.ends))
          asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(seen_cnt) : "l"(p.counters + it.s * p.Hkv + it.kvh) : "memory");
      }
      // this warp's tiles are gt = warp, warp + 8, ... over the whole stream, so stage / parity advance
      // incrementally (no division by the runtime ring depth)
      MBAR_WAIT(full_u32 + 8u * stage, parity, 2, item, (int)(tile_base + (uint32_t)ti));
      TL_FIRST(2);
      do_tile(ring_u32 + stage * kStage, it.t0 + ti * kTileTokens);
      // All lanes have consumed the stage: hand it back to the producer.  The fence is REQUIRED: ptxas
      // issues the mbarrier arrive without waiting for the LDSMs still in flight (it even hoists it above the
      // last HMMAs), and with bank-conflicted V fragments (B = 32 bf16: 64-byte rows, 4-way conflicts) the
      // shared-memory queue is deep enough for the refill of the stage to land before those reads are
      // performed (seen as run-to-run differences in output dims 112-127).  MEMBAR.CTA orders them.
      __threadfence_block();
      __syncwarp();
      if (lane == 0) mbar_arrive(empty_u32 + 8u * stage);
      stage += kConsumerWarps;
      if (stage >= (uint32_t)nst) {
        stage -= (uint32_t)nst;
        parity ^= 1u;
      }
    }
    tile_base += (uint32_t)ntiles;
    TL(3);
    if constexpr (kPrefetchQ) {
      has = next_item(nx);
      if (has) q_issue(nx);  // in flight during the merge below
    }

    if constexpr (PW) {
      // ---- per-warp items: the warp finishes its partition alone ----------------------------------------
      float lsum[2], inv[2];
#pragma unroll
      for (int e = 0; e < 2; ++e) {  // sum over the 8 lanes that share a head column
        float l = lrun[0][e];
        l += __shfl_xor_sync(0xffffffffu, l, 4);
        l += __shfl_xor_sync(0xffffffffu, l, 8);
        l += __shfl_xor_sync(0xffffffffu, l, 16);
        lsum[e] = l;
        inv[e] = vsc * __fdividef(1.f, l + 1e-6f);
      }
      // normalised rows staged in the warp's scratch [8 heads][128 dims] (the merge area is unused in this mode),
      // then written with 16-byte stores: a head's row is 256 contiguous bytes in out and in tmp_out
      __nv_bfloat16* scr = reinterpret_cast<__nv_bfloat16*>(merge_o) + warp * (8 * kD);
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int e = 0; e < 4; ++e)
          scr[(2 * c + (e & 1)) * kD + 16 * i + g + 8 * (e >> 1)] = __float2bfloat16_rn(oacc[0][i][e] * inv[e & 1]);
      __syncwarp();
      const bool single = it.L <= p.so.chunk;
      const bool want_partials = p.so.o != nullptr;  // normalised format (the host selects this mode only for it)
      for (int ch = lane; ch < 8 * 16; ch += 32) {
        const int h = ch >> 4, part16 = ch & 15;
        if (h >= p.G) break;
        const uint4 v = *reinterpret_cast<const uint4*>(scr + h * kD + part16 * 8);
        const int64_t row = (int64_t)it.s * p.H + it.kvh * p.G + h;
        if (single) *reinterpret_cast<uint4*>(p.out + row * kD + part16 * 8) = v;
        if (want_partials)
          *reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(p.so.o) + (row * p.so.num_partitions + it.part) * kD + part16 * 8) = v;
      }
      if (g == 0) {
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int h = 2 * c + e;
          if (h < p.G) {
            const int head = it.kvh * p.G + h;
            if (want_partials) {
              const int64_t pi = ((int64_t)it.s * p.H + head) * p.so.num_partitions + it.part;
              p.so.m[pi] = mrun[0][e] * kLn2;
              p.so.l[pi] = lsum[e];
            }
            if (single && p.lse != nullptr)
              p.lse[(int64_t)head * p.S + it.s] = lsum[e] > 0.f ? fmaf(mrun[0][e], kLn2, logf(lsum[e])) : -INFINITY;
          }
        }
      }
      __syncwarp();  // the scratch is reused by this warp's next item
      TL(4);
      continue;
    }
    // ---- merge the consumer warps' partial results ------------------------------------------
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
#pragma unroll
      for (int e = 0; e < 2; ++e) {  // sum over the 8 lanes that share a head column
        lrun[nt][e] += __shfl_xor_sync(0xffffffffu, lrun[nt][e], 4);
        lrun[nt][e] += __shfl_xor_sync(0xffffffffu, lrun[nt][e], 8);
        lrun[nt][e] += __shfl_xor_sync(0xffffffffu, lrun[nt][e], 16);
      }
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int h = nt * 8 + 2 * c + (e & 1);
          const int d = 16 * i + g + 8 * (e >> 1);
          merge_o[(warp * NH + h) * kMergeStride + d] = oacc[nt][i][e];
        }
      if (g == 0) {
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          merge_m[warp * NH + nt * 8 + 2 * c + e] = mrun[nt][e];
          merge_l[warp * NH + nt * 8 + 2 * c + e] = lrun[nt][e];
        }
      }
    }
    if constexpr (FLAT) {
      if (tid == 0 && !(it.starts && it.ends)) {
        // c0 .. c1: the ranges (static shares and dynamic chunks, numbered in tile order) that hold a piece of this pair
        const int c0 = range_of_tile(sched, it.pb);
        const int c1 = range_of_tile(sched, it.pb + it.n - 1);
        merge_flag[1] = seen_cnt == c1 - c0;
        merge_flag[2] = c0;
        merge_flag[3] = c1;
      }
    }
    consumer_bar_sync();
    const int d = tid & (kD - 1);
    if constexpr (!FLAT) {
      // ---- item schedule: one partition of (seq, kv head) --------------------------------------------
      const bool single = it.L <= p.so.chunk;
      const bool want_partials = p.so.o != nullptr && (!single || p.so.normalized);
      for (int h = tid >> 7; h < NH; h += (kConsumerWarps * kWarpSize) >> 7) {
        if (h >= p.G) break;
        float M = -INFINITY;
#pragma unroll
        for (int w = 0; w < kConsumerWarps; ++w) M = fmaxf(M, merge_m[w * NH + h]);
        float lsum = 0.f, o = 0.f;
#pragma unroll
        for (int w = 0; w < kConsumerWarps; ++w) {
          const float wgt = exp2f(merge_m[w * NH + h] - M);  // idle warps: exp2(-inf) = 0
          lsum = fmaf(merge_l[w * NH + h], wgt, lsum);
          o = fmaf(merge_o[(w * NH + h) * kMergeStride + d], wgt, o);
        }
        o *= vsc;
        const int head = it.kvh * p.G + h;
        const float on = o * __fdividef(1.f, lsum + 1e-6f);
        if (single) {
          p.out[((int64_t)it.s * p.H + head) * kD + d] = __float2bfloat16_rn(on);
          if (d == 0 && p.lse != nullptr)
            p.lse[(int64_t)head * p.S + it.s] = lsum > 0.f ? fmaf(M, kLn2, logf(lsum)) : -INFINITY;
        }
        if (want_partials) {
          const int64_t pi = ((int64_t)it.s * p.H + head) * p.so.num_partitions + it.part;
          if (p.so.normalized)
            reinterpret_cast<__nv_bfloat16*>(p.so.o)[pi * kD + d] = __float2bfloat16_rn(on);
          else
            reinterpret_cast<float*>(p.so.o)[pi * kD + d] = o;
          if (d == 0) {
            p.so.m[pi] = M * kLn2;
            p.so.l[pi] = lsum;
          }
        }
      }
    } else {
      // ---- flat schedule: a whole pair is finished here; a split pair leaves a partial in this CTA's
      // head / tail slot and the CTA that delivers the pair's last segment merges them (same arithmetic
      // as the reference reduce, csrc/attention/attention_kernels.cuh:594-657) ----------------------------
      const bool whole = it.starts && it.ends;
      const int slot = (it.rid * 2 + (it.starts ? 1 : 0)) * p.G;
      // split pair (the values were written by thread 0 when the segment began, barriers in between)
      const bool fast = !whole && merge_flag[1] != 0;
      const int c0 = merge_flag[2], c1 = merge_flag[3];
      if (fast) __threadfence();  // acquire side: the other pieces' partials, published before the counter moved
      for (int h = tid >> 7; h < NH; h += (kConsumerWarps * kWarpSize) >> 7) {
        if (h >= p.G) break;
        float M = -INFINITY;
#pragma unroll
        for (int w = 0; w < kConsumerWarps; ++w) M = fmaxf(M, merge_m[w * NH + h]);
        float lsum = 0.f, o = 0.f;
#pragma unroll
        for (int w = 0; w < kConsumerWarps; ++w) {
          const float wgt = exp2f(merge_m[w * NH + h] - M);
          lsum = fmaf(merge_l[w * NH + h], wgt, lsum);
          o = fmaf(merge_o[(w * NH + h) * kMergeStride + d], wgt, o);
        }
        o *= vsc;
        if (fast) {
          // the merge of the `last` path below, with this CTA's piece taken from registers instead of its slot:
          // same operands, same order (ranges c0 .. c1), hence the same bits
          float Mg = M;
          for (int c = c0; c <= c1; ++c)
            if (c != it.rid) Mg = fmaxf(Mg, __ldcg(p.ws_m + (c * 2 + (c == c0 ? 1 : 0)) * p.G + h));
          float lg = 0.f, og = 0.f;
          for (int c = c0; c <= c1; ++c) {
            const int sl = (c * 2 + (c == c0 ? 1 : 0)) * p.G + h;
            const bool mine = c == it.rid;
            const float wgt = exp2f((mine ? M : __ldcg(p.ws_m + sl)) - Mg);
            lg = fmaf(mine ? lsum : __ldcg(p.ws_l + sl), wgt, lg);
            og = fmaf(mine ? o : __ldcg(p.ws_o + (int64_t)sl * kD + d), wgt, og);
          }
          p.out[((int64_t)it.s * p.H + it.kvh * p.G + h) * kD + d] = __float2bfloat16_rn(og * __fdividef(1.f, lg + 1e-6f));
          if (d == 0 && p.lse != nullptr)
            p.lse[(int64_t)(it.kvh * p.G + h) * p.S + it.s] = fmaf(Mg, kLn2, logf(lg));
        } else if (whole) {
          p.out[((int64_t)it.s * p.H + it.kvh * p.G + h) * kD + d] =
              __float2bfloat16_rn(o * __fdividef(1.f, lsum + 1e-6f));
          if (d == 0 && p.lse != nullptr)
            p.lse[(int64_t)(it.kvh * p.G + h) * p.S + it.s] = lsum > 0.f ? fmaf(M, kLn2, logf(lsum)) : -INFINITY;
        } else {
          p.ws_o[(int64_t)(slot + h) * kD + d] = o;
          if (d == 0) {
            p.ws_m[slot + h] = M;  // log2 units
            p.ws_l[slot + h] = lsum;
          }
        }
      }
      if (fast) {
        if (tid == 0) p.counters[it.s * p.Hkv + it.kvh] = 0;  // leave the counters zeroed for the next launch
      } else if (!whole) {
        __threadfence();       // this thread's partials are visible device-wide
        consumer_bar_sync();   // ... and so are everyone's
        if (tid == 0) {
          int* cnt = p.counters + it.s * p.Hkv + it.kvh;
          const int old = atomicAdd(cnt, 1);
          const int last = old == c1 - c0;
          if (last) *cnt = 0;  // leave the counters zeroed for the next launch
          *merge_flag = last;
        }
        consumer_bar_sync();
        if (*merge_flag) {
          __threadfence();
          for (int h = tid >> 7; h < p.G; h += (kConsumerWarps * kWarpSize) >> 7) {
            // range c0 holds the segment that starts the pair (tail slot), the others their head slot
            float M = -INFINITY;
            for (int c = c0; c <= c1; ++c)
              M = fmaxf(M, __ldcg(p.ws_m + (c * 2 + (c == c0 ? 1 : 0)) * p.G + h));
            float lsum = 0.f, o = 0.f;
            for (int c = c0; c <= c1; ++c) {
              const int sl = (c * 2 + (c == c0 ? 1 : 0)) * p.G + h;
              const float wgt = exp2f(__ldcg(p.ws_m + sl) - M);
              lsum = fmaf(__ldcg(p.ws_l + sl), wgt, lsum);
              o = fmaf(__ldcg(p.ws_o + (int64_t)sl * kD + d), wgt, o);
            }
            p.out[((int64_t)it.s * p.H + it.kvh * p.G + h) * kD + d] =
                __float2bfloat16_rn(o * __fdividef(1.f, lsum + 1e-6f));
            if (d == 0 && p.lse != nullptr)
              p.lse[(int64_t)(it.kvh * p.G + h) * p.S + it.s] = fmaf(M, kLn2, logf(lsum));
          }
        }
      }
    }
    consumer_bar_sync();  // merge buffers are free again
    TL(4);
    if constexpr (!kPrefetchQ) has = next_item(nx);
  }
}

template <int B, bool FP8, int NT>
size_t stream_smem_bytes(int nstages) {
  using C = Cfg<B, FP8>;
  return 128 + (size_t)nstages * C::kStageBytes +
         (size_t)kConsumerWarps * 8 * NT * (kMergeStride + 2) * sizeof(float) + (size_t)nstages * 16 + sizeof(Sched) + 16;
}

constexpr size_t kMaxSmem = 227 * 1024;

template <int B, bool FP8, int NT, bool FLAT, int NP, bool PW = false>
int launch_stream_np(StreamParams& sp, cudaStream_t s) {
  using C = Cfg<B, FP8>;
  // ring depth: a multiple of the consumer-warp count (see the tile -> warp mapping in the kernel), capped at
  // 128 KiB in flight — the random-chunk gather does not get faster beyond that, and with 2 KiB chunks it gets
  // slower (profiles/r1_gather_read_bench.log, r1_sweep.log: fp8 16 stages 5.67 TB/s vs 24 stages 5.23 TB/s).
  // Per-warp items: every warp's slice is its own pipeline, three stages deep (two leave the warp waiting on HBM
  // latency: 8 warps x 2 x 8 KiB / ~3 us < the 47 GB/s an SM has to sustain).
  int nst = kConsumerWarps;
  while (stream_smem_bytes<B, FP8, NT>(nst + kConsumerWarps) <= kMaxSmem &&
         (size_t)(nst + kConsumerWarps) * C::kStageBytes <= (PW ? 192 : 128) * 1024)
    nst += kConsumerWarps;
  if (const int want = env_int("B200_STREAM_NSTAGES")) nst = max(kConsumerWarps, min(nst, want / kConsumerWarps * kConsumerWarps));
  sp.nstages = nst;
  const size_t smem = stream_smem_bytes<B, FP8, NT>(nst);
  auto kern = paged_attention_stream_kernel<B, FP8, NT, FLAT, NP, false, PW>;
  static thread_local int configured_device = -1;
  int dev = 0;
  cudaGetDevice(&dev);
  if (configured_device != dev) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxSmem);
    if (e != cudaSuccess) {
      set_error("cudaFuncSetAttribute(smem) failed: %s", cudaGetErrorString(e));
      return B200_ERR_CUDA;
    }
    configured_device = dev;
  }
  const int grid = min(min(device_sm_count(), kFlatMaxCtas), sp.total_items);
  kern<<<grid, (kConsumerWarps + NP) * kWarpSize, smem, s>>>(sp);
  count_launch();
  return check_launch("paged_attention_stream");
}

// Issuing warps: random 2 KiB chunks (fp8, block_size 16) need 8 issuing warps to reach the gather ceiling,
// 4 KiB chunks are saturated by 4 (profiles/r1_gather_read_bench.log).
template <int B, bool FP8, int NT, bool FLAT>
int launch_stream_t(StreamParams& sp, cudaStream_t s) {
  // 8 producer warps leave 128 registers per thread: enough for one head tile (NT = 1), spills with two
  if constexpr (FP8 && NT == 1 && !FLAT) {
    // v2 partitions in the reference's normalised format: per-warp items (B200_STREAM_PW=-1 switches it off)
    if (sp.so.normalized && sp.so.o != nullptr && reinterpret_cast<uintptr_t>(sp.so.o) % 16 == 0 && env_int("B200_STREAM_PW") >= 0)
      return launch_stream_np<B, FP8, NT, FLAT, 8, true>(sp, s);
  }
  if constexpr (FP8 && NT == 1) {
    int np = B == 16 ? 8 : 4;
    if (const int want = env_int("B200_STREAM_NPROD")) np = want >= 8 ? 8 : 4;
    if (np == 8) return launch_stream_np<B, FP8, NT, FLAT, 8>(sp, s);
  }
  return launch_stream_np<B, FP8, NT, FLAT, 4>(sp, s);
}

}  // namespace

bool stream_variant_eligible(const b200_paged_attn_params_t& p, const char** why) {
  const char* w = nullptr;
  const int G = p.num_kv_heads > 0 ? p.num_heads / p.num_kv_heads : 0;
  const int eb = p.kv_dtype == B200_KV_FP8_E4M3 ? 1 : 2;
  if (p.dtype != B200_DTYPE_BF16) w = "query dtype is not bf16";
  else if (p.head_size != kD) w = "head_size is not 128";
  else if (p.block_size != 16 && p.block_size != 32) w = "block_size is not 16 or 32";
  else if (G < 1 || G > kFlatMaxGroup) w = "GQA group size above 16";
  else if ((uint64_t)p.num_seqs * p.num_kv_heads >= kFlatCounterSlots) w = "more (seq, kv head) pairs than arrival counters";
  else if ((p.kv_block_stride * eb) % 16 != 0 || (p.kv_head_stride * eb) % 16 != 0) w = "KV strides not 16-byte multiples";
  else if ((reinterpret_cast<uintptr_t>(p.key_cache) | reinterpret_cast<uintptr_t>(p.value_cache)) % 16 != 0) w = "KV cache not 16-byte aligned";
  else if (p.q_stride % 8 != 0 || reinterpret_cast<uintptr_t>(p.query) % 16 != 0) w = "query rows not 16-byte aligned";
  else if (reinterpret_cast<uintptr_t>(p.out) % 16 != 0) w = "out not 16-byte aligned";
  if (why) *why = w;
  return w == nullptr;
}

namespace {
template <bool FLAT>
int dispatch_stream(const b200_paged_attn_params_t& p, StreamParams& sp, cudaStream_t s) {
  const bool fp8 = p.kv_dtype == B200_KV_FP8_E4M3;
  const int nt = (sp.G + 7) / 8;
#define GO(B_, FP8_) \
  (nt == 1 ? launch_stream_t<B_, FP8_, 1, FLAT>(sp, s) : launch_stream_t<B_, FP8_, 2, FLAT>(sp, s))
  if (p.block_size == 16) return fp8 ? GO(16, true) : GO(16, false);
  return fp8 ? GO(32, true) : GO(32, false);
#undef GO
}
}  // namespace

// ---- flash layout (paged_attention_flash) -------------------------------------------------------------------
bool stream_flash_eligible(const b200_paged_attn_params_t& p, const char** why) {
  const char* w = nullptr;
  const int G = p.num_kv_heads > 0 ? p.num_heads / p.num_kv_heads : 0;
  if (p.dtype != B200_DTYPE_BF16) w = "query dtype is not bf16";
  else if (p.kv_dtype != B200_KV_AUTO) w = "flash-layout decode takes a bf16 cache (the reference's flash-attention path has no fp8 cache)";
  else if (p.head_size != kD) w = "head_size is not 128";
  else if (p.block_size < 16 || p.block_size % 16 != 0) w = "block_size is not a multiple of 16";
  else if (G < 1 || G > kFlatMaxGroup) w = "GQA group size above 16";
  else if ((uint64_t)p.num_seqs * p.num_kv_heads >= kFlatCounterSlots) w = "more (seq, kv head) pairs than arrival counters";
  else if ((p.kv_block_stride * 2) % 16 != 0 || (p.kv_head_stride * 2) % 16 != 0 || (p.kv_page_stride * 2) % 16 != 0)
    w = "KV strides not 16-byte multiples";
  else if ((reinterpret_cast<uintptr_t>(p.key_cache) | reinterpret_cast<uintptr_t>(p.value_cache)) % 16 != 0) w = "KV cache not 16-byte aligned";
  else if (p.q_stride % 8 != 0 || reinterpret_cast<uintptr_t>(p.query) % 16 != 0) w = "query rows not 16-byte aligned";
  else if (reinterpret_cast<uintptr_t>(p.out) % 16 != 0) w = "out not 16-byte aligned";
  if (why) *why = w;
  return w == nullptr;
}

namespace {
// cuTensorMapEncodeTiled through the runtime's driver entry point lookup (no link-time dependency on libcuda).
// Pure host computation: nothing is enqueued, allocated or read from the device.
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = [] {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess) {
      cudaGetLastError();
      f = nullptr;
    }
    return reinterpret_cast<EncodeTiledFn>(f);
  }();
  return fn;
}

// 4-D map over (dim, kv head, token in block, block); box = 64 dims x 1 head x 16 tokens x 1 block, 128-byte swizzle.
int encode_flash_map(TmaDesc* out, const void* cache, const b200_paged_attn_params_t& p) {
  EncodeTiledFn enc = encode_tiled_fn();
  B200_REQUIRE(enc != nullptr, B200_ERR_CUDA, "paged_attention_flash: cuTensorMapEncodeTiled is not available from the driver");
  const cuuint64_t nb = p.kv_num_blocks > 0 ? (cuuint64_t)p.kv_num_blocks : (cuuint64_t)1 << 24;
  const cuuint64_t dims[4] = {(cuuint64_t)kD, (cuuint64_t)p.num_kv_heads, (cuuint64_t)p.block_size, nb};
  const cuuint64_t strides[3] = {(cuuint64_t)p.kv_head_stride * 2, (cuuint64_t)p.kv_page_stride * 2,
                                 (cuuint64_t)p.kv_block_stride * 2};
  const cuuint32_t box[4] = {64, 1, 16, 1};
  const cuuint32_t estr[4] = {1, 1, 1, 1};
  CUtensorMap m;
  const CUresult r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(cache), dims, strides, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  B200_REQUIRE(r == CUDA_SUCCESS, B200_ERR_INVALID_ARGUMENT,
               "paged_attention_flash: cuTensorMapEncodeTiled failed (%d) for strides %lld/%lld/%lld elements", (int)r,
               (long long)p.kv_block_stride, (long long)p.kv_page_stride, (long long)p.kv_head_stride);
  static_assert(sizeof(CUtensorMap) == sizeof(TmaDesc), "CUtensorMap is 128 bytes");
  memcpy(out, &m, sizeof(m));
  return B200_OK;
}

template <int NT>
int launch_stream_flash_t(StreamParams& sp, cudaStream_t s) {
  constexpr int NP = 4;
  // ring depth: a multiple of the consumer-warp count; 8 stages = 256 tokens = 128 KiB of rows in flight
  int nst = kConsumerWarps;
  auto smem_of = [](int n) {
    return 1024 + (size_t)n * kFlashStageBytes + (size_t)kConsumerWarps * 8 * NT * (kMergeStride + 2) * sizeof(float) +
           (size_t)n * 16 + sizeof(Sched) + 16;
  };
  if (const int e = env_int("B200_STREAM_NSTAGES")) {
    const int want = e / kConsumerWarps * kConsumerWarps;
    if (want > nst && smem_of(want) <= kMaxSmem) nst = want;
  }
  sp.nstages = nst;
  const size_t smem = smem_of(nst);
  auto kern = paged_attention_stream_kernel<16, false, NT, true, NP, true>;
  static thread_local int configured_device = -1;
  int dev = 0;
  cudaGetDevice(&dev);
  if (configured_device != dev) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxSmem);
    if (e != cudaSuccess) {
      set_error("cudaFuncSetAttribute(smem) failed: %s", cudaGetErrorString(e));
      return B200_ERR_CUDA;
    }
    configured_device = dev;
  }
  const int grid = min(min(device_sm_count(), kFlatMaxCtas), sp.total_items);
  kern<<<grid, (kConsumerWarps + NP) * kWarpSize, smem, s>>>(sp);
  count_launch();
  return check_launch("paged_attention_flash");
}
}  // namespace

namespace {
int g_dyn_override = -1;  // b200_set_stream_dynamic(): test hook, overrides the environment knob
// flat schedule: arrival counters, the claim counter of the dynamic region, partial-result slots packed by G
void set_flat_workspace(StreamParams& sp, void* workspace, bool fp8) {
  sp.counters = static_cast<int*>(workspace);
  sp.claim_counter = sp.counters + (kFlatCounterSlots - 1);
  float* slots = reinterpret_cast<float*>(static_cast<char*>(workspace) + kFlatCounterBytes);
  const size_t nslots = (size_t)kFlatMaxRanges * 2 * sp.G;
  sp.ws_o = slots;
  sp.ws_m = slots + nslots * kD;
  sp.ws_l = sp.ws_m + nslots;
  // The dynamic tail region is an EXPERIMENT that measured slower than the static split and is off unless
  // B200_STREAM_DYN=<tiles of the smallest chunk> asks for it (profiles/r2_dyn_bench.log: C2 168 us static, 211 / 197 /
  // 183 us with 4 / 8 / 16-tile minimum chunks): every claimed chunk costs a global atomic, a block-table fetch and a
  // partial-result hand-over (fence + arrival counter) back to back, ~4 us under load, against the ~9 us of tail
  // imbalance it was meant to recover.  Parity of the mechanism is covered by tests/test_paged_attention_gpu.py
  // (b200_set_stream_dynamic).
  (void)fp8;
  const int knob = g_dyn_override >= 0 ? g_dyn_override : env_int("B200_STREAM_DYN");
  sp.dyn_min = knob > 0 ? knob : 0;
}
}  // namespace

void set_stream_dynamic(int min_chunk_tiles) { g_dyn_override = min_chunk_tiles; }

int launch_attention_stream_flash(const b200_paged_attn_params_t& p, void* workspace, cudaStream_t s) {
  const char* why = nullptr;
  B200_REQUIRE(stream_flash_eligible(p, &why), B200_ERR_UNSUPPORTED, "paged_attention_flash: %s", why);
  B200_REQUIRE(workspace != nullptr, B200_ERR_WORKSPACE, "paged_attention_flash: workspace missing");
  StreamParams sp = fill_stream_params(p);
  sp.kv_page_stride = p.kv_page_stride * 2;
  sp.block_size = p.block_size;
  int rc = encode_flash_map(&sp.tm_k, p.key_cache, p);
  if (rc != B200_OK) return rc;
  rc = encode_flash_map(&sp.tm_v, p.value_cache, p);
  if (rc != B200_OK) return rc;
  set_flat_workspace(sp, workspace, p.kv_dtype == B200_KV_FP8_E4M3);
  const int64_t tiles = (int64_t)p.num_seqs * p.num_kv_heads * ((p.max_seq_len + kTileTokens - 1) / kTileTokens);
  sp.total_items = (int)(tiles < (int64_t)kFlatMaxCtas ? tiles : (int64_t)kFlatMaxCtas);
  if (sp.total_items == 0) return B200_OK;
  return (sp.G + 7) / 8 == 1 ? launch_stream_flash_t<1>(sp, s) : launch_stream_flash_t<2>(sp, s);
}

#ifdef B200_TIMELINE
extern "C" int b200_debug_timeline(unsigned long long* host_dst) {
  return cudaMemcpyFromSymbol(host_dst, g_timeline, sizeof(unsigned long long) * 256 * 8) == cudaSuccess ? 0 : 1;
}
#endif

int launch_attention_stream(const b200_paged_attn_params_t& p, const SplitOut& so, cudaStream_t s) {
  const char* why = nullptr;
  B200_REQUIRE(stream_variant_eligible(p, &why), B200_ERR_UNSUPPORTED, "stream kernel not eligible: %s", why);
  B200_REQUIRE(so.chunk > 0 && so.chunk % kTileTokens == 0, B200_ERR_INVALID_ARGUMENT,
               "stream kernel: chunk %d is not a multiple of %d", so.chunk, kTileTokens);
  StreamParams sp = fill_stream_params(p);
  sp.so = so;
  sp.total_items = so.num_partitions * p.num_seqs * p.num_kv_heads;
  if (sp.total_items == 0) return B200_OK;
  return dispatch_stream<false>(p, sp, s);
}

int launch_attention_stream_flat(const b200_paged_attn_params_t& p, void* workspace, cudaStream_t s) {
  const char* why = nullptr;
  B200_REQUIRE(stream_variant_eligible(p, &why), B200_ERR_UNSUPPORTED, "stream kernel not eligible: %s", why);
  B200_REQUIRE(workspace != nullptr, B200_ERR_WORKSPACE, "stream kernel: workspace missing");
  StreamParams sp = fill_stream_params(p);
  set_flat_workspace(sp, workspace, p.kv_dtype == B200_KV_FP8_E4M3);
  // upper bound of the tile count (the kernel derives the real one from seq_lens on the device)
  const int64_t tiles = (int64_t)p.num_seqs * p.num_kv_heads * ((p.max_seq_len + kTileTokens - 1) / kTileTokens);
  sp.total_items = (int)(tiles < (int64_t)kFlatMaxCtas ? tiles : (int64_t)kFlatMaxCtas);
  if (sp.total_items == 0) return B200_OK;
  return dispatch_stream<true>(p, sp, s);
}

}  // namespace b200
