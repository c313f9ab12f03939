// Streaming paged-attention decode kernel for B200 (sm_100a): head_size 128,
// block_size 16/32, bf16 queries, bf16 or fp8-e4m3 KV cache, any GQA group
// size up to 16.  Serves paged_attention_v1 and paged_attention_v2
// (reference: csrc/attention/attention_kernels.cuh:79-485).
//
// Design:
//  * persistent grid, one CTA per SM; work item = (partition, seq, kv head);
//    the CTA serves ALL query heads of the kv head, so every KV byte is read
//    from HBM once (the reference re-reads it once per query head);
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
#include <float.h>
#include <stdio.h>
#include <stdlib.h>

#include "common.cuh"

namespace b200 {
namespace {

constexpr int kD = 128;             // head size served by this kernel
constexpr int kTileTokens = 32;     // tokens per ring stage
constexpr int kConsumerWarps = 8;
constexpr int kProducerWarps = 4;
constexpr int kThreads = (kConsumerWarps + kProducerWarps) * kWarpSize;
constexpr int kMergeStride = kD + 4;  // fp32 words per (warp, head) row
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;

struct StreamParams {
  const __nv_bfloat16* q;
  const uint8_t* kc;
  const uint8_t* vc;
  const int32_t* bt;
  const int32_t* seq_lens;
  const float* alibi;
  const float* k_scale;
  const float* v_scale;
  __nv_bfloat16* out;
  SplitOut so;
  int64_t q_stride;         // elements
  int64_t kv_block_stride;  // bytes
  int64_t kv_head_stride;   // bytes
  int S, H, Hkv, G, max_blocks, nstages, total_items;
  float scale;
  int debug_num_blocks;  // B200_DEBUG builds only: bound for block-table entries (0 = unchecked)
};

// ---- PTX wrappers ------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
#ifdef B200_DEBUG
__device__ __forceinline__ bool mbar_try(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}
// debug build: a wait that gives up after ~2 s, reports who was waiting for what, and traps
__device__ __noinline__ void mbar_wait_dbg(uint32_t bar, uint32_t parity, int tag, int a, int b) {
  const long long t0 = clock64();
  while (!mbar_try(bar, parity)) {
    if (clock64() - t0 > 4000000000LL) {
      printf("[b200 debug] mbarrier timeout: cta %d warp %d lane %d tag %d a %d b %d bar 0x%x parity %u\n",
             (int)blockIdx.x, (int)(threadIdx.x >> 5), (int)(threadIdx.x & 31), tag, a, b, bar, parity);
      __trap();
    }
  }
}
#define MBAR_WAIT(bar, parity, tag, a, b) mbar_wait_dbg(bar, parity, tag, a, b)
#else
#define MBAR_WAIT(bar, parity, tag, a, b) mbar_wait(bar, parity)
#endif
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(bar),
      "r"(parity)
      : "memory");
}
// 1-D TMA bulk copy global -> shared, completion counted on an mbarrier.
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
      "l"(src), "r"(bytes), "r"(bar)
      : "memory");
}
__device__ __forceinline__ void ldmatrix_x4(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
__device__ __forceinline__ uint32_t movmatrix_trans(uint32_t a) {
  uint32_t d;
  asm volatile("movmatrix.sync.aligned.m8n8.trans.b16 %0, %1;" : "=r"(d) : "r"(a));
  return d;
}
template <bool F16>
__device__ __forceinline__ void mma16816(float (&d)[4], const uint32_t (&a)[4], const uint32_t (&b)[2]) {
  if constexpr (F16) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
        "{%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
  } else {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
        "{%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
  }
}
// two e4m3 bytes -> f16x2 (exact)
__device__ __forceinline__ uint32_t cvt_e4m3x2_f16x2(uint16_t v) {
  uint32_t d;
  asm("cvt.rn.f16x2.e4m3x2 %0, %1;" : "=r"(d) : "h"(v));
  return d;
}
template <bool F16>
__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
  uint32_t d;
  if constexpr (F16) {
    asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
  } else {
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
  }
  return d;
}
__device__ __forceinline__ float fast_exp2(float x) {  // ex2.approx: 2 ulp, exp2(-inf) = 0
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ void consumer_bar_sync() {
  asm volatile("bar.sync 1, %0;" ::"n"(kConsumerWarps * kWarpSize) : "memory");
}

// ---- work items ----------------------------------------------------------------
struct Item {
  int s, kvh, part, t0, t1, L;
};
__device__ __forceinline__ bool decode_item(const StreamParams& p, int item, Item& it) {
  const int SH = p.S * p.Hkv;
  it.part = item / SH;
  const int r = item - it.part * SH;
  it.s = r / p.Hkv;
  it.kvh = r - it.s * p.Hkv;
  it.L = __ldg(p.seq_lens + it.s);
  it.t0 = it.part * p.so.chunk;
  if (it.t0 >= it.L) return false;
  it.t1 = min(it.t0 + p.so.chunk, it.L);
  return true;
}

template <int B, bool FP8>
struct Cfg {
  static constexpr int kElemBytes = FP8 ? 1 : 2;
  static constexpr int kBlkBytes = kD * B * kElemBytes;  // one K (or V) chunk of a block
  static constexpr int kBlocksPerTile = kTileTokens / B;
  static constexpr int kStageBytes = 2 * kBlocksPerTile * kBlkBytes;
  static constexpr int kVOffset = kBlocksPerTile * kBlkBytes;  // V chunks follow the K chunks
};

// ---- the kernel ------------------------------------------------------------------
template <int B, bool FP8, int NT>
__global__ void __launch_bounds__(kThreads, 1) paged_attention_stream_kernel(const StreamParams p) {
  using C = Cfg<B, FP8>;
  constexpr bool F16 = FP8;  // MMA operand type: f16 for fp8 caches, bf16 otherwise
  constexpr int NH = 8 * NT;  // padded heads per kv head
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~uintptr_t(127));
  const int nst = p.nstages;
  uint8_t* ring = smem;
  float* merge_o = reinterpret_cast<float*>(ring + (size_t)nst * C::kStageBytes);
  float* merge_m = merge_o + kConsumerWarps * NH * kMergeStride;
  float* merge_l = merge_m + kConsumerWarps * NH;
  uint64_t* bars = reinterpret_cast<uint64_t*>(merge_l + kConsumerWarps * NH);
  int* merge_flag = reinterpret_cast<int*>(bars + 2 * nst);
  const uint32_t ring_u32 = smem_u32(ring);
  const uint32_t full_u32 = smem_u32(bars);
  const uint32_t empty_u32 = full_u32 + 8u * nst;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) {
    for (int i = 0; i < nst; ++i) {
      mbar_init(full_u32 + 8u * i, 1);
      mbar_init(empty_u32 + 8u * i, 1);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncthreads();

  if (warp >= kConsumerWarps) {
    // ===================== producer warps ========================================
    // Producer warp pw owns the tiles gt == pw (mod kProducerWarps) of the CTA's tile stream; one
    // lane walks the block table and issues the bulk copies (a single issuing thread costs ~1000
    // cycles per tile, so four of them share the stream).
    if (lane != 0) return;
    const int pw = warp - kConsumerWarps;
    uint32_t g_next = (uint32_t)pw;   // next tile of the stream this warp issues
    int stage = pw;                   // g_next % nst (nst >= kProducerWarps)
    uint32_t phase = 0;               // (g_next / nst) & 1
    uint32_t tile_base = 0;
    for (int item = blockIdx.x; item < p.total_items; item += gridDim.x) {
      Item it;
      if (!decode_item(p, item, it)) continue;
      const int nblk = (it.t1 - it.t0 + B - 1) / B;
      const uint32_t ntiles = (uint32_t)((it.t1 - it.t0 + kTileTokens - 1) / kTileTokens);
      const int32_t* bt = p.bt + (int64_t)it.s * p.max_blocks + it.t0 / B;
      const uint8_t* kbase = p.kc + (int64_t)it.kvh * p.kv_head_stride;
      const uint8_t* vbase = p.vc + (int64_t)it.kvh * p.kv_head_stride;
      const uint32_t tile_end = tile_base + ntiles;
      // block ids of the next tile are fetched one step ahead
      int nid0 = 0, nid1 = 0;
      if (g_next < tile_end) {
        const int b0 = (int)(g_next - tile_base) * C::kBlocksPerTile;
        nid0 = __ldg(bt + b0);
        if (C::kBlocksPerTile == 2 && b0 + 1 < nblk) nid1 = __ldg(bt + b0 + 1);
      }
      while (g_next < tile_end) {
        const int b0 = (int)(g_next - tile_base) * C::kBlocksPerTile;
        const int id0 = nid0, id1 = nid1;
        const int nvalid = min(C::kBlocksPerTile, nblk - b0);
        const int bn = b0 + kProducerWarps * C::kBlocksPerTile;
        if (bn < nblk) {
          nid0 = __ldg(bt + bn);
          if (C::kBlocksPerTile == 2 && bn + 1 < nblk) nid1 = __ldg(bt + bn + 1);
        }
#ifdef B200_DEBUG
        if (p.debug_num_blocks > 0 && ((unsigned)id0 >= (unsigned)p.debug_num_blocks ||
                                       (nvalid == 2 && (unsigned)id1 >= (unsigned)p.debug_num_blocks))) {
          printf("[b200 debug] bad block id: cta %d item %d s %d part %d b0 %d id0 %d id1 %d nblk %d\n",
                 (int)blockIdx.x, item, it.s, it.part, b0, id0, id1, nblk);
          __trap();
        }
#endif
        MBAR_WAIT(empty_u32 + 8u * stage, phase ^ 1u, 1, item, stage);
        const uint32_t fb = full_u32 + 8u * stage;
        mbar_expect_tx(fb, (uint32_t)(nvalid * 2 * C::kBlkBytes));
        const uint32_t dst = ring_u32 + (uint32_t)stage * C::kStageBytes;
        bulk_g2s(dst, kbase + (int64_t)id0 * p.kv_block_stride, C::kBlkBytes, fb);
        bulk_g2s(dst + C::kVOffset, vbase + (int64_t)id0 * p.kv_block_stride, C::kBlkBytes, fb);
        if (C::kBlocksPerTile == 2 && nvalid == 2) {
          bulk_g2s(dst + C::kBlkBytes, kbase + (int64_t)id1 * p.kv_block_stride, C::kBlkBytes, fb);
          bulk_g2s(dst + C::kVOffset + C::kBlkBytes, vbase + (int64_t)id1 * p.kv_block_stride, C::kBlkBytes, fb);
        }
        g_next += kProducerWarps;
        stage += kProducerWarps;
        if (stage >= nst) {
          stage -= nst;
          phase ^= 1u;
        }
      }
      tile_base = tile_end;
    }
    return;
  }

  // ===================== consumer warps ============================================
  const int g = lane >> 2, c = lane & 3;
  // ldmatrix lane addressing (byte offsets inside a K / V chunk)
  const int mi = lane >> 3, r8 = lane & 7;
  uint32_t k_lane_off, v_lane_off;
  if constexpr (!FP8) {
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

  uint32_t tile_base = 0;  // tiles issued to the ring before this item
  for (int item = blockIdx.x; item < p.total_items; item += gridDim.x) {
    Item it;
    if (!decode_item(p, item, it)) continue;
    const int ntiles = (it.t1 - it.t0 + kTileTokens - 1) / kTileTokens;

    // ---- Q fragments (B operand: k = dims, n = heads) ------------------------------
    uint32_t qf[NT][8][2];
    float sc2[NT][2];   // logit multiplier (log2 units) of heads 2c, 2c+1
    float al2[NT][2];   // alibi slope * log2(e)
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      const int hl = nt * 8 + g;  // head inside the group owned by this lane's B column
      const bool hv = hl < p.G;
      const __nv_bfloat16* qrow = p.q + (int64_t)it.s * p.q_stride + (int64_t)(it.kvh * p.G + hl) * kD;
      float qscale_inv = 1.f;
      if constexpr (!FP8) {
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
          qf[nt][ks][0] = hv ? __ldg(reinterpret_cast<const uint32_t*>(qrow + 16 * ks + 2 * c)) : 0u;
          qf[nt][ks][1] = hv ? __ldg(reinterpret_cast<const uint32_t*>(qrow + 16 * ks + 8 + 2 * c)) : 0u;
        }
      } else {
        // bf16 -> f16 with a per-head power-of-two prescale so that no value
        // overflows f16; the prescale is undone in the logit multiplier.
        float qv[8][4];
        float amax = 0.f;
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
          uint2 raw = hv ? __ldg(reinterpret_cast<const uint2*>(qrow + 16 * ks + 4 * c)) : make_uint2(0u, 0u);
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
        al2[nt][e] = (p.alibi != nullptr && hc < p.G) ? __ldg(p.alibi + it.kvh * p.G + hc) * kLog2e : 0.f;
      }
    }
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
    for (int ti = (int)((warp - tile_base) & (kConsumerWarps - 1)); ti < ntiles; ti += kConsumerWarps) {
      const uint32_t gt = tile_base + (uint32_t)ti;
      const uint32_t stage = gt % (uint32_t)nst;
      const uint32_t parity = (gt / (uint32_t)nst) & 1u;
      MBAR_WAIT(full_u32 + 8u * stage, parity, 2, item, (int)gt);
      const uint32_t sb = ring_u32 + stage * C::kStageBytes;
      const int tok0 = it.t0 + ti * kTileTokens;   // first token of the tile
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
        const uint32_t kb = sb + (B == 16 ? j * C::kBlkBytes : j * 16 * 16) + k_lane_off;
        if constexpr (!FP8) {
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
            uint32_t a0[4] = {cvt_e4m3x2_f16x2((uint16_t)(r[0] & 0xffffu)), cvt_e4m3x2_f16x2((uint16_t)(r[1] & 0xffffu)),
                              cvt_e4m3x2_f16x2((uint16_t)(r[0] >> 16)), cvt_e4m3x2_f16x2((uint16_t)(r[1] >> 16))};
            uint32_t a1[4] = {cvt_e4m3x2_f16x2((uint16_t)(r[2] & 0xffffu)), cvt_e4m3x2_f16x2((uint16_t)(r[3] & 0xffffu)),
                              cvt_e4m3x2_f16x2((uint16_t)(r[2] >> 16)), cvt_e4m3x2_f16x2((uint16_t)(r[3] >> 16))};
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
      uint32_t vmask[2][2];
      if (partial) {
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const int tl = 16 * j + vtok_lo, th = 16 * j + vtok_hi;
          vmask[j][0] = (tl < nvalid ? 0x0000ffffu : 0u) | (tl + 1 < nvalid ? 0xffff0000u : 0u);
          vmask[j][1] = (th < nvalid ? 0x0000ffffu : 0u) | (th + 1 < nvalid ? 0xffff0000u : 0u);
        }
      }
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const uint32_t vb = sb + C::kVOffset + (B == 16 ? j * C::kBlkBytes : j * 16 * C::kElemBytes) + v_lane_off;
        if constexpr (!FP8) {
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            uint32_t a[4];
            ldmatrix_x4(a, vb + i * 16 * B * 2);
            if (partial) {  // stale cache past seq_len may hold NaN/Inf: zero it
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
            uint32_t a0[4] = {cvt_e4m3x2_f16x2((uint16_t)(r[0] & 0xffffu)), cvt_e4m3x2_f16x2((uint16_t)(r[1] & 0xffffu)),
                              cvt_e4m3x2_f16x2((uint16_t)(r[0] >> 16)), cvt_e4m3x2_f16x2((uint16_t)(r[1] >> 16))};
            uint32_t a1[4] = {cvt_e4m3x2_f16x2((uint16_t)(r[2] & 0xffffu)), cvt_e4m3x2_f16x2((uint16_t)(r[3] & 0xffffu)),
                              cvt_e4m3x2_f16x2((uint16_t)(r[2] >> 16)), cvt_e4m3x2_f16x2((uint16_t)(r[3] >> 16))};
            if (partial) {
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
      // all lanes have consumed the stage: hand it back to the producer
      __syncwarp();
      if (lane == 0) mbar_arrive(empty_u32 + 8u * stage);
    }
    tile_base += (uint32_t)ntiles;

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
    consumer_bar_sync();
    {
      const bool single = it.L <= p.so.chunk;
      const bool want_partials = p.so.o != nullptr && (!single || p.so.normalized);
      const int d = tid & (kD - 1);
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
        if (single) p.out[((int64_t)it.s * p.H + head) * kD + d] = __float2bfloat16_rn(on);
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
    }
    if (p.so.counters != nullptr && it.L > p.so.chunk) {
      // ---- fused split-KV reduce: the CTA that delivers the last partition of (seq, kv head) merges them
      // (same arithmetic as split_reduce_kernel / reference attention_kernels.cuh:594-657) --------------
      __threadfence();       // this thread's partials are visible device-wide
      consumer_bar_sync();   // ... and so are everyone's
      if (tid == 0) {
        const int nparts = (it.L + p.so.chunk - 1) / p.so.chunk;
        int* cnt = p.so.counters + it.s * p.Hkv + it.kvh;
        const int old = atomicAdd(cnt, 1);
        const int last = old == nparts - 1;
        if (last) *cnt = 0;  // leave the counters zeroed for the next launch
        *merge_flag = last;
      }
      consumer_bar_sync();
      if (*merge_flag) {
        __threadfence();
        const int nparts = (it.L + p.so.chunk - 1) / p.so.chunk;
        const int d = tid & (kD - 1);
        for (int h = tid >> 7; h < p.G; h += (kConsumerWarps * kWarpSize) >> 7) {
          const int head = it.kvh * p.G + h;
          const int64_t base = ((int64_t)it.s * p.H + head) * p.so.num_partitions;
          float M = -INFINITY;
          for (int j = 0; j < nparts; ++j) M = fmaxf(M, __ldcg(p.so.m + base + j));
          float lsum = 0.f, o = 0.f;
          for (int j = 0; j < nparts; ++j) {
            const float wgt = __expf(__ldcg(p.so.m + base + j) - M);
            lsum = fmaf(__ldcg(p.so.l + base + j), wgt, lsum);
            o = fmaf(__ldcg(reinterpret_cast<const float*>(p.so.o) + (base + j) * kD + d), wgt, o);
          }
          p.out[((int64_t)it.s * p.H + head) * kD + d] = __float2bfloat16_rn(o * __fdividef(1.f, lsum + 1e-6f));
        }
      }
    }
    consumer_bar_sync();  // merge buffers are free again
  }
}

template <int B, bool FP8, int NT>
size_t stream_smem_bytes(int nstages) {
  using C = Cfg<B, FP8>;
  return 128 + (size_t)nstages * C::kStageBytes +
         (size_t)kConsumerWarps * 8 * NT * (kMergeStride + 2) * sizeof(float) + (size_t)nstages * 16 + 16;
}

constexpr size_t kMaxSmem = 227 * 1024;

template <int B, bool FP8, int NT>
int launch_stream_t(StreamParams& sp, cudaStream_t s) {
  using C = Cfg<B, FP8>;
  // deepest ring that fits (capped: beyond ~190 KiB in flight nothing is gained)
  // (a multiple of the consumer-warp count: see the tile -> warp mapping in the kernel)
  int nst = kConsumerWarps;
  while (stream_smem_bytes<B, FP8, NT>(nst + kConsumerWarps) <= kMaxSmem &&
         (size_t)(nst + kConsumerWarps) * C::kStageBytes <= 192 * 1024)
    nst += kConsumerWarps;
  if (const char* e = getenv("B200_STREAM_NSTAGES"))
    nst = max(kConsumerWarps, min(nst, atoi(e) / kConsumerWarps * kConsumerWarps));
  sp.nstages = nst;
  const size_t smem = stream_smem_bytes<B, FP8, NT>(nst);
  auto kern = paged_attention_stream_kernel<B, FP8, NT>;
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
  const int grid = min(device_sm_count(), sp.total_items);
  kern<<<grid, kThreads, smem, s>>>(sp);
  count_launch();
  return check_launch("paged_attention_stream");
}

}  // namespace

bool stream_variant_eligible(const b200_paged_attn_params_t& p, const char** why) {
  const char* w = nullptr;
  const int G = p.num_kv_heads > 0 ? p.num_heads / p.num_kv_heads : 0;
  const int eb = p.kv_dtype == B200_KV_FP8_E4M3 ? 1 : 2;
  if (p.dtype != B200_DTYPE_BF16) w = "query dtype is not bf16";
  else if (p.head_size != kD) w = "head_size is not 128";
  else if (p.block_size != 16 && p.block_size != 32) w = "block_size is not 16 or 32";
  else if (G < 1 || G > 16) w = "GQA group size above 16";
  else if ((p.kv_block_stride * eb) % 16 != 0 || (p.kv_head_stride * eb) % 16 != 0) w = "KV strides not 16-byte multiples";
  else if ((reinterpret_cast<uintptr_t>(p.key_cache) | reinterpret_cast<uintptr_t>(p.value_cache)) % 16 != 0) w = "KV cache not 16-byte aligned";
  else if (p.q_stride % 8 != 0 || reinterpret_cast<uintptr_t>(p.query) % 16 != 0) w = "query rows not 16-byte aligned";
  if (why) *why = w;
  return w == nullptr;
}

int launch_attention_stream(const b200_paged_attn_params_t& p, const SplitOut& so, cudaStream_t s) {
  const char* why = nullptr;
  B200_REQUIRE(stream_variant_eligible(p, &why), B200_ERR_UNSUPPORTED, "stream kernel not eligible: %s", why);
  B200_REQUIRE(so.chunk > 0 && so.chunk % kTileTokens == 0, B200_ERR_INVALID_ARGUMENT,
               "stream kernel: chunk %d is not a multiple of %d", so.chunk, kTileTokens);
  const bool fp8 = p.kv_dtype == B200_KV_FP8_E4M3;
  const int eb = fp8 ? 1 : 2;
  StreamParams sp;
  sp.q = static_cast<const __nv_bfloat16*>(p.query);
  sp.kc = static_cast<const uint8_t*>(p.key_cache);
  sp.vc = static_cast<const uint8_t*>(p.value_cache);
  sp.bt = p.block_tables; sp.seq_lens = p.seq_lens; sp.alibi = p.alibi_slopes;
  sp.k_scale = p.k_scale; sp.v_scale = p.v_scale;
  sp.out = static_cast<__nv_bfloat16*>(p.out);
  sp.so = so;
  sp.q_stride = p.q_stride;
  sp.kv_block_stride = p.kv_block_stride * eb;
  sp.kv_head_stride = p.kv_head_stride * eb;
  sp.S = p.num_seqs; sp.H = p.num_heads; sp.Hkv = p.num_kv_heads; sp.G = p.num_heads / p.num_kv_heads;
  sp.max_blocks = p.max_num_blocks_per_seq;
  sp.nstages = 0;
  sp.total_items = so.num_partitions * p.num_seqs * p.num_kv_heads;
  sp.scale = p.scale;
  sp.debug_num_blocks = 0;
  if (const char* e = getenv("B200_DEBUG_NUM_BLOCKS")) sp.debug_num_blocks = atoi(e);
  if (sp.total_items == 0) return B200_OK;
  const int nt = (sp.G + 7) / 8;
#define GO(B_, FP8_)                                                     \
  (nt == 1 ? launch_stream_t<B_, FP8_, 1>(sp, s) : launch_stream_t<B_, FP8_, 2>(sp, s))
  if (p.block_size == 16) return fp8 ? GO(16, true) : GO(16, false);
  return fp8 ? GO(32, true) : GO(32, false);
#undef GO
}

}  // namespace b200
