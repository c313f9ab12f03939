// tcgen05 paged-attention decode kernel for FP8-e4m3 KV caches (B200, sm_100a): head_size 128, block_size 16,
// bf16 queries, GQA group <= 8.  Serves paged_attention_v1 (flat schedule) and paged_attention_v2 (512-token items).
// Reference semantics: csrc/attention/attention_kernels.cuh:75-485 with the fp8 dequantisation of its call sites
// (:274-277, :403-407); SURVEY.md §3.3-3.4, §8(a) a2/a7.
//
// Why a second streaming kernel: with an fp8 cache the mma.sync kernel (attention_stream.cu) spends its issue slots
// expanding e4m3 to f16 on the CUDA cores (132 F2FP per 32-token tile and warp; profiles/r1_c3_fp8_ncu_summary.txt)
// and mma.sync ...e4m3 is expanded by ptxas to the same F2FP + HMMA.  The 5th-generation tensor core reads e4m3
// operands straight from shared memory, and the paged cache layouts ARE canonical no-swizzle K-major UMMA operands:
//   K chunk [D/16][16 tok][16 B]  -> B operand of S = Q.K^T: core matrix = 8 tokens x 16 dims, LBO 256 B, SBO 128 B
//   V chunk [D][16 tok]           -> B operand of O = P.V  : core matrix = 8 dims x 16 tokens, LBO = next block, SBO 128 B
// (verified bit-exact on a B200 by tools/umma_probe.cu).  So the 1-D TMA bulk copies of attention_stream.cu are kept
// and the tensor core consumes the ring in place; the CUDA cores only see G x 32 logits per tile.
//
// Operand precision: the bf16 query and the fp32 probabilities are split into TWO e4m3 terms, x ~= hi + lo/16
// (hi = rn_e4m3(x), lo = rn_e4m3((x - hi) * 16)), stacked along M: accumulator row h holds the hi product, row 8+h
// the lo product, recombined in fp32.  Relative error <= 2^-9 per element (bf16 itself is 2^-9), K and V are exact.
//
// Warp roles (448 threads, one CTA per SM, persistent):
//   warps 0-3   "class" warps.  M = 64 accumulators occupy TMEM lanes 32q..32q+15 (q = 0..3); the Q operand is
//               replicated in all four 16-row groups, so every quadrant q sees the full S tile and class warp q owns
//               the tiles gt == q (mod 4): softmax with its own running reference, P written to its own P buffer
//               (rows of the other quadrants stay zero, so P.V adds nothing to their accumulator lanes).
//   warp 4      Q.K^T issuer (one elected lane): S = Q.K^T per block (8 x M64 N16 K32 per tile)
//   warp 5      P.V issuer: O += P.V (M64 N128 K32 per tile); tcgen05.commit hands ring stages back to the producers
//   warps 6-13  producers: block-table walk + 1-D TMA bulk copies into the ring (stream_common.cuh).
// The four class partials of a (seq, kv head) segment are merged through shared memory at its end (exact: each class
// keeps its true running max), then finished exactly like attention_stream.cu (flat arrival counters / v2 partials).
#include "stream_common.cuh"

namespace b200 {
namespace {

constexpr int kUB = 16;                // block size served
constexpr int kClassWarps = 4;
constexpr int kQkWarp = kClassWarps;      // warp issuing S = Q.K^T
constexpr int kPvWarp = kClassWarps + 1;  // warp issuing O += P.V
constexpr int kUProducers = 8;
constexpr int kUThreads = (kClassWarps + 2 + kUProducers) * kWarpSize;
using UC = Cfg<kUB, true>;             // ring stage: K chunk 0, K chunk 1, V chunk 0, V chunk 1 (2 KiB each)
constexpr int kQBufBytes = 8 * 1024;   // A operand of Q.K^T: 8 chunks (16 dims) x 64 rows x 16 B
constexpr int kPBufBytes = 2 * 1024;   // A operand of P.V  : 2 chunks (16 tokens) x 64 rows x 16 B
constexpr int kAChunkStride = 1024;    // 64 rows x 16 B
constexpr int kMergeRow = kD + 4;      // fp32 words per merged accumulator row
constexpr uint32_t kTmemCols = 512;
constexpr uint32_t kTmemO = 0;         // 2 x 128 columns: output accumulators, double-buffered by segment parity
constexpr uint32_t kTmemS = 256;       // 4 x 32 columns: logits of one tile per class
constexpr float kPShift = 1.f;         // p' = 2^(s - m_ref + 1): 2 at the reference
constexpr float kTau = 7.f;            // re-reference when a logit exceeds m_ref by 2^7 (p' would pass 256; e4m3 max 448)
constexpr float kLoScale = 16.f;

// barrier slots after full[nst] / empty[nst]
enum : int { kBarSReady = 0, kBarPReady = 4, kBarPFree = 8, kBarQReady = 12, kBarOReady = 14, kBarOFree = 16,
             kBarDrained = 18, kBarRescGo = 19, kBarRescDone = 23, kBarSFree = 27, kNumExtraBars = 31 };

// ---- tcgen05 wrappers -------------------------------------------------------------------------------------
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n.reg .pred p;\nelect.sync _|p, 0xffffffff;\nselp.u32 %0, 1, 0, p;\n}" : "=r"(pred));
  return pred != 0;
}
__host__ __device__ constexpr uint32_t umma_idesc_e4m3(int M, int N) {  // f32 accumulate, e4m3 x e4m3, both K-major
  return (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void umma_f8(uint32_t d_tmem, uint64_t a, uint64_t b, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::f8f6f4 [%0], %1, %2, %3, p;\n}\n" ::"r"(d_tmem),
      "l"(a), "l"(b), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {  // arrives on `bar` when all MMAs issued so far are done
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
// 16 TMEM lanes x 32 columns in the m16n8 accumulator-fragment distribution:
// thread t, register 4j + 2a + b = [lane t/4 + 8a][column 8j + 2(t%4) + b]   (verified by tools/umma_probe.cu)
__device__ __forceinline__ void tmem_ld_frag16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.16x256b.x4.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
      "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]),
      "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]),
      "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
      : "memory");
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}
// two floats -> two e4m3 bytes (lo operand in the low byte), round-to-nearest-even, saturating
__device__ __forceinline__ uint32_t cvt_f32x2_e4m3x2(float lo, float hi) {
  uint16_t d;
  asm("cvt.rn.satfinite.e4m3x2.f32 %0, %1, %2;" : "=h"(d) : "f"(hi), "f"(lo));
  return (uint32_t)d;
}
// two e4m3 bytes -> two floats (exact)
__device__ __forceinline__ float2 cvt_e4m3x2_f32x2(uint32_t v) {
  uint32_t h2;
  asm("cvt.rn.f16x2.e4m3x2 %0, %1;" : "=r"(h2) : "h"((uint16_t)v));
  return __half22float2(*reinterpret_cast<__half2*>(&h2));
}
// x[0..3] -> 4 hi bytes and 4 lo bytes (x ~= hi + lo / 16)
__device__ __forceinline__ void split4_e4m3(const float (&x)[4], uint32_t& hi4, uint32_t& lo4) {
  const uint32_t h01 = cvt_f32x2_e4m3x2(x[0], x[1]), h23 = cvt_f32x2_e4m3x2(x[2], x[3]);
  const float2 f01 = cvt_e4m3x2_f32x2(h01), f23 = cvt_e4m3x2_f32x2(h23);
  const uint32_t l01 = cvt_f32x2_e4m3x2((x[0] - f01.x) * kLoScale, (x[1] - f01.y) * kLoScale);
  const uint32_t l23 = cvt_f32x2_e4m3x2((x[2] - f23.x) * kLoScale, (x[3] - f23.y) * kLoScale);
  hi4 = h01 | (h23 << 16);
  lo4 = l01 | (l23 << 16);
}
__device__ __forceinline__ void class_bar_sync() {
  asm volatile("bar.sync 1, %0;" ::"n"(kClassWarps * kWarpSize) : "memory");
}
__device__ __forceinline__ int ld_volatile_smem(const int* p) {
  int v;
  asm volatile("ld.volatile.shared.s32 %0, [%1];" : "=r"(v) : "r"(smem_u32(p)) : "memory");
  return v;
}
__device__ __forceinline__ void st_volatile_smem(int* p, int v) {
  asm volatile("st.volatile.shared.s32 [%0], %1;" ::"r"(smem_u32(p)), "r"(v) : "memory");
}

struct USmem {
  uint8_t* ring;
  uint8_t* qbuf;       // [2][kQBufBytes]
  uint8_t* pbuf;       // [4][kPBufBytes]
  float* merge_o;      // [4 classes][16 rows][kMergeRow]
  float* merge_mref;   // [4][8]
  float* merge_mtrue;  // [4][8]
  float* merge_l;      // [4][8]
  uint64_t* bars;      // full[nst] empty[nst] extra[kNumExtraBars]
  Sched* sched;
  int* resc_req;       // [4]
  int* merge_flag;
  uint32_t* tmem_slot;
};
__host__ __device__ inline size_t usmem_bytes(int nst) {
  return 128 + (size_t)nst * UC::kStageBytes + 2 * kQBufBytes + 4 * kPBufBytes +
         (size_t)kClassWarps * 16 * kMergeRow * sizeof(float) + 3 * kClassWarps * 8 * sizeof(float) +
         (size_t)(2 * nst + kNumExtraBars) * 8 + sizeof(Sched) + 64;
}
__device__ __forceinline__ USmem carve_smem(uint8_t* raw, int nst) {
  USmem m;
  uint8_t* p = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 127) & ~uintptr_t(127));
  m.ring = p;                                   p += (size_t)nst * UC::kStageBytes;
  m.qbuf = p;                                   p += 2 * kQBufBytes;
  m.pbuf = p;                                   p += 4 * kPBufBytes;
  m.merge_o = reinterpret_cast<float*>(p);      p += (size_t)kClassWarps * 16 * kMergeRow * sizeof(float);
  m.merge_mref = reinterpret_cast<float*>(p);   p += kClassWarps * 8 * sizeof(float);
  m.merge_mtrue = reinterpret_cast<float*>(p);  p += kClassWarps * 8 * sizeof(float);
  m.merge_l = reinterpret_cast<float*>(p);      p += kClassWarps * 8 * sizeof(float);
  m.bars = reinterpret_cast<uint64_t*>(p);      p += (size_t)(2 * nst + kNumExtraBars) * 8;
  m.sched = reinterpret_cast<Sched*>(p);        p += sizeof(Sched);
  m.resc_req = reinterpret_cast<int*>(p);       p += 4 * sizeof(int);
  m.merge_flag = reinterpret_cast<int*>(p);     p += sizeof(int);
  m.tmem_slot = reinterpret_cast<uint32_t*>(p);
  return m;
}

// ---- MMA issuers (whole warp converged, one elected lane issues) ------------------------------------------------
// Two issuing warps, because one warp's dependent chain of uniform-datapath instructions (~230 per tile when a
// single warp did both jobs, ~1200 cycles) was the bottleneck of the whole kernel (profiles/r1_c3_umma_v0_*):
//   Q.K^T warp: waits for a ring stage and for the class's S buffer, issues 4 MMAs per block, commits to s_ready;
//   P.V warp  : waits for the class's P, issues one MMA, commits the ring stage back to the producers.
// A P.V of tile t is only issued after class t&3 saw S(t), i.e. after Q.K^T(t) completed, so the commit that frees
// the stage also covers the other warp's reads of its K chunks.
// Descriptor low words are kept as "base + offset/16" sums: addresses stay below 256 KiB, so adding to the 14-bit
// address field never carries into the LBO field.
__device__ __forceinline__ uint64_t umma_desc_from(uint32_t lo, uint32_t hi) { return ((uint64_t)hi << 32) | lo; }

template <bool FLAT>
__device__ __forceinline__ void qk_warp(const StreamParams& p, const USmem& sm, int nst, uint32_t tmem) {
  const uint32_t full_u32 = smem_u32(sm.bars), xb = full_u32 + 16u * nst;
  constexpr uint32_t idesc_s = umma_idesc_e4m3(64, 16);
  constexpr uint32_t desc_hi = (128u >> 4) | (1u << 14);  // SBO 128 B, descriptor version 1
  const uint32_t q_lo0 = ((smem_u32(sm.qbuf) >> 4) & 0x3fffu) | ((kAChunkStride >> 4) << 16);
  const uint32_t k_lo0 = ((smem_u32(sm.ring) >> 4) & 0x3fffu) | ((256u >> 4) << 16);
  uint32_t gt = 0, stage = 0, full_par = 0, sidx = 0;
  WorkIter<FLAT> wi;
  wi.init(p, sm.sched);
  Seg it;
  while (wi.next(p, it)) {
    const int ntiles = (it.t1 - it.t0 + kTileTokens - 1) / kTileTokens;
    const int last_blocks = min(2, (it.t1 - (it.t0 + (ntiles - 1) * kTileTokens) + kUB - 1) / kUB);
    const uint32_t q_lo = q_lo0 + (sidx & 1u) * (kQBufBytes >> 4);
    MBAR_WAIT(xb + 8u * (kBarQReady + (sidx & 1u)), (sidx >> 1) & 1u, 14, (int)sidx, 0);
    for (int ti = 0; ti < ntiles; ++ti) {
      const uint32_t c = gt & 3u;
      // the class has read S of its previous tile (first use passes: parity trick)
      MBAR_WAIT(xb + 8u * (kBarSFree + c), ((gt >> 2) & 1u) ^ 1u, 16, (int)gt, (int)c);
      MBAR_WAIT(full_u32 + 8u * stage, full_par, 15, (int)gt, (int)stage);
      tc_fence_after();
      const int nblk = ti == ntiles - 1 ? last_blocks : 2;
      const uint32_t k_lo = k_lo0 + stage * (UC::kStageBytes >> 4);
      const uint32_t d_s = tmem + kTmemS + 32u * c;
      if (elect_one()) {
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)
          umma_f8(d_s, umma_desc_from(q_lo + kk * (2 * kAChunkStride >> 4), desc_hi),
                  umma_desc_from(k_lo + kk * (512 >> 4), desc_hi), idesc_s, kk > 0);
        if (nblk == 2) {
#pragma unroll
          for (int kk = 0; kk < 4; ++kk)
            umma_f8(d_s + 16u, umma_desc_from(q_lo + kk * (2 * kAChunkStride >> 4), desc_hi),
                    umma_desc_from(k_lo + (UC::kBlkBytes >> 4) + kk * (512 >> 4), desc_hi), idesc_s, kk > 0);
        }
        umma_commit(xb + 8u * (kBarSReady + c));
      }
      __syncwarp();
      ++gt;
      if (++stage == (uint32_t)nst) {
        stage = 0;
        full_par ^= 1u;
      }
    }
    ++sidx;
  }
}

template <bool FLAT>
__device__ __forceinline__ void pv_warp(const StreamParams& p, const USmem& sm, int nst, uint32_t tmem) {
  const uint32_t full_u32 = smem_u32(sm.bars), empty_u32 = full_u32 + 8u * nst, xb = empty_u32 + 8u * nst;
  constexpr uint32_t idesc_o = umma_idesc_e4m3(64, 128);
  constexpr uint32_t desc_hi = (128u >> 4) | (1u << 14);
  const uint32_t p_lo0 = ((smem_u32(sm.pbuf) >> 4) & 0x3fffu) | ((kAChunkStride >> 4) << 16);
  const uint32_t v_lo0 = (((smem_u32(sm.ring) + UC::kVOffset) >> 4) & 0x3fffu) | ((UC::kBlkBytes >> 4) << 16);
  uint32_t gt = 0, stage = 0, sidx = 0, drained_par = 0, resc_par = 0;
  WorkIter<FLAT> wi;
  wi.init(p, sm.sched);
  Seg it;
  while (wi.next(p, it)) {
    const int ntiles = (it.t1 - it.t0 + kTileTokens - 1) / kTileTokens;
    const uint32_t ob = sidx & 1u;
    const uint32_t d_o = tmem + kTmemO + 128u * ob;
    for (int ti = 0; ti < ntiles; ++ti) {
      const uint32_t c = gt & 3u;
      MBAR_WAIT(xb + 8u * (kBarPReady + c), (gt >> 2) & 1u, 10, (int)gt, (int)c);
      if (ti == 0 && sidx >= 2)  // the class warps have drained this accumulator (two segments ago)
        MBAR_WAIT(xb + 8u * (kBarOFree + ob), ((sidx >> 1) - 1u) & 1u, 11, (int)sidx, 0);
      if (ld_volatile_smem(sm.resc_req + c)) {
        // class c moved its reference: every P.V issued so far must finish, then it rescales its accumulator lanes
        if (elect_one()) umma_commit(xb + 8u * kBarDrained);
        MBAR_WAIT(xb + 8u * kBarDrained, drained_par, 12, (int)c, 0);
        drained_par ^= 1u;
        if (elect_one()) {
          st_volatile_smem(sm.resc_req + c, 0);
          mbar_arrive(xb + 8u * (kBarRescGo + c));
        }
        MBAR_WAIT(xb + 8u * (kBarRescDone + c), (resc_par >> c) & 1u, 13, (int)c, 0);
        resc_par ^= 1u << c;
      }
      tc_fence_after();
      if (elect_one()) {
        umma_f8(d_o, umma_desc_from(p_lo0 + c * (kPBufBytes >> 4), desc_hi),
                umma_desc_from(v_lo0 + stage * (UC::kStageBytes >> 4), desc_hi), idesc_o, ti > 0);
        umma_commit(empty_u32 + 8u * stage);           // ring stage back to the producers
        umma_commit(xb + 8u * (kBarPFree + c));         // P buffer of the class reusable
        if (ti == ntiles - 1) umma_commit(xb + 8u * (kBarOReady + ob));
      }
      __syncwarp();
      ++gt;
      if (++stage == (uint32_t)nst) stage = 0;
    }
    ++sidx;
  }
}

// ---- class (softmax) warps ---------------------------------------------------------------------------------------
template <bool FLAT>
__device__ __forceinline__ void class_warp(const StreamParams& p, const USmem& sm, int nst, uint32_t tmem, int q, int lane) {
  const uint32_t ring_u32 = smem_u32(sm.ring), qbuf_u32 = smem_u32(sm.qbuf);
  const uint32_t pbuf_u32 = smem_u32(sm.pbuf) + (uint32_t)q * kPBufBytes;
  const uint32_t xb = smem_u32(sm.bars) + 16u * nst;
  // S is read with the 16x256b fragment shape: thread t gets, for head h = t / 4, the hi row (TMEM lane h) and the lo
  // row (lane 8 + h) of the 8 tokens 8j + 2(t % 4) + {0, 1}, j = 0..3 — all 32 lanes work, no cross-lane exchange.
  const int h = lane >> 2, c4 = lane & 3;
  const bool row_lane = lane < 16;       // 32x32b accesses (O rows): lanes 0-7 hold hi rows, 8-15 lo rows
  const bool active = h < p.G;
  const uint32_t lane_base = (uint32_t)(32 * q) << 16;
  const float ksc = __ldg(p.k_scale), vsc = __ldg(p.v_scale);
  const int tid = q * kWarpSize + lane;  // 0..127 over the class warps
  constexpr float kInvLo = 1.f / kLoScale;

  // Write the hi / lo e4m3 rows of this quadrant's copy of the Q operand for segment g into qbuf[buf]; returns the
  // logit multiplier (log2 units, k_scale and the per-head power-of-two prescale folded in) of head lane >> 2.
  auto prep_q = [&](const Seg& g, uint32_t buf) -> float {
    const int hh = lane >> 2, pp = lane & 3;  // head, 32-dim part
    const bool hv = hh < p.G;
    const __nv_bfloat16* qrow = p.q + (int64_t)g.s * p.q_stride + (int64_t)(g.kvh * p.G + hh) * kD + 32 * pp;
    float x[32];
    float amax = 0.f;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const uint4 raw = hv ? __ldg(reinterpret_cast<const uint4*>(qrow) + i) : make_uint4(0u, 0u, 0u, 0u);
      const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const __nv_bfloat162 b2 = *reinterpret_cast<const __nv_bfloat162*>(&w[k]);
        x[8 * i + 2 * k] = __low2float(b2);
        x[8 * i + 2 * k + 1] = __high2float(b2);
      }
    }
#pragma unroll
    for (int i = 0; i < 32; ++i) amax = fmaxf(amax, fabsf(x[i]));
    amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 1));
    amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 2));
    float qs = 1.f;
    if (amax > 0.f && amax < INFINITY) {  // amax * qs in [128, 256)
      int ex;
      frexpf(amax, &ex);
      qs = ldexpf(1.f, min(8 - ex, 100));
    }
    const uint32_t row_hi = qbuf_u32 + buf * kQBufBytes + (uint32_t)(16 * q + hh) * 16u;
#pragma unroll
    for (int cc = 0; cc < 2; ++cc) {
      uint32_t hi4[4], lo4[4];
#pragma unroll
      for (int g4 = 0; g4 < 4; ++g4) {
        const float v4[4] = {x[16 * cc + 4 * g4] * qs, x[16 * cc + 4 * g4 + 1] * qs, x[16 * cc + 4 * g4 + 2] * qs,
                             x[16 * cc + 4 * g4 + 3] * qs};
        split4_e4m3(v4, hi4[g4], lo4[g4]);
      }
      const uint32_t a = row_hi + (uint32_t)(2 * pp + cc) * kAChunkStride;
      asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(a), "r"(hi4[0]), "r"(hi4[1]), "r"(hi4[2]), "r"(hi4[3]) : "memory");
      asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(a + 128u), "r"(lo4[0]), "r"(lo4[1]), "r"(lo4[2]), "r"(lo4[3]) : "memory");
    }
    fence_async_smem();
    __syncwarp();
    if (lane == 0) mbar_arrive(xb + 8u * (kBarQReady + buf));
    const float mine = p.scale * kLog2e * ksc / qs;
    return mine;  // lanes 4hh..4hh+3 computed head hh = lane >> 2
  };

  uint32_t s_par = 0, pfree_par = 0, resc_go_par = 0, sidx = 0, tile_base = 0;
  uint32_t stage = (uint32_t)q;  // ring stage of this class's next tile (gt = q, q + 4, ... over the whole stream)
  WorkIter<FLAT> wi;
  wi.init(p, sm.sched);
  Seg it, nxt;
  float sc2_next = 0.f;
  {
    WorkIter<FLAT> w2 = wi;
    if (w2.next(p, nxt)) sc2_next = prep_q(nxt, 0);
  }
  while (wi.next(p, it)) {
    const float sc2 = sc2_next;
    {  // the other Q buffer was last read by the previous segment, whose MMAs are all complete: fill it for the next one
      WorkIter<FLAT> w2 = wi;
      if (w2.next(p, nxt)) sc2_next = prep_q(nxt, (sidx + 1u) & 1u);
    }
    const int ntiles = (it.t1 - it.t0 + kTileTokens - 1) / kTileTokens;
    const uint32_t ob = sidx & 1u;
    const float al2 = (p.alibi != nullptr && active) ? __ldg(p.alibi + it.kvh * p.G + h) * kLog2e : 0.f;
    const bool use_alibi = p.alibi != nullptr;
    float m_ref = -INFINITY, m_true = -INFINITY, l = 0.f;
    bool fresh = true;
    uint32_t r[32];

    for (int ti = (int)(((uint32_t)q - tile_base) & 3u); ti < ntiles; ti += 4) {
      MBAR_WAIT(xb + 8u * (kBarSReady + q), s_par, 20, q, ti);
      s_par ^= 1u;
      tc_fence_after();
      uint32_t sr[16];
      tmem_ld_frag16(tmem + lane_base + kTmemS + 32u * q, sr);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(xb + 8u * (kBarSFree + q));  // S is in registers: the next Q.K^T of this class may land
      const int tok0 = it.t0 + ti * kTileTokens;
      const int nvalid = min(kTileTokens, it.t1 - tok0);
      float sv[8];
      float mx = -INFINITY;
#pragma unroll
      for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int b = 0; b < 2; ++b) {
          const int tt = 8 * j + 2 * c4 + b;  // token inside the tile
          float v = fmaf(__uint_as_float(sr[4 * j + 2 + b]), kInvLo, __uint_as_float(sr[4 * j + b])) * sc2;
          if (use_alibi) v = fmaf(al2, (float)(tok0 + tt - it.L + 1), v);
          if (tt >= nvalid) v = -INFINITY;
          sv[2 * j + b] = v;
          mx = fmaxf(mx, v);
        }
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
      m_true = fmaxf(m_true, mx);
      bool need = false;
      if (fresh) {
        m_ref = mx;
        fresh = false;
      } else if (mx > m_ref + kTau) {
        need = true;
      }
      const bool any_need = __any_sync(0xffffffffu, need && active);
      float f = 1.f;
      if (any_need && active) {
        const float m_new = fmaxf(m_ref, mx);
        f = fast_exp2(m_ref - m_new);
        l *= f;
        m_ref = m_new;
      }
      uint32_t hi4[2], lo4[2];
#pragma unroll
      for (int g4 = 0; g4 < 2; ++g4) {
        float pq[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          pq[e] = fast_exp2(sv[4 * g4 + e] - m_ref + kPShift);  // masked tokens: exp2(-inf) = 0
          l += pq[e];
        }
        split4_e4m3(pq, hi4[g4], lo4[g4]);
      }
      MBAR_WAIT(xb + 8u * (kBarPFree + q), pfree_par ^ 1u, 21, q, ti);  // P.V of this class's previous tile is done
      pfree_par ^= 1u;
      if (active) {
        // token pair j (tokens 8j + 2c4, +1) sits in chunk j / 2 at byte 8 (j & 1) + 2 c4 of row 16q + h (hi) / + 8 (lo)
        const uint32_t a = pbuf_u32 + (uint32_t)(16 * q + h) * 16u + 2u * (uint32_t)c4;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const uint32_t aj = a + (uint32_t)(j >> 1) * kAChunkStride + 8u * (uint32_t)(j & 1);
          const uint16_t hv = (uint16_t)(hi4[j >> 1] >> (16 * (j & 1))), lv = (uint16_t)(lo4[j >> 1] >> (16 * (j & 1)));
          asm volatile("st.shared.u16 [%0], %1;" ::"r"(aj), "h"(hv) : "memory");
          asm volatile("st.shared.u16 [%0], %1;" ::"r"(aj + 128u), "h"(lv) : "memory");
        }
      }
      if (nvalid < kTileTokens) {
        // Last, partial tile of a sequence: V bytes past seq_len are stale cache (possibly NaN patterns) or, for a missing
        // second block, whatever the ring held; 0 * NaN would poison the accumulator, so zero them in place
        // (reference: csrc/attention/attention_kernels.cuh:409-419).
        fence_async_smem();
        const uint32_t vb = ring_u32 + stage * UC::kStageBytes + UC::kVOffset;
        for (int row = lane; row < 2 * kD; row += kWarpSize) {
          const int keep = nvalid - 16 * (row >> 7);  // valid tokens of this row's block
          if (keep >= 16) continue;
          uint32_t w[4];
          asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]) : "r"(vb + row * 16u));
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const int kw = keep - 4 * k;
            w[k] &= kw >= 4 ? 0xffffffffu : (kw <= 0 ? 0u : ((1u << (8 * kw)) - 1u));
          }
          asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(vb + row * 16u), "r"(w[0]), "r"(w[1]), "r"(w[2]), "r"(w[3]) : "memory");
        }
      }
      fence_async_smem();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (any_need) st_volatile_smem(sm.resc_req + q, 1);
        mbar_arrive(xb + 8u * (kBarPReady + q));
      }
      if (any_need) {
        // the MMA warp drains the tensor pipe, then this warp rescales its accumulator lanes to the new reference
        MBAR_WAIT(xb + 8u * (kBarRescGo + q), resc_go_par, 22, q, ti);
        resc_go_par ^= 1u;
        tc_fence_after();
        // accumulator rows are TMEM lanes 0-7 (hi) and 8-15 (lo) of head lane & 7, whose factor lives in thread 4 (lane & 7)
        const float f_src = __shfl_sync(0xffffffffu, f, (lane & 7) * 4);
        const float f_row = row_lane ? f_src : 1.f;
#pragma unroll 1
        for (int ch = 0; ch < 4; ++ch) {
          const uint32_t ta = tmem + lane_base + kTmemO + 128u * ob + 32u * ch;
          tmem_ld32(ta, r);
#pragma unroll
          for (int i = 0; i < 32; ++i) r[i] = __float_as_uint(__uint_as_float(r[i]) * f_row);
          tmem_st32(ta, r);
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(xb + 8u * (kBarRescDone + q));
      }
      stage += 4u;
      if (stage >= (uint32_t)nst) stage -= (uint32_t)nst;
    }
    tile_base += (uint32_t)ntiles;

    // ---- end of the segment: park this class's partial in shared memory ---------------------------------------
    l += __shfl_xor_sync(0xffffffffu, l, 1);  // the four token quarters of a head
    l += __shfl_xor_sync(0xffffffffu, l, 2);
    MBAR_WAIT(xb + 8u * (kBarOReady + ob), (sidx >> 1) & 1u, 23, q, (int)sidx);
    tc_fence_after();
#pragma unroll 1
    for (int ch = 0; ch < 4; ++ch) {
      tmem_ld32(tmem + lane_base + kTmemO + 128u * ob + 32u * ch, r);
      if (row_lane) {
        float* dst = sm.merge_o + (size_t)(q * 16 + lane) * kMergeRow + 32 * ch;
#pragma unroll
        for (int i = 0; i < 32; i += 4)
          *reinterpret_cast<uint4*>(dst + i) = make_uint4(r[i], r[i + 1], r[i + 2], r[i + 3]);
      }
    }
    tc_fence_before();
    __syncwarp();
    if (lane == 0) mbar_arrive(xb + 8u * (kBarOFree + ob));
    if (c4 == 0) {
      sm.merge_mref[q * 8 + h] = fresh ? -INFINITY : m_ref;
      sm.merge_mtrue[q * 8 + h] = fresh ? -INFINITY : m_true;
      sm.merge_l[q * 8 + h] = fresh ? 0.f : l;
    }
    class_bar_sync();
    // ---- merge the four class partials (exact: every class kept its true running max) --------------------------
    const int d = tid;  // 128 class threads, one head dim each
    const bool whole = FLAT && it.starts && it.ends;
    const int slot = FLAT ? ((int)blockIdx.x * 2 + (it.starts ? 1 : 0)) * kFlatMaxGroup : 0;
    const bool single = !FLAT && it.L <= p.so.chunk;
    const bool want_partials = !FLAT && p.so.o != nullptr && (!single || p.so.normalized);
    for (int hh = 0; hh < p.G; ++hh) {
      float M = -INFINITY;
#pragma unroll
      for (int c = 0; c < kClassWarps; ++c) M = fmaxf(M, sm.merge_mtrue[c * 8 + hh]);
      float lsum = 0.f, o = 0.f;
#pragma unroll
      for (int c = 0; c < kClassWarps; ++c) {
        const float mr = sm.merge_mref[c * 8 + hh];
        if (mr == -INFINITY) continue;  // class without a tile in this segment
        const float w = exp2f(mr - kPShift - M);
        lsum = fmaf(sm.merge_l[c * 8 + hh], w, lsum);
        o = fmaf(fmaf(sm.merge_o[(c * 16 + 8 + hh) * kMergeRow + d], kInvLo, sm.merge_o[(c * 16 + hh) * kMergeRow + d]), w, o);
      }
      o *= vsc;
      const int head = it.kvh * p.G + hh;
      if constexpr (!FLAT) {
        // one 512-token partition of (seq, kv head): reference v2 format
        // (csrc/attention/attention_kernels.cuh:337-346, 472-484)
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
      } else {
        if (whole) {
          p.out[((int64_t)it.s * p.H + head) * kD + d] = __float2bfloat16_rn(o * __fdividef(1.f, lsum + 1e-6f));
        } else {
          p.ws_o[(int64_t)(slot + hh) * kD + d] = o;
          if (d == 0) {
            p.ws_m[slot + hh] = M;  // log2 units
            p.ws_l[slot + hh] = lsum;
          }
        }
      }
    }
    if constexpr (FLAT) {
      if (!whole) {
        // a pair split over CTAs: the CTA that delivers the last segment merges the partials
        // (same arithmetic as the reference reduce, csrc/attention/attention_kernels.cuh:594-657)
        __threadfence();
        class_bar_sync();
        const int c0 = cta_of_tile(it.pb, sm.sched->T, sm.sched->neff);
        const int c1 = cta_of_tile(it.pb + it.n - 1, sm.sched->T, sm.sched->neff);
        if (tid == 0) {
          int* cnt = p.counters + it.s * p.Hkv + it.kvh;
          const int old = atomicAdd(cnt, 1);
          const int last = old == c1 - c0;
          if (last) *cnt = 0;  // leave the counters zeroed for the next launch
          *sm.merge_flag = last;
        }
        class_bar_sync();
        if (*sm.merge_flag) {
          __threadfence();
          for (int hh = 0; hh < p.G; ++hh) {
            float M = -INFINITY;
            for (int c = c0; c <= c1; ++c)
              M = fmaxf(M, __ldcg(p.ws_m + (c * 2 + (c == c0 ? 1 : 0)) * kFlatMaxGroup + hh));
            float lsum = 0.f, o = 0.f;
            for (int c = c0; c <= c1; ++c) {
              const int sl = (c * 2 + (c == c0 ? 1 : 0)) * kFlatMaxGroup + hh;
              const float wgt = exp2f(__ldcg(p.ws_m + sl) - M);
              lsum = fmaf(__ldcg(p.ws_l + sl), wgt, lsum);
              o = fmaf(__ldcg(p.ws_o + (int64_t)sl * kD + d), wgt, o);
            }
            p.out[((int64_t)it.s * p.H + it.kvh * p.G + hh) * kD + d] =
                __float2bfloat16_rn(o * __fdividef(1.f, lsum + 1e-6f));
          }
        }
      }
    }
    class_bar_sync();  // merge buffers are free again
    ++sidx;
  }
}

// ---- the kernel ----------------------------------------------------------------------------------------------------
template <bool FLAT>
__global__ void __launch_bounds__(kUThreads, 1) paged_attention_umma_fp8_kernel(const StreamParams p) {
  extern __shared__ uint8_t smem_raw[];
  const int nst = p.nstages;
  const USmem sm = carve_smem(smem_raw, nst);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t full_u32 = smem_u32(sm.bars), empty_u32 = full_u32 + 8u * nst, xb = empty_u32 + 8u * nst;
  if (tid == 0) {
    for (int i = 0; i < 2 * nst; ++i) mbar_init(full_u32 + 8u * i, 1);
    for (int i = 0; i < kNumExtraBars; ++i) {
      const bool four = (i >= kBarQReady && i < kBarQReady + 2) || (i >= kBarOFree && i < kBarOFree + 2);
      mbar_init(xb + 8u * i, four ? kClassWarps : 1);
    }
    for (int i = 0; i < 4; ++i) sm.resc_req[i] = 0;
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  // P and Q operand buffers start all-zero: rows a class never writes must contribute nothing
  for (int i = tid; i < (2 * kQBufBytes + 4 * kPBufBytes) / 16; i += kUThreads)
    reinterpret_cast<uint4*>(sm.qbuf)[i] = make_uint4(0u, 0u, 0u, 0u);
  fence_async_smem();
  if (warp == kQkWarp) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(sm.tmem_slot)), "n"(kTmemCols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  if constexpr (FLAT) {
    if (warp == 0) compute_sched(p, sm.sched, lane);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *sm.tmem_slot;

  if (warp < kClassWarps) {
    class_warp<FLAT>(p, sm, nst, tmem, warp, lane);
  } else if (warp == kQkWarp) {
    qk_warp<FLAT>(p, sm, nst, tmem);
  } else if (warp == kPvWarp) {
    pv_warp<FLAT>(p, sm, nst, tmem);
  } else if (lane == 0) {
    producer_loop<kUB, true, FLAT, kUProducers>(p, sm.sched, warp - kPvWarp - 1, smem_u32(sm.ring), full_u32, empty_u32, nst);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kQkWarp)
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(kTmemCols) : "memory");
}

template <bool FLAT>
int launch_umma_t(StreamParams& sp, cudaStream_t s) {
  int nst = 16;  // 128 KiB in flight (profiles/r1_gather_read_bench.log); a multiple of 8 producers and 4 classes
  if (const int e = env_int("B200_STREAM_NSTAGES")) nst = max(8, min(16, e / 8 * 8));
  sp.nstages = nst;
  const size_t smem = usmem_bytes(nst);
  auto kern = paged_attention_umma_fp8_kernel<FLAT>;
  static thread_local int configured_device = -1;
  int dev = 0;
  cudaGetDevice(&dev);
  if (configured_device != dev) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) {
      set_error("cudaFuncSetAttribute(smem) failed: %s", cudaGetErrorString(e));
      return B200_ERR_CUDA;
    }
    configured_device = dev;
  }
  const int grid = min(min(device_sm_count(), kFlatMaxCtas), sp.total_items);
  kern<<<grid, kUThreads, smem, s>>>(sp);
  count_launch();
  return check_launch("paged_attention_umma_fp8");
}

}  // namespace

bool umma_fp8_variant_eligible(const b200_paged_attn_params_t& p, const char** why) {
  const char* w = nullptr;
  const int G = p.num_kv_heads > 0 ? p.num_heads / p.num_kv_heads : 0;
  if (!stream_variant_eligible(p, &w)) {
  } else if (p.kv_dtype != B200_KV_FP8_E4M3) w = "KV cache is not fp8";
  else if (p.block_size != kUB) w = "block_size is not 16";
  else if (G > 8) w = "GQA group size above 8";
  if (why) *why = w;
  return w == nullptr;
}

int launch_attention_umma_fp8(const b200_paged_attn_params_t& p, const SplitOut& so, cudaStream_t s) {
  const char* why = nullptr;
  B200_REQUIRE(umma_fp8_variant_eligible(p, &why), B200_ERR_UNSUPPORTED, "tcgen05 fp8 kernel not eligible: %s", why);
  B200_REQUIRE(so.chunk > 0 && so.chunk % kTileTokens == 0, B200_ERR_INVALID_ARGUMENT,
               "tcgen05 fp8 kernel: chunk %d is not a multiple of %d", so.chunk, kTileTokens);
  StreamParams sp = fill_stream_params(p);
  sp.so = so;
  sp.total_items = so.num_partitions * p.num_seqs * p.num_kv_heads;
  if (sp.total_items == 0) return B200_OK;
  return launch_umma_t<false>(sp, s);
}

int launch_attention_umma_fp8_flat(const b200_paged_attn_params_t& p, void* workspace, cudaStream_t s) {
  const char* why = nullptr;
  B200_REQUIRE(umma_fp8_variant_eligible(p, &why), B200_ERR_UNSUPPORTED, "tcgen05 fp8 kernel not eligible: %s", why);
  B200_REQUIRE(workspace != nullptr, B200_ERR_WORKSPACE, "tcgen05 fp8 kernel: workspace missing");
  StreamParams sp = fill_stream_params(p);
  sp.counters = static_cast<int*>(workspace);
  float* slots = reinterpret_cast<float*>(static_cast<char*>(workspace) + kFlatCounterBytes);
  sp.ws_o = slots;
  sp.ws_m = slots + (size_t)kFlatMaxCtas * 2 * kFlatMaxGroup * kD;
  sp.ws_l = sp.ws_m + (size_t)kFlatMaxCtas * 2 * kFlatMaxGroup;
  const int64_t tiles = (int64_t)p.num_seqs * p.num_kv_heads * ((p.max_seq_len + kTileTokens - 1) / kTileTokens);
  sp.total_items = (int)(tiles < (int64_t)kFlatMaxCtas ? tiles : (int64_t)kFlatMaxCtas);
  if (sp.total_items == 0) return B200_OK;
  return launch_umma_t<true>(sp, s);
}

}  // namespace b200
