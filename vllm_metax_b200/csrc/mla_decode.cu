// MLA decode attention for B200 (sm_100a): tcgen05 tensor cores, TMEM accumulators, TMA-fed paged latent rows.
// Replaces torch.ops._C.cutlass_mla_decode (csrc/attention/mla/cutlass_mla_entry.cu:28-38, checks and launch
// csrc/attention/mla/cutlass_mla_kernels.cu:154-225, schema csrc/torch_bindings.cpp:287-290) and the Triton split-K
// MLA decode (vllm_metax/attention/ops/triton_decode_attention.py:673-721).  DESIGN.md §10.
//
//   out[s,h,:512] = softmax_t(scale * [q_nope|q_pe][s,h] . kv[t,:576]) . kv[t,:512]       (one shared latent head)
//
// Formulation (everything transposed so that the head count is the MMA N extent and the accumulators fit one SM):
//   S^T[128 tokens x NG heads] = Ktile[128 x 576] . Q^T      9 chunks of 64 dims, 4 K-steps each  (A, B K-major SW128)
//   O^T[512 dv x NG heads]    += V^T[512 x 128] . P^T        4 M=128 blocks, 8 K-steps of 16 tokens (A MN-major SW128)
// A TMA box [64 dims x rows] with the 128-byte swizzle is the K-major A operand of the first product and the MN-major
// A operand of the second, so K chunks and V chunks are the same 16 KiB ring stages, loaded once for each product.
// TMEM: S^T double-buffered (2 x NG columns) + O^T (4 x NG columns).
//
// One CTA per (sequence, head group of NG <= 64, token split).  Warps: 0..3 = softmax / epilogue (one TMEM lane = one
// token row each), 4 = MMA issuer, 5..8 = TMA producers (one issuing lane each, ring fills dealt round-robin).  Softmax uses a lazily updated per-head reference
// m_ref (log2 units): p = 2^(s - m_ref); the column-wise (cross-lane) maximum is only taken when some logit passes
// m_ref by kTau, after which O^T and the running sums are rescaled.  Row sums are kept per thread and reduced once.
#include <cuda.h>
#include <string.h>

#include "stream_common.cuh"

namespace b200 {
namespace {

constexpr int kTile = 128;            // tokens per tile (MMA M of S^T)
constexpr int kChunkBytes = 16384;    // ring stage: 128 rows x 128 B
constexpr int kKChunks = 9;           // 576 / 64
constexpr int kVChunks = 8;           // 512 / 64
constexpr int kProducers = 4;         // issuing warps: one thread tops out near 47 GB/s, four reach ~120 GB/s per SM
constexpr int kMmaWarp = 4;
constexpr int kMlaThreads = (4 + 1 + kProducers) * 32;  // warps 0-3 softmax, 4 MMA issuer, 5-8 producers
// 64-head groups: four more softmax warps (12-15; warps 9-11 idle so that every role sits in whole warpgroups).  The
// main warp and the helper warp of a TMEM lane quadrant (warp % 4) split the 64 S^T columns in two, thread = (token
// row, head half): one softmax warp per scheduler alone left every dependent-instruction latency of a 64-column row
// exposed (7.7 k cycles per tile against ~4 k of MMA issue).  512 threads launch with 128 registers each; the MMA /
// producer warpgroups (warps 4-11) hand registers to the two softmax warpgroups with setmaxnreg (96 / 152).
constexpr int kMlaThreadsWide = 16 * 32;
constexpr int kHelperWarp0 = 12;
__host__ __device__ constexpr int mla_threads(int NG) { return NG == 64 ? kMlaThreadsWide : kMlaThreads; }
constexpr int kMaxSlots = 16;
constexpr float kTau = 16.f;          // re-reference when a logit passes m_ref by 2^16
constexpr int kMaxSplits = 256;
constexpr int kMinSplitTiles = 4;

struct MlaParams {
  void* out;
  float* lse;
  const int32_t* seq_lens;
  const int32_t* page_table;
  float* part_out;   // [S][H][max_splits][512] fp32 (normalised partial outputs)
  float* part_lse;   // [S][H][max_splits]
  float* debug_logits;
  long long out_seq_stride, out_head_stride, page_table_stride, lse_stride;
  float scale_log2;
  int S, H, groups, page_size, page_shift, box_rows, max_splits, split_tokens, nslots, f16;
  int flat;                 // persistent grid over the flat (sequence, group, tile) list
  int stamp_cta;            // development: the CTA whose timeline is stamped
  long long* cta_times;     // development: [cta][3] = globaltimer at start, at end, SM id
  long long* flat_prefix;   // [S + 2] unit prefix per sequence, the total and the mode, written by CTA 0 for the merge kernel
};

// ---- tcgen05 / TMA wrappers ---------------------------------------------------------------------------------
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// The issuing lane's own instruction stream is what bounds these kernels (profiles/r2_mla_*: ~30 scalar instructions of
// 64-bit descriptor arithmetic per MMA cost more than the MMA), so descriptors travel as two 32-bit halves: the high
// half (SBO, version, swizzle mode) is a constant, the low half (start address, LBO) takes one 32-bit add per MMA.
__device__ __forceinline__ void umma_f16(uint32_t d_tmem, uint32_t a_lo, uint32_t a_hi, uint32_t b_lo, uint32_t b_hi,
                                         uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n.reg .pred p;\n.reg .b64 da, db;\nmov.b64 da, {%1, %2};\nmov.b64 db, {%3, %4};\nsetp.ne.b32 p, %6, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %5, p;\n}\n" ::"r"(d_tmem),
      "r"(a_lo), "r"(a_hi), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(accumulate)
      : "memory");
}
// e4m3 x e4m3 -> f32 (kind::f8f6f4, K = 32 elements = the same 32 bytes per instruction as the 16-bit kinds)
__device__ __forceinline__ void umma_f8(uint32_t d_tmem, uint32_t a_lo, uint32_t a_hi, uint32_t b_lo, uint32_t b_hi,
                                        uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n.reg .pred p;\n.reg .b64 da, db;\nmov.b64 da, {%1, %2};\nmov.b64 db, {%3, %4};\nsetp.ne.b32 p, %6, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f8f6f4 [%0], da, db, %5, p;\n}\n" ::"r"(d_tmem),
      "r"(a_lo), "r"(a_hi), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(accumulate)
      : "memory");
}
// kind::f8f6f4 instruction descriptor: f32 accumulate, e4m3 (format 0) operands, A major (0 K, 1 MN), B K-major
__host__ __device__ constexpr uint32_t idesc_f8(int M, int N, int a_mn_major) {
  return (1u << 4) | ((uint32_t)a_mn_major << 15) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
// two floats -> two e4m3 bytes (first argument in the low byte), round-to-nearest-even, saturating; and back (exact)
__device__ __forceinline__ uint32_t cvt_e4m3x2(float lo, float hi) {
  uint16_t d;
  asm("cvt.rn.satfinite.e4m3x2.f32 %0, %1, %2;" : "=h"(d) : "f"(hi), "f"(lo));
  return (uint32_t)d;
}
__device__ __forceinline__ float2 e4m3x2_to_f32x2(uint32_t v) {
  uint32_t h2;
  asm("cvt.rn.f16x2.e4m3x2 %0, %1;" : "=r"(h2) : "h"((uint16_t)v));
  return __half22float2(*reinterpret_cast<__half2*>(&h2));
}
// 4 x 4 byte transpose across the four lanes of a quad: lane j ends with byte j of every lane's word, in lane order
__device__ __forceinline__ uint32_t quad_transpose_bytes(uint32_t w, int lane) {
  uint32_t y = __shfl_xor_sync(0xffffffffu, w, 1);
  w = (lane & 1) ? __byte_perm(w, y, 0x3715) : __byte_perm(w, y, 0x6240);
  y = __shfl_xor_sync(0xffffffffu, w, 2);
  return (lane & 2) ? __byte_perm(w, y, 0x3276) : __byte_perm(w, y, 0x5410);
}
// descriptor halves for the 128-byte swizzle: lo = (addr >> 4) | (LBO >> 4) << 16, hi = (SBO >> 4) | version 1 | layout 2
__device__ __forceinline__ uint32_t desc_lo(uint32_t addr, uint32_t lbo_bytes) {
  return ((addr & 0x3FFFF) >> 4) | ((lbo_bytes >> 4) << 16);
}
constexpr uint32_t kDescHiSw128 = (1024u >> 4) | (1u << 14) | (2u << 29);   // SBO = 1024 B (8 rows x 128 B)
// kind::f16 instruction descriptor: f32 accumulate, A/B element format (0 f16, 1 bf16), A major (0 K, 1 MN), B K-major
__host__ __device__ constexpr uint32_t idesc_f16(int M, int N, int fmt, int a_mn_major) {
  return (1u << 4) | ((uint32_t)fmt << 7) | ((uint32_t)fmt << 10) | ((uint32_t)a_mn_major << 15) |
         ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const void* tmap, int c0, int c1, int c2, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
      ::"r"(dst), "l"(tmap), "r"(bar), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
template <int N>
__device__ __forceinline__ void softmax_bar_sync() { asm volatile("bar.sync 1, %0;" ::"n"(N) : "memory"); }
template <int N>
__device__ __forceinline__ bool softmax_bar_or(bool v) {
  uint32_t r;
  asm volatile("{\n.reg .pred p, q;\nsetp.ne.b32 q, %1, 0;\nbar.red.or.pred p, 1, %2, q;\nselp.u32 %0, 1, 0, p;\n}\n"
               : "=r"(r)
               : "r"((uint32_t)v), "n"(N)
               : "memory");
  return r != 0;
}
__device__ __forceinline__ int f32_sortable(float x) {
  const int i = __float_as_int(x);
  return i ^ ((i >> 31) & 0x7fffffff);
}
__device__ __forceinline__ float sortable_f32(int i) { return __int_as_float(i ^ ((i >> 31) & 0x7fffffff)); }

// 32 TMEM lanes x N consecutive 32-bit columns -> N registers per thread (thread = lane)
template <int N>
__device__ __forceinline__ void tmem_ld(uint32_t taddr, uint32_t* r);
template <>
__device__ __forceinline__ void tmem_ld<16>(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr));
}
template <>
__device__ __forceinline__ void tmem_ld<32>(uint32_t taddr, uint32_t* r) {
  tmem_ld<16>(taddr, r);
  tmem_ld<16>(taddr + 16, r + 16);
}
template <>
__device__ __forceinline__ void tmem_ld<64>(uint32_t taddr, uint32_t* r) {
  tmem_ld<32>(taddr, r);
  tmem_ld<32>(taddr + 32, r + 32);
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t* r) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
      "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// ---- shared-memory carve-up ----------------------------------------------------------------------------------
struct MlaMisc {
  uint64_t full[kMaxSlots], empty[kMaxSlots];
  uint64_t s_full[2], s_free[2], p_full, p_free, o_full, q_full, q_free, o_free;
  uint32_t tmem_base;
  uint32_t pad[3];
  float mref[64];      // per-head softmax reference, log2 units
  float fac[64];       // rescale factor of the last re-reference; 1 / row sum in the epilogue
  int colmax[4][64];   // per-warp column maxima (sortable-int encoding)
};
__host__ __device__ constexpr size_t mla_smem_bytes(int NG, int nslots, int kchunks = kKChunks) {
  return 1024 + (size_t)NG * 128 * (kchunks + 2) + (size_t)nslots * kChunkBytes + sizeof(MlaMisc);
}

// A segment = consecutive tiles of one (sequence, head group).  Two ways to get them:
//  * split mode (p.flat == 0): the CTA's one segment comes from blockIdx (sequence, group, token split);
//  * flat mode  (p.flat == 1): the grid is persistent (one CTA per SM); the list of all (sequence, group, tile)
//    units is cut into gridDim.x equal ranges and CTA c walks range c, which may end one item, cover several whole
//    ones and begin another.  Partial items go to the CTA's two partial slots (first / last segment of its range)
//    and mla_merge_flat_kernel folds them; whole items are written directly.  No wave quantisation, and ragged
//    batches are balanced by construction.
struct MlaSeg {
  int seq, group, tile0, tile1, tiles_seq, L;
  int slot;      // flat mode: -1 = whole item (direct), else the CTA's partial slot 0 / 1; split mode: unused
  bool valid;
};

struct MlaFlatStart {   // computed once per CTA (warp 0), shared through smem
  long long u0, u1;     // this CTA's unit range (range mode)
  int seq, group, tile; // position of u0
  int item_mode;        // 1: whole items dealt round-robin (item c, c + P, ...): no partials; 0: equal unit ranges
  int sums[160];        // item mode's per-CTA tile totals (scratch for the decision)
};

__device__ __forceinline__ int mla_tiles_of(int L) { return L > 0 ? (L + kTile - 1) / kTile : 0; }
__device__ __forceinline__ int mla_units_of(int L, int unit) { return L > 0 ? (L + unit - 1) / unit : 0; }

// Flat-schedule prologue, run by every thread of the CTA: zero-fill of empty sequences, the item-mode / range-mode
// decision and this worker's start position (worker = CTA, or CTA pair; `unit` = tokens per scheduling unit).
__device__ void mla_flat_prologue(const MlaParams& p, MlaFlatStart& fs, int unit, int groups, int P, int worker,
                                  int slack_pct, bool header_writer, int tid, int nthreads, int nzero, int zero_rank) {
  const int warp = tid >> 5, lane = tid & 31;
  // empty sequences never show up in the unit list: their zeros / -inf are written here
  for (int s = worker; s < p.S; s += P) {
    if (p.seq_lens[s] > 0) continue;
    for (int i = tid + zero_rank * nthreads; i < p.H * 512; i += nzero * nthreads) {
      const long long o = (long long)s * p.out_seq_stride + (long long)(i / 512) * p.out_head_stride + i % 512;
      if (p.f16) reinterpret_cast<__half*>(p.out)[o] = __float2half_rn(0.f);
      else reinterpret_cast<__nv_bfloat16*>(p.out)[o] = __float2bfloat16_rn(0.f);
    }
    if (p.lse != nullptr && zero_rank == 0 && tid < p.H) p.lse[(long long)s * p.lse_stride + tid] = -INFINITY;
  }
  // Two ways to deal the work: whole items round-robin (no partial results, but only balanced when the items are
  // alike) or equal unit ranges (always balanced; items cut across workers go through fp32 partials and the merge
  // kernel).  Every worker evaluates both makespans from seq_lens and picks the same one.
  for (int i = tid; i < 160; i += nthreads) fs.sums[i] = 0;
  __syncthreads();
  const int Pc = min(P, 160);
  for (int i = tid; i < p.S * groups; i += nthreads) {
    const int t = mla_units_of(p.seq_lens[i / groups], unit);
    if (t > 0) atomicAdd(&fs.sums[i % Pc], t);
  }
  __syncthreads();
  if (warp == 0) {
    long long run = 0;
    for (int s0 = 0; s0 < p.S; s0 += 32) {
      const int s = s0 + lane;
      const long long v = s < p.S ? (long long)mla_units_of(p.seq_lens[s], unit) * groups : 0;
      long long inc = v;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const long long up = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += up;
      }
      if (header_writer && s < p.S) p.flat_prefix[s] = run + inc - v;
      run += __shfl_sync(0xffffffffu, inc, 31);
    }
    const long long W = run;
    int worst = 0;
    for (int c = lane; c < Pc; c += 32) worst = max(worst, fs.sums[c]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) worst = max(worst, __shfl_xor_sync(0xffffffffu, worst, o));
    const long long quota = (W + P - 1) / P;
    const int item_mode = P <= 160 && (long long)worst * 100 <= quota * (100 + slack_pct) + 200;
    if (header_writer && lane == 0) {
      p.flat_prefix[p.S] = W;
      p.flat_prefix[p.S + 1] = item_mode;
    }
    if (lane == 0) fs.item_mode = item_mode;
    if (item_mode) {
      // first non-empty item of this worker: items w, w + P, ... in (sequence, group) order
      int it = worker;
      const int n_items = p.S * groups;
      while (it < n_items && p.seq_lens[it / groups] <= 0) it += P;
      if (lane == 0) {
        fs.u0 = 0, fs.u1 = it < n_items ? 1 : 0;
        fs.seq = it < n_items ? it / groups : -1, fs.group = it < n_items ? it % groups : 0, fs.tile = 0;
      }
    } else {
      const long long u0 = W * worker / P, u1 = W * (worker + 1) / P;
      int fseq = -1, fgroup = 0, ftile = 0;
      long long acc = 0;
      if (u0 < u1) {
        for (int s0 = 0; s0 < p.S && fseq < 0; s0 += 32) {
          const int s = s0 + lane;
          const long long v = s < p.S ? (long long)mla_units_of(p.seq_lens[s], unit) * groups : 0;
          long long inc = v;
#pragma unroll
          for (int o = 1; o < 32; o <<= 1) {
            const long long up = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += up;
          }
          const bool here = v > 0 && acc + inc - v <= u0 && u0 < acc + inc;
          const unsigned m = __ballot_sync(0xffffffffu, here);
          if (m != 0) {
            const int src = __ffs(m) - 1;
            const long long before = __shfl_sync(0xffffffffu, acc + inc - v, src);
            const long long span = __shfl_sync(0xffffffffu, v, src);
            fseq = s0 + src;
            const int ts = (int)(span / groups);
            fgroup = (int)((u0 - before) / ts);
            ftile = (int)((u0 - before) % ts);
          }
          acc += __shfl_sync(0xffffffffu, inc, 31);
        }
      }
      if (lane == 0) {
        fs.u0 = u0, fs.u1 = u1;
        fs.seq = fseq, fs.group = fgroup, fs.tile = ftile;
      }
    }
  }
  __syncthreads();
}

// the flat walk (tile0 / tile1 / tiles_seq in scheduling units of `unit` tokens)
__device__ __forceinline__ MlaSeg mla_flat_first(const MlaParams& p, const MlaFlatStart& fs, int unit) {
  MlaSeg g;
  g.seq = fs.seq, g.group = fs.group;
  g.L = p.seq_lens[g.seq];
  g.tiles_seq = mla_units_of(g.L, unit);
  if (fs.item_mode) {
    g.tile0 = 0, g.tile1 = g.tiles_seq, g.slot = -1, g.valid = true;
    return g;
  }
  g.tile0 = fs.tile;
  g.tile1 = (int)min((long long)g.tiles_seq, g.tile0 + (fs.u1 - fs.u0));
  g.slot = (g.tile0 == 0 && g.tile1 == g.tiles_seq) ? -1 : 0;
  g.valid = true;
  return g;
}
__device__ __forceinline__ MlaSeg mla_flat_next(const MlaParams& p, const MlaFlatStart& fs, int unit, int groups, int P,
                                                MlaSeg g, long long& done) {
  if (fs.item_mode) {
    int it = g.seq * groups + g.group;
    const int n_items = p.S * groups;
    do { it += P; } while (it < n_items && p.seq_lens[it / groups] <= 0);
    if (it >= n_items) { g.valid = false; return g; }
    g.seq = it / groups, g.group = it % groups;
    g.L = p.seq_lens[g.seq];
    g.tiles_seq = mla_units_of(g.L, unit);
    g.tile0 = 0, g.tile1 = g.tiles_seq, g.slot = -1;
    return g;
  }
  done += g.tile1 - g.tile0;
  const long long left = (fs.u1 - fs.u0) - done;
  if (left <= 0 || g.tile1 < g.tiles_seq) { g.valid = false; return g; }
  if (++g.group == groups) {
    g.group = 0;
    do { ++g.seq; g.L = p.seq_lens[g.seq]; } while (g.L <= 0);
    g.tiles_seq = mla_units_of(g.L, unit);
  }
  g.tile0 = 0;
  g.tile1 = (int)min((long long)g.tiles_seq, left);
  g.slot = g.tile1 == g.tiles_seq ? -1 : 1;
  return g;
}

// FP8: q and cache are e4m3 bytes (the reference op's float8_e4m3fn form): rows of 576 bytes = four 128-byte chunks + a
// fifth whose upper half lies outside the tensor and is zero-filled by TMA; K = 32 per instruction, so half as many S^T
// products; the probabilities go to the tensor core as TWO e4m3 terms (hi = rn(p), lo = rn(p - hi): ~7 bits, where one
// term would miss the reference's own fp8 tolerance on short contexts) with p scaled to at most 2^8.
template <int NG, bool FP8>
__global__ void __launch_bounds__(mla_threads(NG), 1)
mla_decode_kernel(const MlaParams p, const __grid_constant__ TmaDesc tm_kv, const __grid_constant__ TmaDesc tm_qn,
                  const __grid_constant__ TmaDesc tm_qp) {
  constexpr int KC = FP8 ? 5 : kKChunks;           // K chunks per tile (16 KiB each)
  constexpr int VC = FP8 ? 4 : kVChunks;           // V chunks per tile
  constexpr int KH = FP8 ? 6 : 10;                 // ring halves of the K phase (an even count: chunks + unused)
  // e4m3 probabilities must stay below 448: p = 2^(x - m_ref + kShift) <= 2^(kShift + kTauX) = 2^8.  A re-reference
  // puts m_ref kHead above the column maximum, so the next one needs a logit 5 log2 units above everything seen so
  // far (with m_ref = the maximum itself and a small kTauX, 64 heads re-referenced every few tiles: ~10 k cycles each).
  constexpr float kTauX = FP8 ? 3.f : kTau;
  constexpr float kShift = FP8 ? 5.f : 0.f;
  constexpr float kHead = FP8 ? 2.f : 0.f;
  constexpr int NTHR = mla_threads(NG);              // threads of the CTA
  constexpr bool SW2 = NG == 64;                     // two softmax threads per token row (helper warps 12-15)
  constexpr int NS = SW2 ? 256 : 128;                // softmax threads
  constexpr int CT = SW2 ? 32 : NG;                  // S^T columns (heads) per softmax thread
  // fp8, groups of <= 32 heads: the two e4m3 terms of P^T ([NG rows hi | NG rows lo], contiguous in shared memory) go to
  // the tensor core as ONE B operand of 2 NG rows, so an O^T step is one instruction instead of two (an instruction
  // with a shared-memory A operand costs the same for any N); the hi and lo products land in separate TMEM columns and
  // are added in the epilogue.  64-head groups have no TMEM columns to spare for that (2 NG + 4 * 2 NG > 512).
  constexpr bool MT = FP8 && NG <= 32;
  constexpr int NO = MT ? 2 * NG : NG;               // TMEM columns of one 128-row O^T block
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  __shared__ MlaFlatStart fstart;
  if (!p.flat) {
    const int seq = blockIdx.y, group = blockIdx.x % p.groups, split = blockIdx.x / p.groups;
    const int L = p.seq_lens[seq];
    if (L <= 0) {  // empty sequence: zeros and lse = -inf, written by the split-0 CTAs
      if (split == 0) {
        const int h0 = group * NG, nh = min(NG, p.H - h0);
        for (int i = tid; i < nh * 512; i += NTHR) {
          const long long o = (long long)seq * p.out_seq_stride + (long long)(h0 + i / 512) * p.out_head_stride + i % 512;
          if (p.f16) reinterpret_cast<__half*>(p.out)[o] = __float2half_rn(0.f);
          else reinterpret_cast<__nv_bfloat16*>(p.out)[o] = __float2bfloat16_rn(0.f);
        }
        if (p.lse != nullptr && tid < nh) p.lse[(long long)seq * p.lse_stride + h0 + tid] = -INFINITY;
      }
      return;
    }
    if (split * p.split_tokens >= L) return;
  }
  if (p.flat) {
    mla_flat_prologue(p, fstart, kTile, p.groups, (int)gridDim.x, (int)blockIdx.x, 20,
                      blockIdx.x == 0, tid, NTHR, 1, 0);
    if (fstart.u0 >= fstart.u1) return;   // more CTAs than units
  }
  // the segment walk: every role runs the same deterministic sequence
  auto first_seg = [&]() {
    if (p.flat) return mla_flat_first(p, fstart, kTile);
    MlaSeg g;
    g.seq = blockIdx.y, g.group = blockIdx.x % p.groups;
    g.L = p.seq_lens[g.seq];
    g.tiles_seq = mla_tiles_of(g.L);
    const int split = blockIdx.x / p.groups;
    g.tile0 = split * (p.split_tokens / kTile);
    g.tile1 = min(g.tiles_seq, g.tile0 + p.split_tokens / kTile);
    g.slot = split;
    g.valid = g.tile0 < g.tile1;
    return g;
  };
  // `done` = units consumed so far by this CTA (flat range mode)
  auto next_seg = [&](MlaSeg g, long long& done) {
    if (!p.flat) { g.valid = false; return g; }
    return mla_flat_next(p, fstart, kTile, p.groups, (int)gridDim.x, g, done);
  };

  long long cta_t0 = 0;
  if (p.cta_times != nullptr && tid == 0) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(cta_t0));
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* const base_ptr = smem_raw + (base - smem_u32(smem_raw));
  const uint32_t q_u32 = base;                                  // 9 chunks [NG rows x 128 B]
  const uint32_t p_u32 = q_u32 + KC * NG * 128;                 // 2 chunks [NG rows x 128 B]: 64 tokens each (16-bit) / hi and lo terms (fp8)
  const uint32_t ring_u32 = p_u32 + 2 * NG * 128;
  uint8_t* const ring_ptr = base_ptr + (ring_u32 - base);
  MlaMisc* const misc = reinterpret_cast<MlaMisc*>(ring_ptr + (size_t)p.nslots * kChunkBytes);
  const int nslots = p.nslots;

  if (tid == 0) {
    for (int i = 0; i < (nslots >> 1); ++i) {
      mbar_init(smem_u32(&misc->full[i]), 2);   // the two producers that fill the pair's halves
      mbar_init(smem_u32(&misc->empty[i]), 1);  // one tcgen05.commit
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(smem_u32(&misc->s_full[i]), 1);
      mbar_init(smem_u32(&misc->s_free[i]), NS);
    }
    mbar_init(smem_u32(&misc->p_full), NS);
    mbar_init(smem_u32(&misc->p_free), 1);
    mbar_init(smem_u32(&misc->o_full), 1);
    mbar_init(smem_u32(&misc->q_full), 1);
    mbar_init(smem_u32(&misc->q_free), 1);               // the segment's last S^T product has read the query
    mbar_init(smem_u32(&misc->o_free), NS); // the segment's O^T has been read out
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  constexpr uint32_t kColsNeed = 2 * NG + 4 * NO;
  constexpr uint32_t kCols = kColsNeed <= 128 ? 128 : (kColsNeed <= 256 ? 256 : 512);  // power of two
  static_assert(kColsNeed <= 512, "TMEM columns");
  if (warp == kMmaWarp) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&misc->tmem_base)),
                 "r"(kCols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = misc->tmem_base;
  const uint32_t tm_s[2] = {tmem, tmem + NG};
  const uint32_t tm_o = tmem + 2 * NG;
  // development timeline (tools/mla_debug.py --timeline): clock64 stamps of CTA 0's first segment, [role][tile][4]
  long long* const stamps = (p.debug_logits != nullptr && blockIdx.x == p.stamp_cta && blockIdx.y == 0)
                                ? reinterpret_cast<long long*>(p.debug_logits + 128 * 64) : nullptr;
  auto stamp = [&](int role, int tile, int k) {
    if (stamps != nullptr && tile < 64) stamps[(role * 64 + tile) * 4 + k] = clock64();
  };

  // Ring protocol: the ring is npairs pairs of 16 KiB chunks.  A pair has ONE full barrier (two producer arrivals,
  // each announcing its chunk's bytes) and ONE empty barrier (one tcgen05.commit), so the MMA lane pays one wait and
  // one commit per 32 KiB.  Per tile: 5 K pairs (chunks 0..8 + an unused half) for S^T, then 4 V pairs for O^T.
  // Tile-indexed barriers (S^T buffers, P^T) run on a tile counter that continues across segments; segment-indexed
  // ones (query, O^T) on the segment counter.
  const int npairs = nslots >> 1;
  // (setmaxnreg sits at the head of each role's own branch: ptxas sizes the register file of a region by the
  // setmaxnreg that dominates it, and code after a join of a 96- and a 152-register path gets the smaller one)
  if (warp > kMmaWarp && warp <= kMmaWarp + kProducers) {
    if constexpr (SW2) asm volatile("setmaxnreg.dec.sync.aligned.u32 96;\n");
    // =================== TMA producers: half-pair f of the ring sequence is filled by producer f % kProducers ===================
    const int pw = warp - kMmaWarp - 1;
    if (lane == 0) {
      const int box_rows = p.box_rows;
      const uint32_t box_bytes = (uint32_t)box_rows * 128u;
      int pair = 0, turn = 0;
      uint32_t phase = 0;  // parity of the ring pass
      long long done = 0;
      int nseg = 0;
      for (MlaSeg g = first_seg(); g.valid; g = next_seg(g, done), ++nseg) {
        const int L = g.L, t0 = g.tile0 * kTile, T = g.tile1 - g.tile0;
        if (pw == 0) {
          // query group: 8 latent chunks + the rope chunk, heads beyond H zero-filled by the tensor map bounds
          if (nseg > 0) mbar_wait(smem_u32(&misc->q_free), (nseg - 1) & 1);
          mbar_expect_tx(smem_u32(&misc->q_full), (uint32_t)(KC * NG * 128));
          for (int c = 0; c < VC; ++c)
            tma_load_3d(q_u32 + c * NG * 128, &tm_qn, c * (FP8 ? 128 : 64), g.group * NG, g.seq, smem_u32(&misc->q_full));
          tma_load_3d(q_u32 + VC * NG * 128, &tm_qp, 0, g.group * NG, g.seq, smem_u32(&misc->q_full));
        }
        const int32_t* table = p.page_table + (long long)g.seq * p.page_table_stride;
        // halves 0 .. nhalves-1 of the tile starting at token tok0: chunk index = half (the 10th K half is left unused)
        auto load_halves = [&](int tok0, int nhalves, int nchunks) {
          const int valid = min(kTile, L - tok0);                  // rows of this tile that hold cache entries
          const int vbox = (valid + box_rows - 1) / box_rows;
          int page[8], row[8];
#pragma unroll
          for (int b = 0; b < 8; ++b) {
            if (b < vbox) {
              const int tok = tok0 + b * box_rows;
              page[b] = table[tok >> p.page_shift];
              row[b] = tok & (p.page_size - 1);
            }
          }
          for (int c = 0; c < nhalves; ++c) {
            if (turn == pw) {
              mbar_wait(smem_u32(&misc->empty[pair]), phase ^ 1);
              const uint32_t full = smem_u32(&misc->full[pair]);
              if (c < nchunks) {
                mbar_expect_tx(full, box_bytes * (uint32_t)vbox);
                const uint32_t dst = ring_u32 + (uint32_t)(2 * pair + (c & 1)) * kChunkBytes;
#pragma unroll
                for (int b = 0; b < 8; ++b)
                  if (b < vbox) tma_load_3d(dst + b * box_bytes, &tm_kv, c * (FP8 ? 128 : 64), row[b], page[b], full);
              } else {
                mbar_arrive(full);
              }
            }
            if (++turn == kProducers) turn = 0;
            if (c & 1) {
              if (++pair == npairs) { pair = 0; phase ^= 1; }
            }
          }
        };
        load_halves(t0, KH, KC);
        for (int i = 0; i < T; ++i) {
          if (i + 1 < T) load_halves(t0 + (i + 1) * kTile, KH, KC);
          load_halves(t0 + i * kTile, VC, VC);
        }
      }
    }
  } else if (SW2 && warp > kMmaWarp + kProducers && warp < kHelperWarp0) {
    // idle warps of the MMA / producer warpgroups
    if constexpr (SW2) asm volatile("setmaxnreg.dec.sync.aligned.u32 96;\n");
  } else if (warp == kMmaWarp) {
    if constexpr (SW2) asm volatile("setmaxnreg.dec.sync.aligned.u32 96;\n");
    // =================== MMA issuer (one lane) ===================
    if (lane == 0) {
      const int fmt = p.f16 ? 0 : 1;
      const uint32_t idesc_s = FP8 ? idesc_f8(128, NG, 0) : idesc_f16(128, NG, fmt, 0);
      const uint32_t idesc_o = FP8 ? idesc_f8(128, NO, 1) : idesc_f16(128, NG, fmt, 1);
      int pair = 0;
      uint32_t phase = 0;
      auto advance = [&]() {
        if (++pair == npairs) { pair = 0; phase ^= 1; }
      };
      long long acc[4] = {0, 0, 0, 0}, tmark = 0;   // development: cycles in waits / S issue / PV issue / commits
      auto tick = [&]() { if (stamps != nullptr) tmark = clock64(); };
      auto tock = [&](int k) { if (stamps != nullptr) acc[k] += clock64() - tmark; };
      const uint32_t q_lo = desc_lo(q_u32, 0), p_lo = desc_lo(p_u32, 0);
      long long done = 0;
      int nseg = 0;
      int gt = 0;   // tile counter across segments
      for (MlaSeg g = first_seg(); g.valid; g = next_seg(g, done), ++nseg) {
        const int t0 = g.tile0 * kTile, t1 = min(g.L, g.tile1 * kTile), T = g.tile1 - g.tile0;
        const int gt0 = gt;
        mbar_wait(smem_u32(&misc->q_full), nseg & 1);
        auto issue_s = [&](int i) {
          const int x = gt0 + i, b = x & 1;
          if (x >= 2) mbar_wait(smem_u32(&misc->s_free[b]), ((x >> 1) & 1) ^ 1);
          tc_fence_after();
          stamp(0, x, 0);
          for (int c2 = 0; c2 < KH / 2; ++c2) {
            tick();
            mbar_wait(smem_u32(&misc->full[pair]), phase);
            tc_fence_after();
            tock(0);
            const uint32_t a_lo = desc_lo(ring_u32 + (uint32_t)pair * 2 * kChunkBytes, 0);
            const uint32_t b_lo = q_lo + (uint32_t)(c2 * 2 * NG * 128 >> 4);
            tick();
#pragma unroll
            for (int h = 0; h < 2; ++h) {
              if (2 * c2 + h < KC) {
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                  if (FP8 && 2 * c2 + h == 4 && k >= 2) continue;   // the rope chunk's upper 64 bytes are TMA zero fill
                  if constexpr (FP8)
                    umma_f8(tm_s[b], a_lo + ((h * kChunkBytes + k * 32) >> 4), kDescHiSw128,
                            b_lo + ((h * NG * 128 + k * 32) >> 4), kDescHiSw128, idesc_s, (c2 | h | k) != 0);
                  else
                    umma_f16(tm_s[b], a_lo + ((h * kChunkBytes + k * 32) >> 4), kDescHiSw128,
                             b_lo + ((h * NG * 128 + k * 32) >> 4), kDescHiSw128, idesc_s, (c2 | h | k) != 0);
                }
              }
            }
            tock(1);
            tick();
            umma_commit(smem_u32(&misc->empty[pair]));
            tock(3);
            advance();
          }
          umma_commit(smem_u32(&misc->s_full[b]));
          if (i == T - 1) umma_commit(smem_u32(&misc->q_free));   // the query buffer may be reloaded once this is done
          stamp(0, gt0 + i, 1);
        };
        auto issue_pv = [&](int i) {
          const int x = gt0 + i;
          const int rows = min(kTile, t1 - (t0 + i * kTile));   // valid token rows of this tile
          constexpr int KS = FP8 ? 32 : 16;                      // tokens per K step
          const int ksteps = (rows + KS - 1) / KS;
          const int zero_from = rows, zero_to = ksteps * KS;     // rows of the last K step that are past the sequence
          mbar_wait(smem_u32(&misc->p_full), x & 1);
          if (i == 0 && nseg > 0) mbar_wait(smem_u32(&misc->o_free), (nseg - 1) & 1);  // previous O^T has been read out
          tc_fence_after();
          stamp(0, x, 2);
          for (int j = 0; j < VC / 2; ++j) {
            tick();
            mbar_wait(smem_u32(&misc->full[pair]), phase);
            tc_fence_after();
            tock(0);
            const uint32_t a_addr = ring_u32 + (uint32_t)pair * 2 * kChunkBytes;
            if (zero_from < zero_to) {
              // V rows past the sequence end may hold anything (0 * NaN = NaN): zero them before the tensor core reads
              for (int hlf = 0; hlf < 2; ++hlf)
                for (int u = zero_from * 8; u < zero_to * 8; ++u)
                  *reinterpret_cast<uint4*>(ring_ptr + (size_t)(2 * pair + hlf) * kChunkBytes + (size_t)u * 16) = make_uint4(0, 0, 0, 0);
              fence_async_smem();
            }
            tick();
            if constexpr (FP8) {
              // one chunk = one 128-row dv block (128 one-byte elements per swizzled row): blocks 2j and 2j + 1
#pragma unroll
              for (int hlf = 0; hlf < 2; ++hlf) {
                const uint32_t a_lo = desc_lo(a_addr + hlf * kChunkBytes, 0);
                const uint32_t d_o = tm_o + (2 * j + hlf) * NO;
#pragma unroll
                for (int ks = 0; ks < 4; ++ks)
                  if (ks < ksteps) {
                    umma_f8(d_o, a_lo + ((ks * 4096) >> 4), kDescHiSw128, p_lo + ((ks * 32) >> 4), kDescHiSw128, idesc_o, (i | ks) != 0);
                    if constexpr (!MT)   // the lo term as a second product into the same columns
                      umma_f8(d_o, a_lo + ((ks * 4096) >> 4), kDescHiSw128, p_lo + ((NG * 128 + ks * 32) >> 4), kDescHiSw128, idesc_o, 1);
                  }
              }
            } else {
              const uint32_t a_lo = desc_lo(a_addr, kChunkBytes);
              const uint32_t d_o = tm_o + j * NO;
#pragma unroll
              for (int ks = 0; ks < 8; ++ks)
                if (ks < ksteps)
                  umma_f16(d_o, a_lo + ((ks * 2048) >> 4), kDescHiSw128,
                           p_lo + (((ks >> 2) * NG * 128 + (ks & 3) * 32) >> 4), kDescHiSw128, idesc_o, (i | ks) != 0);
            }
            tock(2);
            tick();
            umma_commit(smem_u32(&misc->empty[pair]));
            tock(3);
            advance();
          }
          umma_commit(smem_u32(&misc->p_free));
          stamp(0, x, 3);
        };
        issue_s(0);
        for (int i = 0; i < T; ++i) {
          if (i + 1 < T) issue_s(i + 1);
          issue_pv(i);
        }
        umma_commit(smem_u32(&misc->o_full));
        gt += T;
        if (nseg == 0 && stamps != nullptr) {
          for (int k = 0; k < 4; ++k) stamps[(0 * 64 + 63) * 4 + k] = acc[k];
          stamps[(1 * 64 + 63) * 4 + 0] = T;
        }
      }
    }
    __syncwarp();
  } else {
    // =================== softmax / epilogue warps: thread = one token row = one TMEM lane ===================
    // warps 0..3 and, for 64-head groups, the helper warps 12..15: the main warp of a quadrant takes columns 0..31,
    // the helper columns 32..63
    if constexpr (SW2) asm volatile("setmaxnreg.inc.sync.aligned.u32 152;\n");
    const bool helper = SW2 && warp >= kHelperWarp0;
    const int quad = warp & 3;                   // TMEM lane quadrant this warp may access
    const int row = quad * 32 + lane;            // token row within the tile / dv row within an O^T block
    const int st = helper ? 128 + (warp - kHelperWarp0) * 32 + lane : tid;   // 0..NS-1; per-head work: st < NG
    const int cbase = helper ? 32 : 0;           // first S^T / O^T column of this thread
    const uint32_t lane_base = (uint32_t)(quad * 32) << 16;
    constexpr int HC = NG < 32 ? NG : 32;        // columns handled per pass (register budget)
    long long done = 0;
    int nseg = 0;
    int gt = 0;   // tile counter across segments
    for (MlaSeg g = first_seg(); g.valid; g = next_seg(g, done), ++nseg) {
    const int t0 = g.tile0 * kTile, t1 = min(g.L, g.tile1 * kTile), T = g.tile1 - g.tile0;
    const int h0 = g.group * NG;
    const int nh = min(NG, p.H - h0);           // real heads of this group
    float lsum[CT];
#pragma unroll
    for (int h = 0; h < CT; ++h) lsum[h] = 0.f;
    if (st < 64) misc->mref[st] = -INFINITY;
    softmax_bar_sync<NS>();

    for (int i = 0; i < T; ++i, ++gt) {
      const int b = gt & 1;
      mbar_wait(smem_u32(&misc->s_full[b]), (gt >> 1) & 1);
      tc_fence_after();
      const bool valid = t0 + i * kTile + row < t1;
      if (tid == 0) stamp(1, gt, 0);
      // ---- pass 1: does any logit pass its reference by kTau?  (S^T is read again in pass 2: TMEM reads are cheap) ----
      bool over = false;
#pragma unroll
      for (int cc = 0; cc < CT; cc += HC) {
        const int c0 = cbase + cc;
        uint32_t r[HC];
        tmem_ld<HC>(tm_s[b] + c0 + lane_base, r);
        tmem_ld_wait();
        if (p.debug_logits != nullptr && i == 0 && nseg == 0 && blockIdx.x == 0 && blockIdx.y == 0) {
#pragma unroll
          for (int h = 0; h < HC; ++h) p.debug_logits[row * 64 + c0 + h] = __uint_as_float(r[h]);
        }
#pragma unroll
        for (int h = 0; h < HC; ++h)
          over |= valid && __uint_as_float(r[h]) * p.scale_log2 > misc->mref[c0 + h] + kTauX;
      }
      if (softmax_bar_or<NS>(over)) {
        // ---- rare path: raise the references to this tile's column maxima, rescale O^T and the running sums ----
#pragma unroll
        for (int cc = 0; cc < CT; cc += HC) {
          const int c0 = cbase + cc;
          uint32_t r[HC];
          tmem_ld<HC>(tm_s[b] + c0 + lane_base, r);
          tmem_ld_wait();
#pragma unroll
          for (int h = 0; h < HC; ++h) {
            const float x = valid ? __uint_as_float(r[h]) * p.scale_log2 : -INFINITY;
            const int m = __reduce_max_sync(0xffffffffu, f32_sortable(x));
            if (lane == 0) misc->colmax[quad][c0 + h] = m;
          }
        }
        softmax_bar_sync<NS>();
        if (st < NG) {
          const int m = max(max(misc->colmax[0][st], misc->colmax[1][st]), max(misc->colmax[2][st], misc->colmax[3][st]));
          const float m_old = misc->mref[st];
          float m_new = fmaxf(m_old, sortable_f32(m) + kHead);
          if (m_new == -INFINITY) m_new = 0.f;
          misc->fac[st] = m_old == -INFINITY ? 0.f : exp2f(m_old - m_new);
          misc->mref[st] = m_new;
        }
        softmax_bar_sync<NS>();
#pragma unroll
        for (int h = 0; h < CT; ++h) lsum[h] *= misc->fac[cbase + h];
        if (i > 0) {
          mbar_wait(smem_u32(&misc->p_free), (gt - 1) & 1);  // P.V of the previous tile has landed in O^T
          tc_fence_after();
          for (int j = 0; j < 4; ++j) {
#pragma unroll
            for (int cl = 0; cl < CT; cl += 16) {
              const int c = cbase + cl;
#pragma unroll
              for (int t = 0; t < (MT ? 2 : 1); ++t) {   // merged terms: the hi columns, then the lo columns
                uint32_t o[16];
                tmem_ld<16>(tm_o + j * NO + t * NG + c + lane_base, o);
                tmem_ld_wait();
#pragma unroll
                for (int e = 0; e < 16; ++e) o[e] = __float_as_uint(__uint_as_float(o[e]) * misc->fac[c + e]);
                tmem_st16(tm_o + j * NO + t * NG + c + lane_base, o);
              }
            }
          }
          tmem_st_wait();
          tc_fence_before();
        }
      }
      if (tid == 0) stamp(1, gt, 1);
      // ---- pass 2: p = 2^(x - m_ref); 16-bit values into the P^T operand, fp32 into the running sums ----
      // P^T[h][token], K-major SW128.  16-bit: chunk = row / 64, 16-byte unit (row % 64) / 8 XOR (h % 8).
      // fp8: one chunk per term (hi, lo), one byte per token: unit row / 16 XOR (h % 8).
      const uint32_t chunk = FP8 ? p_u32 : p_u32 + (row >> 6) * NG * 128;
      const uint32_t unit = FP8 ? (uint32_t)(row >> 4) : (uint32_t)((row & 63) >> 3);
      const uint32_t within = FP8 ? (uint32_t)(row & 15) : (uint32_t)((row & 7) * 2);
#pragma unroll
      for (int cc = 0; cc < CT; cc += HC) {
        const int c0 = cbase + cc;
        uint32_t r[HC];
        tmem_ld<HC>(tm_s[b] + c0 + lane_base, r);
        tmem_ld_wait();
        if (cc + HC == CT) {   // this thread's last read of this S^T buffer
          tc_fence_before();
          mbar_arrive(smem_u32(&misc->s_free[b]));
        }
        if constexpr (FP8) {
          // two e4m3 terms per probability (hi = rn(p), lo = rn(p - hi)), converted two heads at a time; the bytes of
          // four heads are transposed across the four lanes of a quad (= four consecutive tokens), so every lane
          // stores one 32-bit word per term and head group instead of eight single bytes
#pragma unroll
          for (int h = 0; h < HC; h += 4) {
            float pv[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const float x = valid ? __uint_as_float(r[h + e]) * p.scale_log2 : -INFINITY;
              pv[e] = fast_exp2(x - misc->mref[c0 + h + e] + kShift);
              lsum[cc + h + e] += pv[e];
            }
            const uint32_t hi01 = cvt_e4m3x2(pv[0], pv[1]), hi23 = cvt_e4m3x2(pv[2], pv[3]);
            const float2 f01 = e4m3x2_to_f32x2(hi01), f23 = e4m3x2_to_f32x2(hi23);
            const uint32_t lo01 = cvt_e4m3x2(pv[0] - f01.x, pv[1] - f01.y), lo23 = cvt_e4m3x2(pv[2] - f23.x, pv[3] - f23.y);
            r[h] = hi01 | (hi23 << 16);        // bytes = heads h .. h+3 of this token, hi term
            r[h + 1] = lo01 | (lo23 << 16);    // lo term
          }
          if (cc == 0 && gt > 0) mbar_wait(smem_u32(&misc->p_free), (gt - 1) & 1);
          if (cc == 0 && tid == 0) stamp(1, gt, 2);
#pragma unroll
          for (int h = 0; h < HC; h += 4) {
            const uint32_t thi = quad_transpose_bytes(r[h], lane), tlo = quad_transpose_bytes(r[h + 1], lane);
            const int hd = c0 + h + (lane & 3);   // the head whose four tokens this lane now holds
            const uint32_t addr = chunk + hd * 128 + ((unit ^ (hd & 7)) << 4) + (within & ~3u);
            asm volatile("st.shared.u32 [%0], %1;" ::"r"(addr), "r"(thi) : "memory");
            asm volatile("st.shared.u32 [%0], %1;" ::"r"(addr + NG * 128), "r"(tlo) : "memory");
          }
        } else {
#pragma unroll
          for (int h = 0; h < HC; ++h) {
            const float x = valid ? __uint_as_float(r[h]) * p.scale_log2 : -INFINITY;
            const float pv = fast_exp2(x - misc->mref[c0 + h] + kShift);
            lsum[cc + h] += pv;
            r[h] = p.f16 ? (uint32_t)__half_as_ushort(__float2half_rn(pv)) : (uint32_t)__bfloat16_as_ushort(__float2bfloat16_rn(pv));
          }
          if (cc == 0 && gt > 0) mbar_wait(smem_u32(&misc->p_free), (gt - 1) & 1);
          if (cc == 0 && tid == 0) stamp(1, gt, 2);
#pragma unroll
          for (int h = 0; h < HC; ++h) {
            const uint32_t addr = chunk + (c0 + h) * 128 + ((unit ^ (h & 7)) << 4) + within;
            asm volatile("st.shared.u16 [%0], %1;" ::"r"(addr), "h"((uint16_t)r[h]) : "memory");
          }
        }
      }
      fence_async_smem();
      mbar_arrive(smem_u32(&misc->p_full));
      if (tid == 0) stamp(1, gt, 3);
    }

    // ---- epilogue of the segment: reduce the per-thread row sums over the 128 token rows, normalise O^T, store ----
    if (tid == 0) stamp(2, nseg, 0);
    mbar_wait(smem_u32(&misc->o_full), nseg & 1);   // every MMA of this segment is complete
    tc_fence_after();
    if (tid == 0) stamp(2, nseg, 1);
    {
      float* wsum = reinterpret_cast<float*>(&misc->colmax[0][0]);   // [4 warps][64]: per-warp column sums
#pragma unroll
      for (int h = 0; h < CT; ++h) {
        float v = lsum[h];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0) wsum[quad * 64 + cbase + h] = v;
      }
      softmax_bar_sync<NS>();
      if (st < NG) misc->fac[st] = wsum[st] + wsum[64 + st] + wsum[128 + st] + wsum[192 + st];
      softmax_bar_sync<NS>();
    }
    if (tid == 0) stamp(2, nseg, 2);
    const bool direct = g.tile0 == 0 && g.tile1 == g.tiles_seq;
    // where a partial goes: split mode [seq][head][split], flat mode [cta][slot][head of the group]
    const long long part_row = p.flat ? ((long long)blockIdx.x * 2 + g.slot) * NG : 0;
    if (st < NG) {
      const float l = misc->fac[st];
      if (st < nh) {
        const float lse = (misc->mref[st] - kShift + log2f(l)) * kLn2;
        if (direct) {
          if (p.lse != nullptr) p.lse[(long long)g.seq * p.lse_stride + h0 + st] = lse;
        } else if (p.flat) {
          p.part_lse[part_row + st] = lse;
        } else {
          p.part_lse[((long long)g.seq * p.H + h0 + st) * p.max_splits + g.slot] = lse;
        }
      }
    }
    softmax_bar_sync<NS>();
    if (st < NG) misc->fac[st] = 1.f / misc->fac[st];
    softmax_bar_sync<NS>();
    // O^T -> global: one base pointer and one stride per mode, chosen once (the per-element form of this loop, with
    // its 64-bit index arithmetic and mode branches, took 50-90 k cycles per segment)
    {
      auto store_all = [&](auto put) {
        for (int j = 0; j < 4; ++j) {
          const int dv = j * 128 + row;
#pragma unroll
          for (int cl = 0; cl < CT; cl += 16) {
            const int c = cbase + cl;
            uint32_t o[16];
            tmem_ld<16>(tm_o + j * NO + c + lane_base, o);
            if constexpr (MT) {   // O^T = hi products + lo products
              uint32_t o2[16];
              tmem_ld<16>(tm_o + j * NO + NG + c + lane_base, o2);
              tmem_ld_wait();
#pragma unroll
              for (int e = 0; e < 16; ++e) o[e] = __float_as_uint(__uint_as_float(o[e]) + __uint_as_float(o2[e]));
            } else {
              tmem_ld_wait();
            }
#pragma unroll
            for (int e = 0; e < 16; ++e)
              if (c + e < nh) put(c + e, dv, __uint_as_float(o[e]) * misc->fac[c + e]);
          }
        }
      };
      if (direct) {
        const long long hs = p.out_head_stride;
        if (p.f16) {
          __half* ob = reinterpret_cast<__half*>(p.out) + (long long)g.seq * p.out_seq_stride + (long long)h0 * hs;
          store_all([&](int h, int dv, float v) { ob[h * hs + dv] = __float2half_rn(v); });
        } else {
          __nv_bfloat16* ob = reinterpret_cast<__nv_bfloat16*>(p.out) + (long long)g.seq * p.out_seq_stride + (long long)h0 * hs;
          store_all([&](int h, int dv, float v) { ob[h * hs + dv] = __float2bfloat16_rn(v); });
        }
      } else {
        float* pb = p.flat ? p.part_out + part_row * 512
                           : p.part_out + (((long long)g.seq * p.H + h0) * p.max_splits + g.slot) * 512;
        const long long hs = p.flat ? 512 : (long long)p.max_splits * 512;
        store_all([&](int h, int dv, float v) { pb[h * hs + dv] = v; });
      }
    }
    tc_fence_before();
    if (tid == 0) stamp(2, nseg, 3);
    mbar_arrive(smem_u32(&misc->o_free));   // O^T may be overwritten by the next segment
    softmax_bar_sync<NS>();                  // fac / mref / the scratch are reused by the next segment
    }
  }

  __syncthreads();
  if (p.cta_times != nullptr && tid == 0) {
    long long t1g;
    uint32_t smid;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1g));
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    long long* rec = p.cta_times + ((long long)blockIdx.y * gridDim.x + blockIdx.x) * 3;
    rec[0] = cta_t0, rec[1] = t1g, rec[2] = smid;
  }
  if (warp == kMmaWarp) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(kCols) : "memory");
  }
}

// flat mode: fold the partial slots of every item that more than one CTA worked on.  The decode kernel's CTA 0 left
// the per-sequence unit prefix (and the total) in the workspace header; CTA c's range is [c W / P, (c + 1) W / P).
__global__ void __launch_bounds__(128) mla_merge_flat_kernel(const MlaParams p, int P, int NG, int unit) {
  const int seq = blockIdx.y, h = blockIdx.x;
  const int L = p.seq_lens[seq];
  if (L <= 0) return;
  const int tiles = mla_units_of(L, unit);
  const int group = h / NG, hl = h % NG;
  const long long W = p.flat_prefix[p.S];
  if (p.flat_prefix[p.S + 1] != 0) return;   // item mode: every item was written directly
  const long long a = p.flat_prefix[seq] + (long long)group * tiles, b = a + tiles;
  auto cta_of = [&](long long u) { return (int)(((u + 1) * P + W - 1) / W - 1); };   // the c with c W / P <= u < (c + 1) W / P
  const int c_lo = cta_of(a), c_hi = cta_of(b - 1);
  if (c_lo == c_hi) return;   // one CTA had the whole item and wrote it directly
  float m = -INFINITY;
  for (int c = c_lo; c <= c_hi; ++c) {
    if (W * c / P == W * (c + 1) / P) continue;  // a CTA without units (more CTAs than units)
    const int slot = (W * c / P >= a) ? 0 : 1;   // the item continues from the CTA's range start (slot 0) or ends its range (slot 1)
    m = fmaxf(m, p.part_lse[((long long)c * 2 + slot) * NG + hl]);
  }
  float denom = 0.f;
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int c = c_lo; c <= c_hi; ++c) {
    if (W * c / P == W * (c + 1) / P) continue;
    const int slot = (W * c / P >= a) ? 0 : 1;
    const long long rowi = ((long long)c * 2 + slot) * NG + hl;
    const float w = __expf(p.part_lse[rowi] - m);
    denom += w;
    const float4 v = reinterpret_cast<const float4*>(p.part_out + rowi * 512)[threadIdx.x];
    acc.x += w * v.x, acc.y += w * v.y, acc.z += w * v.z, acc.w += w * v.w;
  }
  const float inv = 1.f / denom;
  acc.x *= inv, acc.y *= inv, acc.z *= inv, acc.w *= inv;
  const long long off = (long long)seq * p.out_seq_stride + (long long)h * p.out_head_stride + threadIdx.x * 4;
  if (p.f16) {
    __half2* o = reinterpret_cast<__half2*>(reinterpret_cast<__half*>(p.out) + off);
    o[0] = __floats2half2_rn(acc.x, acc.y);
    o[1] = __floats2half2_rn(acc.z, acc.w);
  } else {
    __nv_bfloat162* o = reinterpret_cast<__nv_bfloat162*>(reinterpret_cast<__nv_bfloat16*>(p.out) + off);
    o[0] = __floats2bfloat162_rn(acc.x, acc.y);
    o[1] = __floats2bfloat162_rn(acc.z, acc.w);
  }
  if (p.lse != nullptr && threadIdx.x == 0) p.lse[(long long)seq * p.lse_stride + h] = m + __logf(denom);
}

// =====================================================================================================================
// Pair kernel (65..128 heads): a cluster of two CTAs drives the tensor cores of both SMs with cta_group::2 instructions.
//   S^T: M = 256 = the 128-token tiles of BOTH CTAs (CTA r owns tile 2u + r of super-tile u), N = 128 heads; the B operand
//        Q^T is split by the hardware across the pair (CTA r keeps heads 64r..64r+63: 72 KiB each instead of 144).
//   O^T: M = 256 = two 128-row dv slices (CTA r accumulates dv rows 256j + 128r .. +127, j = 0, 1), N = 128 heads,
//        K = the tokens of tile X_t (t = 0, 1 in turn); B operand P^T[X_t] split by heads across the pair, so the
//        softmax threads of CTA t write the probabilities of heads 64..127 / 0..63 into the PEER's shared memory.
// One instruction therefore covers what four instructions of the single-CTA kernel do (two token tiles x two head
// groups): 68 tcgen05.mma per 256 tokens x 128 heads instead of 272, and each CTA streams 17 chunks per TWO tiles
// (9 K chunks of its own tile + its 4-chunk dv slice of V for both tiles).  TMEM per CTA: 2 x 128 (S^T) + 2 x 128 (O^T).
// The softmax reference m_ref[h] must be the same in both CTAs (their probabilities add up in one accumulator): every
// super-tile the CTAs exchange an "a logit passed the reference" flag (and, if set, their column maxima) through
// distributed shared memory and take the same decision.
constexpr int kPairSoftmaxWarps = 8;   // thread = (token row, head half): warps w and w + 4 share a TMEM lane quadrant
constexpr int kPairMmaWarp = 8;
constexpr int kPairProducers = 4;
constexpr int kPairThreads = (kPairSoftmaxWarps + 1 + kPairProducers) * 32;
constexpr int kPairSoftmaxThreads = kPairSoftmaxWarps * 32;

struct PairPost {           // what one CTA tells the pair about a super-tile
  int flag;                 // some logit of its tile passed m_ref + kTau
  int pad[3];
  int colmax[128];          // per-head maximum of its tile (sortable-int encoding), valid when flag != 0
};
struct PairMisc {
  uint64_t full[kMaxSlots / 2], empty[kMaxSlots / 2];
  uint64_t s_full[2], s_free[2], p_full, p_free, o_full, q_full, q_free, o_free;
  uint64_t post_bar[2];     // both CTAs' posts of a super-tile have landed (parity by super-tile)
  uint64_t l_bar[2];        // both CTAs' row sums of a segment have landed (parity by segment)
  uint32_t tmem_base;
  uint32_t pad[3];
  float mref[128];
  float fac[128];
  float lpart[2][2][128];   // [segment parity][CTA]: row sums of each CTA's tokens, exchanged at the end of a segment
  int colmax_w[kPairSoftmaxWarps][64];
  PairPost post[2][2];      // [super-tile parity][source CTA]
};
__host__ __device__ constexpr size_t pair_smem_bytes(int nslots) {
  return 1024 + (size_t)64 * 128 * kKChunks + 3 * 16384 + (size_t)nslots * kChunkBytes + sizeof(PairMisc);
}

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t mapa_u32(uint32_t addr, uint32_t rank) {   // same offset in CTA `rank` of the cluster
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {   // release at cluster scope, possibly remote
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx_cluster(uint32_t cluster_addr, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.release.cluster.shared::cluster.b64 _, [%0], %1;" ::"r"(cluster_addr), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_wait_cluster(uint32_t bar, uint32_t parity) {   // acquire at cluster scope
  asm volatile(
      "{\n.reg .pred p;\nW_%=:\nmbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%0], %1;\n@p bra D_%=;\nbra W_%=;\nD_%=:\n}\n" ::"r"(bar),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void st_cluster_u16(uint32_t cluster_addr, uint16_t v) {
  asm volatile("st.shared::cluster.u16 [%0], %1;" ::"r"(cluster_addr), "h"(v) : "memory");
}
__device__ __forceinline__ void st_cluster_u32(uint32_t cluster_addr, uint32_t v) {
  asm volatile("st.shared::cluster.u32 [%0], %1;" ::"r"(cluster_addr), "r"(v) : "memory");
}
__device__ __forceinline__ void umma_f16_2sm(uint32_t d_tmem, uint32_t a_lo, uint32_t a_hi, uint32_t b_lo, uint32_t b_hi,
                                             uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n.reg .pred p;\n.reg .b64 da, db;\nmov.b64 da, {%1, %2};\nmov.b64 db, {%3, %4};\nsetp.ne.b32 p, %6, 0;\n"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], da, db, %5, p;\n}\n" ::"r"(d_tmem),
      "r"(a_lo), "r"(a_hi), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrives (once all MMAs issued so far are complete) on the barrier at this offset in BOTH CTAs of the pair
__device__ __forceinline__ void umma_commit_2sm(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar),
               "h"((uint16_t)3)
               : "memory");
}
// tiled load whose completion bytes are counted on the LEADER CTA's barrier
__device__ __forceinline__ void tma_load_3d_2sm(uint32_t dst, const void* tmap, int c0, int c1, int c2, uint32_t leader_bar) {
  asm volatile(
      "cp.async.bulk.tensor.3d.cta_group::2.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
      ::"r"(dst), "l"(tmap), "r"(leader_bar), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kPairThreads, 1)
mla_decode_pair_kernel(const MlaParams p, const __grid_constant__ TmaDesc tm_kv, const __grid_constant__ TmaDesc tm_qn,
                       const __grid_constant__ TmaDesc tm_qp) {
  const uint32_t rank = cluster_ctarank();       // 0 = leader (issues the MMAs), 1 = peer
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  constexpr int kUnit = 2 * kTile;               // scheduling unit: one super-tile (a 128-token tile for each CTA)
  __shared__ MlaFlatStart fstart;
  const int npairs_grid = (int)gridDim.x >> 1, my_pair = (int)blockIdx.x >> 1;
  if (!p.flat) {
    const int seq = blockIdx.y, split = my_pair;
    const int L = p.seq_lens[seq];
    if (L <= 0) {  // empty sequence (both CTAs take this branch): zeros and lse = -inf
      if (split == 0) {
        for (int i = tid + rank * kPairThreads; i < p.H * 512; i += 2 * kPairThreads) {
          const long long o = (long long)seq * p.out_seq_stride + (long long)(i / 512) * p.out_head_stride + i % 512;
          if (p.f16) reinterpret_cast<__half*>(p.out)[o] = __float2half_rn(0.f);
          else reinterpret_cast<__nv_bfloat16*>(p.out)[o] = __float2bfloat16_rn(0.f);
        }
        if (p.lse != nullptr && rank == 0 && tid < p.H) p.lse[(long long)seq * p.lse_stride + tid] = -INFINITY;
      }
      return;
    }
    if (split * p.split_tokens >= L) return;
  } else {
    // both CTAs of a pair compute the same walk (worker = pair)
    mla_flat_prologue(p, fstart, kUnit, 1, npairs_grid, my_pair, 20, blockIdx.x == 0, tid, kPairThreads, 2, (int)rank);
    if (fstart.u0 >= fstart.u1) return;   // more pairs than units (both CTAs of the pair return)
  }
  auto first_seg = [&]() {
    if (p.flat) return mla_flat_first(p, fstart, kUnit);
    MlaSeg g;
    g.seq = blockIdx.y, g.group = 0;
    g.L = p.seq_lens[g.seq];
    g.tiles_seq = mla_units_of(g.L, kUnit);
    g.tile0 = my_pair * (p.split_tokens / kUnit);
    g.tile1 = min(g.tiles_seq, g.tile0 + p.split_tokens / kUnit);
    g.slot = my_pair;
    g.valid = g.tile0 < g.tile1;
    return g;
  };
  auto next_seg = [&](MlaSeg g, long long& done) {
    if (!p.flat) { g.valid = false; return g; }
    return mla_flat_next(p, fstart, kUnit, 1, npairs_grid, g, done);
  };

  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* const base_ptr = smem_raw + (base - smem_u32(smem_raw));
  const uint32_t q_u32 = base;                              // 9 chunks [64 heads x 128 B]: heads 64r .. 64r+63
  const uint32_t p_u32 = q_u32 + kKChunks * 64 * 128;       // P^T[X_0], P^T[X_1]: 2 x (2 chunks [64 heads x 128 B])
  const uint32_t stage_u32 = p_u32 + 2 * 16384;             // staging of the probabilities bound for the peer (16 KiB)
  const uint32_t ring_u32 = stage_u32 + 16384;
  uint8_t* const ring_ptr = base_ptr + (ring_u32 - base);
  PairMisc* const misc = reinterpret_cast<PairMisc*>(ring_ptr + (size_t)p.nslots * kChunkBytes);
  const int npairs = p.nslots >> 1;

  if (tid == 0) {
    for (int i = 0; i < npairs; ++i) {
      mbar_init(smem_u32(&misc->full[i]), 2 * 2);   // leader's copy is the one used: two producers in each CTA
      mbar_init(smem_u32(&misc->empty[i]), 1);      // one multicast tcgen05.commit
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(smem_u32(&misc->s_full[i]), 1);
      mbar_init(smem_u32(&misc->s_free[i]), 2 * kPairSoftmaxWarps);   // leader's copy: one arrival per softmax warp of the pair
      mbar_init(smem_u32(&misc->post_bar[i]), 2);                     // one post from each CTA
      mbar_init(smem_u32(&misc->l_bar[i]), 2);                        // both CTAs' row sums of a segment have landed
    }
    mbar_init(smem_u32(&misc->p_full), 2 * kPairSoftmaxWarps);        // leader's copy
    mbar_init(smem_u32(&misc->p_free), 1);
    mbar_init(smem_u32(&misc->o_full), 1);
    mbar_init(smem_u32(&misc->q_full), 2);                            // leader's copy: both CTAs' query loads
    mbar_init(smem_u32(&misc->q_free), 1);                            // multicast commit: the segment's last S^T has read Q
    mbar_init(smem_u32(&misc->o_free), 2 * kPairSoftmaxWarps);        // leader's copy: the segment's O^T has been read out
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == kPairMmaWarp) {   // the same warp in both CTAs, the same destination address
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&misc->tmem_base)),
                 "r"(512u)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem = misc->tmem_base;
  const uint32_t tm_s[2] = {tmem, tmem + 128};
  const uint32_t tm_o = tmem + 256;              // O^T blocks j = 0, 1: 128 columns each
  long long* const stamps = (p.debug_logits != nullptr && (int)blockIdx.x == 2 * p.stamp_cta && blockIdx.y == 0)
                                ? reinterpret_cast<long long*>(p.debug_logits + 128 * 64) : nullptr;
  auto stamp = [&](int role, int tile, int k) {
    if (stamps != nullptr && tile < 64) stamps[(role * 64 + tile) * 4 + k] = clock64();
  };
  // leader-CTA addresses of the barriers that gather arrivals from both CTAs
  const uint32_t lead_q_full = mapa_u32(smem_u32(&misc->q_full), 0);
  const uint32_t lead_p_full = mapa_u32(smem_u32(&misc->p_full), 0);
  const uint32_t lead_o_free = mapa_u32(smem_u32(&misc->o_free), 0);

  if (warp > kPairMmaWarp) {
    // =================== TMA producers (both CTAs): half-pair f of the ring sequence by producer f % kPairProducers ===================
    const int pw = warp - kPairMmaWarp - 1;
    if (lane == 0) {
      const int box_rows = p.box_rows;
      const uint32_t box_bytes = (uint32_t)box_rows * 128u;
      int pair = 0, turn = 0;
      uint32_t phase = 0;
      long long done = 0;
      int nseg = 0;
      for (MlaSeg g = first_seg(); g.valid; g = next_seg(g, done), ++nseg) {
        const int L = g.L, t0 = g.tile0 * kUnit, U = g.tile1 - g.tile0;
        if (pw == 0) {
          // this CTA's 64 query heads; bytes are counted on the leader's barrier
          if (nseg > 0) mbar_wait(smem_u32(&misc->q_free), (nseg - 1) & 1);
          mbar_expect_tx_cluster(lead_q_full, (uint32_t)(kKChunks * 64 * 128));
          for (int c = 0; c < kVChunks; ++c) tma_load_3d_2sm(q_u32 + c * 64 * 128, &tm_qn, c * 64, 64 * (int)rank, g.seq, lead_q_full);
          tma_load_3d_2sm(q_u32 + kVChunks * 64 * 128, &tm_qp, 0, 64 * (int)rank, g.seq, lead_q_full);
        }
        const int32_t* table = p.page_table + (long long)g.seq * p.page_table_stride;
        // `nhalves` ring halves holding chunks chunk0, chunk0 + 1, ... (at most nchunks of them) of token tile `tile`
        auto load_halves = [&](int tile, int chunk0, int nhalves, int nchunks) {
          const int tok0 = t0 + tile * kTile;
          const int valid = max(0, min(kTile, L - tok0));          // rows of this tile that hold cache entries
          const int vbox = (valid + box_rows - 1) / box_rows;
          int page[8], row[8];
#pragma unroll
          for (int b = 0; b < 8; ++b) {
            if (b < vbox) {
              const int tok = tok0 + b * box_rows;
              page[b] = table[tok >> p.page_shift];
              row[b] = tok & (p.page_size - 1);
            }
          }
          for (int c = 0; c < nhalves; ++c) {
            if (turn == pw) {
              mbar_wait(smem_u32(&misc->empty[pair]), phase ^ 1);
              const uint32_t full = mapa_u32(smem_u32(&misc->full[pair]), 0);
              if (c < nchunks) {
                mbar_expect_tx_cluster(full, box_bytes * (uint32_t)vbox);
                const uint32_t dst = ring_u32 + (uint32_t)(2 * pair + (c & 1)) * kChunkBytes;
#pragma unroll
                for (int b = 0; b < 8; ++b)
                  if (b < vbox) tma_load_3d_2sm(dst + b * box_bytes, &tm_kv, (chunk0 + c) * 64, row[b], page[b], full);
              } else {
                mbar_arrive_cluster(full);
              }
            }
            if (++turn == kPairProducers) turn = 0;
            if (c & 1) {
              if (++pair == npairs) { pair = 0; phase ^= 1; }
            }
          }
        };
        // S^T of super-tile u: the 9 K chunks (+ an unused half) of this CTA's own tile 2u + rank
        // O^T of super-tile u: for t = 0, 1 and j = 0, 1 the chunk pair (4j + 2 rank, +1) of tile 2u + t (this CTA's dv slice)
        load_halves((int)rank, 0, 10, kKChunks);
        for (int u = 0; u < U; ++u) {
          if (u + 1 < U) load_halves(2 * (u + 1) + (int)rank, 0, 10, kKChunks);
          for (int t = 0; t < 2; ++t)
            for (int j = 0; j < 2; ++j) load_halves(2 * u + t, 4 * j + 2 * (int)rank, 2, 2);
        }
      }
    }
  } else if (warp == kPairMmaWarp) {
    // =================== MMA issuer: one lane of the leader CTA drives both tensor cores ===================
    if (rank == 0 && lane == 0) {
      const int fmt = p.f16 ? 0 : 1;
      const uint32_t idesc_s = idesc_f16(256, 128, fmt, 0);
      const uint32_t idesc_o = idesc_f16(256, 128, fmt, 1);
      int pair = 0;
      uint32_t phase = 0;
      auto advance = [&]() {
        if (++pair == npairs) { pair = 0; phase ^= 1; }
      };
      const uint32_t q_lo = desc_lo(q_u32, 0);
      long long done = 0;
      int nseg = 0;
      int gu = 0;   // super-tile counter across segments
      for (MlaSeg g = first_seg(); g.valid; g = next_seg(g, done), ++nseg) {
        const int t0 = g.tile0 * kUnit, t1 = min(g.L, g.tile1 * kUnit), U = g.tile1 - g.tile0;
        const int gu0 = gu;
        mbar_wait_cluster(smem_u32(&misc->q_full), nseg & 1);
        auto issue_s = [&](int u) {
          const int x = gu0 + u, b = x & 1;
          if (x >= 2) mbar_wait_cluster(smem_u32(&misc->s_free[b]), ((x >> 1) & 1) ^ 1);
          tc_fence_after();
          stamp(0, x, 0);
          for (int c2 = 0; c2 < 5; ++c2) {
            mbar_wait_cluster(smem_u32(&misc->full[pair]), phase);
            tc_fence_after();
            const uint32_t a_lo = desc_lo(ring_u32 + (uint32_t)pair * 2 * kChunkBytes, 0);
            const uint32_t b_lo = q_lo + (uint32_t)(c2 * 2 * 64 * 128 >> 4);
#pragma unroll
            for (int h = 0; h < 2; ++h) {
              if (2 * c2 + h < kKChunks) {
#pragma unroll
                for (int k = 0; k < 4; ++k)
                  umma_f16_2sm(tm_s[b], a_lo + ((h * kChunkBytes + k * 32) >> 4), kDescHiSw128,
                               b_lo + ((h * 64 * 128 + k * 32) >> 4), kDescHiSw128, idesc_s, (c2 | h | k) != 0);
              }
            }
            umma_commit_2sm(smem_u32(&misc->empty[pair]));
            advance();
          }
          umma_commit_2sm(smem_u32(&misc->s_full[b]));
          if (u == U - 1) umma_commit_2sm(smem_u32(&misc->q_free));   // the query buffers may be reloaded once this is done
          stamp(0, x, 1);
        };
        auto issue_pv = [&](int u) {
          const int x = gu0 + u;
          mbar_wait_cluster(smem_u32(&misc->p_full), x & 1);
          if (u == 0 && nseg > 0) mbar_wait_cluster(smem_u32(&misc->o_free), (nseg - 1) & 1);  // previous O^T has been read out
          tc_fence_after();
          stamp(0, x, 2);
          for (int t = 0; t < 2; ++t) {
            const int rows = max(0, min(kTile, t1 - (t0 + (2 * u + t) * kTile)));   // valid token rows of tile X_t
            const int ksteps = (rows + 15) >> 4;
            const uint32_t p_lo = desc_lo(p_u32 + t * 16384, 0);
            for (int j = 0; j < 2; ++j) {
              mbar_wait_cluster(smem_u32(&misc->full[pair]), phase);
              tc_fence_after();
              if (rows < (ksteps << 4)) {
                // V rows past the sequence end may hold anything (0 * NaN = NaN): zero them in BOTH CTAs' stages
                for (uint32_t r2 = 0; r2 < 2; ++r2)
                  for (int hlf = 0; hlf < 2; ++hlf) {
                    const uint32_t stage = mapa_u32(ring_u32 + (uint32_t)(2 * pair + hlf) * kChunkBytes, r2);
                    for (int un = rows * 8; un < (ksteps << 4) * 8; ++un)
                      asm volatile("st.shared::cluster.v4.b32 [%0], {%1, %1, %1, %1};" ::"r"(stage + un * 16), "r"(0u) : "memory");
                  }
                asm volatile("fence.acq_rel.cluster;" ::: "memory");
                asm volatile("fence.proxy.async;" ::: "memory");
              }
              const uint32_t a_lo = desc_lo(ring_u32 + (uint32_t)pair * 2 * kChunkBytes, kChunkBytes);
              const uint32_t d_o = tm_o + j * 128;
#pragma unroll
              for (int ks = 0; ks < 8; ++ks)
                if (ks < ksteps)
                  umma_f16_2sm(d_o, a_lo + ((ks * 2048) >> 4), kDescHiSw128,
                               p_lo + (((ks >> 2) * 64 * 128 + (ks & 3) * 32) >> 4), kDescHiSw128, idesc_o, (u | t | ks) != 0);
              umma_commit_2sm(smem_u32(&misc->empty[pair]));
              advance();
            }
          }
          umma_commit_2sm(smem_u32(&misc->p_free));
          stamp(0, x, 3);
        };
        issue_s(0);
        for (int u = 0; u < U; ++u) {
          if (u + 1 < U) issue_s(u + 1);
          issue_pv(u);
        }
        umma_commit_2sm(smem_u32(&misc->o_full));
        gu += U;
      }
    }
    __syncwarp();
  } else {
    // =================== softmax / epilogue warps 0..7: thread = (token row of this CTA's tile, head half) ===================
    const int quad = warp & 3, hh = warp >> 2;
    const int row = quad * 32 + lane;
    const uint32_t lane_base = (uint32_t)(quad * 32) << 16;
    const int cbase = 64 * hh;                   // first S^T column (head) of this thread
    auto sm_bar = [&]() { asm volatile("bar.sync 1, 256;" ::: "memory"); };
    // P^T[X_rank] lives in CTA hh (the CTA that supplies heads 64hh.. as its half of the B operand)
    const uint32_t unit = (row & 63) >> 3, within = (row & 7) * 2;
    const uint32_t peer = rank ^ 1;
    // The CTAs agree on the references through "posts": after looking at its S^T tile, each CTA tells both CTAs whether a
    // logit passed m_ref + kTau and, if so, its column maxima.  The post of super-tile u is consumed at the START of
    // super-tile u + 1 (so the exchange is off the critical path); only the first super-tile of a segment, where no
    // reference exists yet, waits for it.  A logit may therefore exceed its reference by more than kTau for one
    // super-tile: harmless in bf16 / fp32 unless it jumps by more than ~2^100 within 256 tokens (the exponent is clamped).
    const bool remote_half = (uint32_t)hh != rank;
    long long done = 0;
    int nseg = 0;
    int gu = 0;   // super-tile counter across segments
    for (MlaSeg g = first_seg(); g.valid; g = next_seg(g, done), ++nseg) {
    const int t0 = g.tile0 * kUnit, t1 = min(g.L, g.tile1 * kUnit), U = g.tile1 - g.tile0;
    float lsum[64];
#pragma unroll
    for (int h = 0; h < 64; ++h) lsum[h] = 0.f;
    if (tid < 128) misc->mref[tid] = -INFINITY;
    sm_bar();

    for (int u = 0; u < U; ++u, ++gu) {
      const int b = gu & 1;
      // ---- references: posts of the previous super-tile of this segment (both CTAs take the same decision) ----
      if (u > 0) {
        const int pb = (gu - 1) & 1;
        mbar_wait_cluster(smem_u32(&misc->post_bar[pb]), ((gu - 1) >> 1) & 1);
        const int f0 = misc->post[pb][0].flag, f1 = misc->post[pb][1].flag;
        if (u > 1 && (f0 | f1)) {   // (the posts of the segment's first super-tile were applied synchronously below)
          if (tid < 128) {
            float m_new = misc->mref[tid];
            const float m_old = m_new;
            if (f0) m_new = fmaxf(m_new, sortable_f32(misc->post[pb][0].colmax[tid]));
            if (f1) m_new = fmaxf(m_new, sortable_f32(misc->post[pb][1].colmax[tid]));
            misc->fac[tid] = exp2f(m_old - m_new);
            misc->mref[tid] = m_new;
          }
          sm_bar();
#pragma unroll
          for (int h = 0; h < 64; ++h) lsum[h] *= misc->fac[cbase + h];
          mbar_wait(smem_u32(&misc->p_free), (gu - 1) & 1);  // the previous super-tile's P.V has landed in O^T
          tc_fence_after();
          for (int j = 0; j < 2; ++j) {
#pragma unroll
            for (int c = 0; c < 64; c += 16) {
              uint32_t o[16];
              tmem_ld<16>(tm_o + j * 128 + cbase + c + lane_base, o);
              tmem_ld_wait();
#pragma unroll
              for (int e = 0; e < 16; ++e) o[e] = __float_as_uint(__uint_as_float(o[e]) * misc->fac[cbase + c + e]);
              tmem_st16(tm_o + j * 128 + cbase + c + lane_base, o);
            }
          }
          tmem_st_wait();
          tc_fence_before();
          sm_bar();   // fac / mref are rewritten by the next rare path only after everyone is done with them
        }
      }
      mbar_wait(smem_u32(&misc->s_full[b]), (gu >> 1) & 1);
      tc_fence_after();
      const bool valid = t0 + (2 * u + (int)rank) * kTile + row < t1;
      if (tid == 0) stamp(1, gu, 0);
      // ---- pass 1: does a logit of this tile pass its reference?  Post the answer (and the column maxima) ----
      bool over = false;
#pragma unroll
      for (int c0 = 0; c0 < 64; c0 += 32) {
        uint32_t r[32];
        tmem_ld<32>(tm_s[b] + cbase + c0 + lane_base, r);
        tmem_ld_wait();
#pragma unroll
        for (int h = 0; h < 32; ++h)
          over |= valid && __uint_as_float(r[h]) * p.scale_log2 > misc->mref[cbase + c0 + h] + kTau;
      }
      uint32_t over_cta;
      asm volatile("{\n.reg .pred p, q;\nsetp.ne.b32 q, %1, 0;\nbar.red.or.pred p, 1, 256, q;\nselp.u32 %0, 1, 0, p;\n}\n"
                   : "=r"(over_cta) : "r"((uint32_t)over) : "memory");
      PairPost* const mine = &misc->post[b][rank];
      if (over_cta) {
#pragma unroll
        for (int c0 = 0; c0 < 64; c0 += 32) {
          uint32_t r[32];
          tmem_ld<32>(tm_s[b] + cbase + c0 + lane_base, r);
          tmem_ld_wait();
#pragma unroll
          for (int h = 0; h < 32; ++h) {
            const float x = valid ? __uint_as_float(r[h]) * p.scale_log2 : -INFINITY;
            const int m = __reduce_max_sync(0xffffffffu, f32_sortable(x));
            if (lane == 0) misc->colmax_w[warp][c0 + h] = m;
          }
        }
        sm_bar();
        if (tid < 128) {
          const int w0 = (tid >> 6) * 4, c = tid & 63;
          const int m = max(max(misc->colmax_w[w0][c], misc->colmax_w[w0 + 1][c]), max(misc->colmax_w[w0 + 2][c], misc->colmax_w[w0 + 3][c]));
          mine->colmax[tid] = m;
          st_cluster_u32(mapa_u32(smem_u32(&mine->colmax[tid]), peer), (uint32_t)m);
        }
        sm_bar();
      }
      if (tid == 0) {
        mine->flag = (int)over_cta;
        st_cluster_u32(mapa_u32(smem_u32(&mine->flag), peer), over_cta);
        mbar_arrive_cluster(mapa_u32(smem_u32(&misc->post_bar[b]), rank));
        mbar_arrive_cluster(mapa_u32(smem_u32(&misc->post_bar[b]), peer));
      }
      if (u == 0) {   // no reference yet: this one exchange is synchronous (O^T and the sums are still empty)
        mbar_wait_cluster(smem_u32(&misc->post_bar[b]), (gu >> 1) & 1);
        const int f0 = misc->post[b][0].flag, f1 = misc->post[b][1].flag;
        if (tid < 128) {
          float m_new = -INFINITY;
          if (f0) m_new = fmaxf(m_new, sortable_f32(misc->post[b][0].colmax[tid]));
          if (f1) m_new = fmaxf(m_new, sortable_f32(misc->post[b][1].colmax[tid]));
          if (m_new == -INFINITY) m_new = 0.f;
          misc->mref[tid] = m_new;
        }
        sm_bar();
      }
      if (tid == 0) stamp(1, gu, 1);
      // ---- pass 2: probabilities of this thread's 64 heads ----
      const uint32_t wr_base = remote_half ? stage_u32 + (row >> 6) * 64 * 128 : p_u32 + rank * 16384 + (row >> 6) * 64 * 128;
#pragma unroll
      for (int c0 = 0; c0 < 64; c0 += 32) {
        uint32_t r[32];
        tmem_ld<32>(tm_s[b] + cbase + c0 + lane_base, r);
        tmem_ld_wait();
        if (c0 == 32) {   // last read of this S^T buffer: one arrival per warp on the leader's barrier
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_cluster(mapa_u32(smem_u32(&misc->s_free[b]), 0));
        }
#pragma unroll
        for (int h = 0; h < 32; ++h) {
          const float x = valid ? __uint_as_float(r[h]) * p.scale_log2 : -INFINITY;
          const float pv = fast_exp2(fminf(x - misc->mref[cbase + c0 + h], 120.f));
          lsum[c0 + h] += pv;
          r[h] = p.f16 ? (uint32_t)__half_as_ushort(__float2half_rn(pv)) : (uint32_t)__bfloat16_as_ushort(__float2bfloat16_rn(pv));
        }
        // the local half goes straight into this CTA's P^T[X_rank] (once the previous P.V released it); the half bound
        // for the peer is staged locally first
        if (c0 == 0 && gu > 0 && !remote_half) mbar_wait(smem_u32(&misc->p_free), (gu - 1) & 1);
        if (c0 == 0 && tid == 0) stamp(1, gu, 2);
#pragma unroll
        for (int h = 0; h < 32; ++h) {
          const uint32_t addr = wr_base + (c0 + h) * 128 + ((unit ^ (h & 7)) << 4) + within;
          asm volatile("st.shared.u16 [%0], %1;" ::"r"(addr), "h"((uint16_t)r[h]) : "memory");
        }
      }
      if (remote_half) {
        // this warp's 32 token rows are 4 sixteen-byte units per head: ship them to the peer's P^T[X_rank], 8 units per lane
        __syncwarp();
        if (gu > 0) mbar_wait(smem_u32(&misc->p_free), (gu - 1) & 1);
        const uint32_t src0 = stage_u32 + (quad >> 1) * 64 * 128;
        const uint32_t dst0 = mapa_u32(p_u32 + rank * 16384 + (quad >> 1) * 64 * 128, peer);
        const uint32_t ubase = (quad & 1) * 4;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const int idx = k * 32 + lane, h = idx >> 2;
          const uint32_t off = h * 128 + (((ubase + (idx & 3)) ^ (h & 7)) << 4);
          uint32_t v0, v1, v2, v3;
          asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v0), "=r"(v1), "=r"(v2), "=r"(v3) : "r"(src0 + off) : "memory");
          asm volatile("st.shared::cluster.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst0 + off), "r"(v0), "r"(v1), "r"(v2), "r"(v3) : "memory");
        }
      }
      asm volatile("fence.proxy.async;" ::: "memory");
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(lead_p_full);
      if (tid == 0) stamp(1, gu, 3);
    }

    // ---- epilogue of the segment: row sums of both CTAs' tokens, normalise this CTA's O^T rows, store ----
    mbar_wait(smem_u32(&misc->o_full), nseg & 1);   // every MMA of the segment is complete
    tc_fence_after();
    {
      float* wsum = reinterpret_cast<float*>(&misc->colmax_w[0][0]);   // [8 warps][64]: per-warp column sums
#pragma unroll
      for (int h = 0; h < 64; ++h) {
        float v = lsum[h];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0) wsum[warp * 64 + h] = v;
      }
      sm_bar();
      const int lb = nseg & 1;
      if (tid < 128) {
        const int w0 = (tid >> 6) * 4, c = tid & 63;
        const float l = wsum[w0 * 64 + c] + wsum[(w0 + 1) * 64 + c] + wsum[(w0 + 2) * 64 + c] + wsum[(w0 + 3) * 64 + c];
        misc->lpart[lb][rank][tid] = l;
        st_cluster_u32(mapa_u32(smem_u32(&misc->lpart[lb][rank][tid]), peer), __float_as_uint(l));
      }
      sm_bar();
      if (tid == 0) {
        mbar_arrive_cluster(mapa_u32(smem_u32(&misc->l_bar[lb]), rank));
        mbar_arrive_cluster(mapa_u32(smem_u32(&misc->l_bar[lb]), peer));
      }
      mbar_wait_cluster(smem_u32(&misc->l_bar[lb]), (nseg >> 1) & 1);
    }
    const bool direct = g.tile0 == 0 && g.tile1 == g.tiles_seq;
    const long long part_row = p.flat ? ((long long)my_pair * 2 + g.slot) * 128 : 0;
    if (tid < 128) {
      const float l = misc->lpart[nseg & 1][0][tid] + misc->lpart[nseg & 1][1][tid];
      misc->fac[tid] = 1.f / l;
      if (rank == 0 && tid < p.H) {
        const float lse = (misc->mref[tid] + log2f(l)) * kLn2;
        if (direct) {
          if (p.lse != nullptr) p.lse[(long long)g.seq * p.lse_stride + tid] = lse;
        } else if (p.flat) {
          p.part_lse[part_row + tid] = lse;
        } else {
          p.part_lse[((long long)g.seq * p.H + tid) * p.max_splits + g.slot] = lse;
        }
      }
    }
    sm_bar();
    {
      auto store_all = [&](auto put) {
        for (int j = 0; j < 2; ++j) {
          const int dv = 256 * j + 128 * (int)rank + row;
#pragma unroll
          for (int c = 0; c < 64; c += 16) {
            uint32_t o[16];
            tmem_ld<16>(tm_o + j * 128 + cbase + c + lane_base, o);
            tmem_ld_wait();
#pragma unroll
            for (int e = 0; e < 16; ++e)
              if (cbase + c + e < p.H) put(cbase + c + e, dv, __uint_as_float(o[e]) * misc->fac[cbase + c + e]);
          }
        }
      };
      if (direct) {
        const long long hs = p.out_head_stride;
        if (p.f16) {
          __half* ob = reinterpret_cast<__half*>(p.out) + (long long)g.seq * p.out_seq_stride;
          store_all([&](int h, int dv, float v) { ob[h * hs + dv] = __float2half_rn(v); });
        } else {
          __nv_bfloat16* ob = reinterpret_cast<__nv_bfloat16*>(p.out) + (long long)g.seq * p.out_seq_stride;
          store_all([&](int h, int dv, float v) { ob[h * hs + dv] = __float2bfloat16_rn(v); });
        }
      } else {
        float* pb = p.flat ? p.part_out + part_row * 512 : p.part_out + ((long long)g.seq * p.H * p.max_splits + g.slot) * 512;
        const long long hs = p.flat ? 512 : (long long)p.max_splits * 512;
        store_all([&](int h, int dv, float v) { pb[h * hs + dv] = v; });
      }
    }
    tc_fence_before();
    __syncwarp();
    if (lane == 0) mbar_arrive_cluster(lead_o_free);   // O^T may be overwritten by the next segment
    sm_bar();                                           // fac / mref / the scratch are reused by the next segment
    }
  }

  // no CTA leaves (or frees TMEM) while its peer may still touch its shared memory or issue pair MMAs
  cluster_sync_all();
  if (warp == kPairMmaWarp) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
  }
}

// ---- merge of the token splits (flash-decoding reduce): out = sum_j w_j O_j, w_j = exp(lse_j - max) / sum ----
__global__ void __launch_bounds__(128) mla_merge_kernel(const MlaParams p) {
  const int seq = blockIdx.y, h = blockIdx.x;
  const int L = p.seq_lens[seq];
  const int n = L > 0 ? (L + p.split_tokens - 1) / p.split_tokens : 1;
  if (n <= 1) return;  // written directly by the decode kernel
  const float* lse = p.part_lse + ((long long)seq * p.H + h) * p.max_splits;
  float m = -INFINITY;
  for (int j = 0; j < n; ++j) m = fmaxf(m, lse[j]);
  float denom = 0.f;
  for (int j = 0; j < n; ++j) denom += __expf(lse[j] - m);
  const float inv = 1.f / denom;
  const float4* po = reinterpret_cast<const float4*>(p.part_out + ((long long)seq * p.H + h) * p.max_splits * 512);
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int j = 0; j < n; ++j) {
    const float w = __expf(lse[j] - m) * inv;
    const float4 v = po[j * 128 + threadIdx.x];
    acc.x += w * v.x, acc.y += w * v.y, acc.z += w * v.z, acc.w += w * v.w;
  }
  const long long off = (long long)seq * p.out_seq_stride + (long long)h * p.out_head_stride + threadIdx.x * 4;
  if (p.f16) {
    __half2* o = reinterpret_cast<__half2*>(reinterpret_cast<__half*>(p.out) + off);
    o[0] = __floats2half2_rn(acc.x, acc.y);
    o[1] = __floats2half2_rn(acc.z, acc.w);
  } else {
    __nv_bfloat162* o = reinterpret_cast<__nv_bfloat162*>(reinterpret_cast<__nv_bfloat16*>(p.out) + off);
    o[0] = __floats2bfloat162_rn(acc.x, acc.y);
    o[1] = __floats2bfloat162_rn(acc.z, acc.w);
  }
  if (p.lse != nullptr && threadIdx.x == 0) p.lse[(long long)seq * p.lse_stride + h] = m + __logf(denom);
}

// ---- host side ---------------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn mla_encode_fn() {
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

// 3-D map (inner dim, rows, outer) with a [64 x box_rows x 1] box and the 128-byte swizzle; strides in elements
int encode_map3(TmaDesc* out, const void* ptr, int etype /* 0 bf16, 1 f16, 2 bytes (e4m3) */, uint64_t d0, uint64_t d1,
                uint64_t d2, int64_t s1, int64_t s2, uint32_t box_rows, const char* what) {
  EncodeTiledFn enc = mla_encode_fn();
  B200_REQUIRE(enc != nullptr, B200_ERR_CUDA, "mla_decode: cuTensorMapEncodeTiled is not available from the driver");
  const int eb = etype == 2 ? 1 : 2;
  const cuuint64_t dims[3] = {d0, d1, d2};
  const cuuint64_t strides[2] = {(cuuint64_t)s1 * eb, (cuuint64_t)s2 * eb};
  const cuuint32_t box[3] = {(cuuint32_t)(128 / eb), box_rows, 1};   // 128-byte rows (a box may reach past dim 0: zero fill)
  const cuuint32_t estr[3] = {1, 1, 1};
  CUtensorMap m;
  const CUtensorMapDataType dt = etype == 2 ? CU_TENSOR_MAP_DATA_TYPE_UINT8
                                            : (etype == 1 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16);
  const CUresult r = enc(&m, dt, 3, const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                         CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  B200_REQUIRE(r == CUDA_SUCCESS, B200_ERR_INVALID_ARGUMENT,
               "mla_decode: cuTensorMapEncodeTiled failed (%d) for %s (dims %llu/%llu/%llu, strides %lld/%lld elements)",
               (int)r, what, (unsigned long long)d0, (unsigned long long)d1, (unsigned long long)d2, (long long)s1,
               (long long)s2);
  static_assert(sizeof(CUtensorMap) == sizeof(TmaDesc), "CUtensorMap is 128 bytes");
  memcpy(out, &m, sizeof(m));
  return B200_OK;
}

int mla_group_width(int H) { return H <= 16 ? 16 : (H <= 32 ? 32 : 64); }

// flat (persistent, tile-balanced) schedule of the single-CTA kernel: whenever the caller leaves the split count to us
bool mla_use_flat(int H, int requested_splits) {
  static const int off = env_int("B200_MLA_FLAT_OFF");
  (void)H;
  return !off && requested_splits <= 0;
}
bool mla_use_pair(int H);
thread_local bool g_mla_fp8 = false;   // the call being planned takes e4m3 inputs (no pair kernel for those)
// workers (CTAs, or CTA pairs) x 2 partial slots x slot width (the group's heads; all 128 for the pair kernel)
uint64_t mla_flat_workspace_bytes(int S, int H) {
  const bool pair = mla_use_pair(H);
  const int width = pair ? 128 : mla_group_width(H);
  const int workers = pair ? device_sm_count() / 2 : device_sm_count();
  const uint64_t header = ((uint64_t)(S + 2) * sizeof(long long) + 255) / 256 * 256;
  return header + (uint64_t)workers * 2 * width * (512 + 1) * sizeof(float);
}

bool mla_use_pair(int H) {   // 65..128 heads: the two-CTA kernel (B200_MLA_PAIR_OFF=1 keeps two independent 64-head groups)
  static const int off = env_int("B200_MLA_PAIR_OFF");
  return H > 64 && !off && !g_mla_fp8;
}

// number of token splits: minimise waves(items) / splits over the allowed range (every split >= kMinSplitTiles tiles)
int mla_choose_splits(int S, int H, long long max_seq_len, int requested) {
  const bool pair = mla_use_pair(H);
  const int unit = pair ? 2 * kTile : kTile;                        // tokens per scheduling unit (tile / super-tile)
  const long long tiles = (max_seq_len + unit - 1) / unit;
  int cap = (int)min((long long)kMaxSplits, max(1LL, tiles / (pair ? kMinSplitTiles / 2 : kMinSplitTiles)));
  if (requested == 1) return 1;
  if (requested > 1) return (int)min((long long)min(requested, kMaxSplits), max(1LL, tiles));
  const int NG = mla_group_width(H);
  const long long base = pair ? (long long)S : (long long)S * ((H + NG - 1) / NG);
  int sms = device_sm_count();  // resident CTAs (or CTA pairs)
  if (pair) sms /= 2;
  int best = 1;
  double best_cost = 1e30;
  for (int s = 1; s <= cap; ++s) {
    const long long per = (tiles + s - 1) / s;                       // units per split
    const long long waves = (base * s + sms - 1) / sms;
    // measured on B200 (profiles/r2_mla_*): a CTA costs ~4 tile-times beyond its tiles (query load, pipeline fill,
    // epilogue), and every extra split adds fp32 partials for the merge kernel to re-read
    const double cost = (double)waves * ((double)per + (pair ? 2.0 : 4.0)) + (s > 1 ? 1.0 + 0.25 * s : 0.0);
    if (cost < best_cost - 1e-9) best_cost = cost, best = s;
  }
  return best;
}

int launch_mla_pair(const MlaParams& mp, const TmaDesc& kv, const TmaDesc& qn, const TmaDesc& qp, dim3 grid, cudaStream_t s) {
  const size_t smem = pair_smem_bytes(mp.nslots);
  static thread_local int configured_device = -1;
  int dev = 0;
  cudaGetDevice(&dev);
  if (configured_device != dev) {
    cudaError_t e = cudaFuncSetAttribute(mla_decode_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448 - 704);
    if (e != cudaSuccess) {
      set_error("mla_decode: cudaFuncSetAttribute(smem, pair kernel) failed: %s", cudaGetErrorString(e));
      return B200_ERR_CUDA;
    }
    configured_device = dev;
  }
  mla_decode_pair_kernel<<<grid, kPairThreads, smem, s>>>(mp, kv, qn, qp);
  count_launch();
  return check_launch("mla_decode (pair kernel)");
}

template <int NG, bool FP8>
int launch_mla(const MlaParams& mp, const TmaDesc& kv, const TmaDesc& qn, const TmaDesc& qp, dim3 grid, cudaStream_t s) {
  auto kern = mla_decode_kernel<NG, FP8>;
  const size_t smem = mla_smem_bytes(NG, mp.nslots, FP8 ? 5 : kKChunks);
  static thread_local int configured_device = -1;
  int dev = 0;
  cudaGetDevice(&dev);
  if (configured_device != dev) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448 - 704);
    if (e != cudaSuccess) {
      set_error("mla_decode: cudaFuncSetAttribute(smem) failed: %s", cudaGetErrorString(e));
      return B200_ERR_CUDA;
    }
    configured_device = dev;
  }
  kern<<<grid, mla_threads(NG), smem, s>>>(mp, kv, qn, qp);
  count_launch();
  return check_launch("mla_decode");
}
}  // namespace
}  // namespace b200

using namespace b200;

extern "C" uint64_t b200_mla_decode_workspace_bytes(int32_t num_seqs, int32_t num_heads, int64_t max_seq_len,
                                                    int32_t num_kv_splits) {
  if (num_seqs <= 0 || num_heads <= 0 || max_seq_len <= 0) return 0;
  if (mla_use_flat(num_heads, num_kv_splits)) return mla_flat_workspace_bytes(num_seqs, num_heads);
  const int splits = mla_choose_splits(num_seqs, num_heads, max_seq_len, num_kv_splits);
  if (splits <= 1) return 0;
  return (uint64_t)num_seqs * num_heads * splits * (512 + 1) * sizeof(float);
}

extern "C" int b200_mla_decode(const b200_mla_decode_params_t* pp, void* stream) {
  B200_REQUIRE(pp != nullptr, B200_ERR_INVALID_ARGUMENT, "mla_decode: null params");
  const b200_mla_decode_params_t& p = *pp;
  B200_REQUIRE(p.out && p.q_nope && p.q_pe && p.kv_cache && p.seq_lens && p.page_table, B200_ERR_INVALID_ARGUMENT,
               "mla_decode: null tensor pointer");
  B200_REQUIRE(p.num_seqs >= 0 && p.num_heads > 0 && p.max_pages > 0 && p.num_blocks > 0, B200_ERR_INVALID_ARGUMENT,
               "mla_decode: negative or zero size");
  B200_REQUIRE(p.dtype == B200_DTYPE_BF16 || p.dtype == B200_DTYPE_F16, B200_ERR_UNSUPPORTED,
               "mla_decode: out must be bf16 or f16 (dtype %d)", p.dtype);
  B200_REQUIRE(p.kv_dtype == B200_KV_AUTO || p.kv_dtype == B200_KV_FP8_E4M3, B200_ERR_UNSUPPORTED,
               "mla_decode: q / cache are either of the out dtype (kv_dtype auto) or e4m3 bytes (fp8_e4m3), got %d", p.kv_dtype);
  B200_REQUIRE(p.num_heads <= 128, B200_ERR_UNSUPPORTED, "mla_decode: num_heads %d > 128", p.num_heads);
  B200_REQUIRE(p.page_size >= 16 && (p.page_size & (p.page_size - 1)) == 0, B200_ERR_UNSUPPORTED,
               "mla_decode: page_size %d must be a power of two >= 16", p.page_size);
  const bool fp8 = p.kv_dtype == B200_KV_FP8_E4M3;
  const int al = fp8 ? 16 : 8;   // elements per 16 bytes
  B200_REQUIRE(p.entry_stride % al == 0 && p.block_stride % al == 0 && p.q_nope_head_stride % al == 0 &&
                   p.q_nope_seq_stride % al == 0 && p.q_pe_head_stride % al == 0 && p.q_pe_seq_stride % al == 0 &&
                   p.out_head_stride % 4 == 0 && p.out_seq_stride % 4 == 0,
               B200_ERR_UNSUPPORTED, "mla_decode: rows must be 16-byte aligned");
  g_mla_fp8 = fp8;
  B200_REQUIRE(((uintptr_t)p.kv_cache | (uintptr_t)p.q_nope | (uintptr_t)p.q_pe) % 16 == 0 && (uintptr_t)p.out % 8 == 0,
               B200_ERR_UNSUPPORTED, "mla_decode: base pointers must be 16-byte aligned");
  if (p.num_seqs == 0) return B200_OK;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const bool f16 = p.dtype == B200_DTYPE_F16;
  const long long max_len = (long long)p.max_pages * p.page_size;
  const bool flat = mla_use_flat(p.num_heads, p.num_kv_splits);
  const int splits = flat ? 1 : mla_choose_splits(p.num_seqs, p.num_heads, max_len, p.num_kv_splits);
  const int NG = mla_group_width(p.num_heads);

  MlaParams mp;
  memset(&mp, 0, sizeof(mp));
  mp.out = p.out, mp.lse = p.lse, mp.seq_lens = p.seq_lens, mp.page_table = p.page_table;
  mp.debug_logits = p.debug_logits;
  mp.stamp_cta = env_int("B200_MLA_STAMP_CTA");
  if (p.debug_logits != nullptr && env_int("B200_MLA_CTA_TIMES"))   // tools/mla_debug.py --cta-times: buffer laid out after the stamps
    mp.cta_times = reinterpret_cast<long long*>(p.debug_logits + 128 * 64 + 3 * 64 * 4 * 2);
  mp.out_seq_stride = p.out_seq_stride, mp.out_head_stride = p.out_head_stride;
  mp.page_table_stride = p.page_table_stride, mp.lse_stride = p.lse_stride;
  mp.scale_log2 = p.scale * kLog2e;
  mp.S = p.num_seqs, mp.H = p.num_heads, mp.groups = (p.num_heads + NG - 1) / NG;
  mp.page_size = p.page_size;
  for (mp.page_shift = 0; (1 << mp.page_shift) < p.page_size; ++mp.page_shift) {}
  mp.box_rows = p.page_size < kTile ? p.page_size : kTile;
  mp.max_splits = splits;
  const bool pair = mla_use_pair(p.num_heads);
  const int unit = pair ? 2 * kTile : kTile;
  const long long tiles = (max_len + unit - 1) / unit;
  mp.split_tokens = (int)((tiles + splits - 1) / splits) * unit;
  mp.f16 = f16 ? 1 : 0;
  // shared memory: one CTA per SM (two CTAs per SM for the narrow groups measured slower: less in flight per CTA)
  const size_t budget = 232448 - 704;  // static shared memory: the flat schedule start record
  int nslots = (int)((budget - 1024 - (size_t)NG * 128 * ((fp8 ? 5 : kKChunks) + 2) - sizeof(MlaMisc)) / kChunkBytes);
  nslots = min(nslots & ~1, kMaxSlots);
  if (pair) nslots = (int)((232448 - 704 - 1024 - (size_t)64 * 128 * kKChunks - 3 * 16384 - sizeof(PairMisc)) / kChunkBytes) & ~1;
  mp.nslots = nslots;
  if (flat) {
    const uint64_t need = mla_flat_workspace_bytes(p.num_seqs, p.num_heads);
    B200_REQUIRE(p.workspace != nullptr && p.workspace_bytes >= need, B200_ERR_WORKSPACE,
                 "mla_decode: the flat schedule needs a workspace of %llu bytes (got %llu)", (unsigned long long)need,
                 (unsigned long long)p.workspace_bytes);
    B200_REQUIRE((uintptr_t)p.workspace % 16 == 0, B200_ERR_WORKSPACE, "mla_decode: workspace must be 16-byte aligned");
    const uint64_t header = ((uint64_t)(p.num_seqs + 2) * sizeof(long long) + 255) / 256 * 256;
    mp.flat = 1;
    mp.flat_prefix = static_cast<long long*>(p.workspace);
    mp.part_out = reinterpret_cast<float*>(static_cast<char*>(p.workspace) + header);
    mp.part_lse = mp.part_out + (pair ? (size_t)(device_sm_count() / 2) * 2 * 128 : (size_t)device_sm_count() * 2 * NG) * 512;
  } else if (splits > 1) {
    const uint64_t need = (uint64_t)p.num_seqs * p.num_heads * splits * (512 + 1) * sizeof(float);
    B200_REQUIRE(p.workspace != nullptr && p.workspace_bytes >= need, B200_ERR_WORKSPACE,
                 "mla_decode: %d token splits need a workspace of %llu bytes (got %llu)", splits, (unsigned long long)need,
                 (unsigned long long)p.workspace_bytes);
    B200_REQUIRE((uintptr_t)p.workspace % 16 == 0, B200_ERR_WORKSPACE, "mla_decode: workspace must be 16-byte aligned");
    mp.part_out = static_cast<float*>(p.workspace);
    mp.part_lse = mp.part_out + (size_t)p.num_seqs * p.num_heads * splits * 512;
  }
  TmaDesc kv, qn, qp;
  const int etype = fp8 ? 2 : (f16 ? 1 : 0);
  int rc = encode_map3(&kv, p.kv_cache, etype, 576, (uint64_t)p.page_size, (uint64_t)p.num_blocks, p.entry_stride, p.block_stride,
                       (uint32_t)mp.box_rows, "kv_cache");
  if (rc != B200_OK) return rc;
  rc = encode_map3(&qn, p.q_nope, etype, 512, (uint64_t)p.num_heads, (uint64_t)p.num_seqs, p.q_nope_head_stride,
                   p.q_nope_seq_stride, (uint32_t)NG, "q_nope");
  if (rc != B200_OK) return rc;
  rc = encode_map3(&qp, p.q_pe, etype, 64, (uint64_t)p.num_heads, (uint64_t)p.num_seqs, p.q_pe_head_stride, p.q_pe_seq_stride,
                   (uint32_t)NG, "q_pe");
  if (rc != B200_OK) return rc;
  const dim3 grid = flat ? dim3((unsigned)(pair ? device_sm_count() / 2 * 2 : device_sm_count()), 1)
                         : dim3((unsigned)((pair ? 2 : mp.groups) * splits), (unsigned)p.num_seqs);
  if (pair) rc = launch_mla_pair(mp, kv, qn, qp, grid, s);
  else if (fp8) rc = NG == 16 ? launch_mla<16, true>(mp, kv, qn, qp, grid, s)
                              : (NG == 32 ? launch_mla<32, true>(mp, kv, qn, qp, grid, s) : launch_mla<64, true>(mp, kv, qn, qp, grid, s));
  else if (NG == 16) rc = launch_mla<16, false>(mp, kv, qn, qp, grid, s);
  else if (NG == 32) rc = launch_mla<32, false>(mp, kv, qn, qp, grid, s);
  else rc = launch_mla<64, false>(mp, kv, qn, qp, grid, s);
  if (rc != B200_OK) return rc;
  if (flat) {
    if (pair) mla_merge_flat_kernel<<<dim3((unsigned)p.num_heads, (unsigned)p.num_seqs), 128, 0, s>>>(mp, (int)grid.x / 2, 128, 2 * kTile);
    else mla_merge_flat_kernel<<<dim3((unsigned)p.num_heads, (unsigned)p.num_seqs), 128, 0, s>>>(mp, (int)grid.x, NG, kTile);
    count_launch();
    rc = check_launch("mla_merge_flat");
  } else if (splits > 1) {
    mla_merge_kernel<<<dim3((unsigned)p.num_heads, (unsigned)p.num_seqs), 128, 0, s>>>(mp);
    count_launch();
    rc = check_launch("mla_merge");
  }
  return rc;
}
