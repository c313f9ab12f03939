// Shared pieces of the streaming paged-attention kernels (attention_stream.cu: mma.sync consumers;
// attention_umma_fp8.cu: tcgen05 consumers): parameters, PTX wrappers, the work decomposition (flat tile-balanced
// schedule / fixed-partition items) and the ring-stage geometry.  Everything lives in an anonymous namespace:
// each translation unit gets its own copy.
#pragma once
#include <float.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>
#include <type_traits>

#include "common.cuh"

namespace b200 {
namespace {


constexpr int kD = 128;             // head size served by this kernel
constexpr int kTileTokens = 32;     // tokens per ring stage
constexpr int kConsumerWarps = 8;
// producer warps: 4 (384 threads, up to 168 registers each) or 8 (512 threads, 128 registers: ptxas fits the
// consumer loop into 128 without spills; setmaxnreg was tried and ptxas does not allocate past the launch bound)
constexpr int kMergeStride = kD + 4;  // fp32 words per (warp, head) row
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;

struct alignas(64) TmaDesc {
  uint64_t w[16];  // a CUtensorMap (128 bytes), filled on the host by cuTensorMapEncodeTiled
};
struct StreamParams {
  TmaDesc tm_k, tm_v;       // flash layout only: 4-D tensor maps of the key / value cache
  const __nv_bfloat16* q;
  const uint8_t* kc;
  const uint8_t* vc;
  const int32_t* bt;
  const int32_t* seq_lens;
  const float* alibi;
  const float* k_scale;
  const float* v_scale;
  __nv_bfloat16* out;
  float* lse;               // optional [H, S] log-sum-exp of the scaled logits (natural log), with the final output
  SplitOut so;              // item schedule only
  float* ws_o;              // flat schedule: slots [range][2][G][kD] (unnormalised partial outputs)
  float* ws_m;              //                      [range][2][G]     (max logit, log2 units)
  float* ws_l;              //                      [range][2][G]     (sum exp)
  int* counters;            //                      [S * Hkv] arrival counters, zero between launches
  int* claim_counter;       //                      next chunk of the dynamic tail region, zero between launches
  int dyn_min;              //                      tiles of the smallest dynamic chunk (0: static schedule only)
  int64_t q_stride;         // elements
  int64_t kv_block_stride;  // bytes
  int64_t kv_head_stride;   // bytes
  int64_t kv_page_stride;   // bytes; flash layout only: stride between the tokens of a block
  int block_size;           // flash layout only (runtime; a multiple of 16)
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
// one non-blocking probe of a phase
__device__ __forceinline__ bool mbar_test(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}
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
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
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
// four e4m3 bytes -> two f16x2 (bytes 0,1 and bytes 2,3); the 16-bit halves are named in PTX so that ptxas
// uses the .H1 operand selector instead of masking / shifting them out with LOP3
__device__ __forceinline__ void cvt_e4m3x4_f16x2x2(uint32_t r, uint32_t& lo, uint32_t& hi) {
  asm("{\n.reg .b16 l, h;\nmov.b32 {l, h}, %2;\ncvt.rn.f16x2.e4m3x2 %0, l;\ncvt.rn.f16x2.e4m3x2 %1, h;\n}"
      : "=r"(lo), "=r"(hi)
      : "r"(r));
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

// ---- work decomposition ----------------------------------------------------------
// A segment is a run of whole tiles of one (seq, kv head) pair processed by one CTA.
struct Seg {
  int s, kvh, t0, t1, L, part;
  bool starts, ends;   // flat: the segment contains the first / last tile of its pair
  long long pb;        // flat: position of the pair's first tile in the flat tile sequence
  int n;               // flat: tiles of the pair
  int rid;             // flat: range (static share of a CTA, or dynamic chunk) the segment belongs to
};
// Flat schedule of this CTA, computed once by warp 0 (shared memory).
// The flat tile sequence [0, T) is cut into a STATIC region [0, Ts): one equal contiguous range per CTA (range id =
// CTA index), and a DYNAMIC tail region [Ts, T) that CTAs claim chunk by chunk as they run out of work (range ids
// neff, neff + 1, ... in tile order), so that SMs that stream faster than others (the spread between the first and the
// last CTA of an equal split is ~10 % of the kernel: profiles/r2_timeline.log) take more of it.  Chunk sizes are "guided":
// nlev levels of neff chunks each, level l made of chunks of dyn_min << (nlev - 1 - l) tiles — big chunks first, dyn_min
// tiles at the very end.
constexpr int kDynMaxClaims = kFlatMaxCtas * kDynMaxLevels / 2 + 2;  // claims one CTA can make (in practice a handful)
constexpr int kSeqBlocks = 256;  // 32-sequence blocks with a prefix entry: dynamic region only for S <= 8192
struct Sched {
  long long f0, f1;    // this CTA's static tile range
  long long base0;     // flat position of the first tile of sequence s0
  long long T;         // tiles in the batch
  long long Ts;        // tiles in the static region
  int s0;              // sequence holding tile f0
  int neff;            // CTAs that own a non-empty static range = min(gridDim.x, T)
  int nlev;            // levels of the dynamic region (0: none)
  int ndyn;            // dynamic chunks = neff * nlev
  int dyn_min;
  int claims[kDynMaxClaims];       // k-th claim of this CTA: -1 not made, -2 being made, else the chunk index
  long long blkpre[kSeqBlocks + 1];  // tiles (per kv head) of the sequences before sequence 32 * b
};
__device__ __forceinline__ int tiles_of(int L) { return L > 0 ? (L + kTileTokens - 1) / kTileTokens : 0; }
__device__ __forceinline__ long long warp_sum_ll(long long v) {
#pragma unroll
  for (int m = 16; m >= 1; m >>= 1) v += __shfl_xor_sync(0xffffffffu, v, m);
  return v;
}
// CTA that owns flat tile x when T tiles are cut into n ranges [c*T/n, (c+1)*T/n)
__device__ __forceinline__ int cta_of_tile(long long x, long long T, int n) {
  return (int)(((x + 1) * n - 1) / T);
}
// first tile / tile count of dynamic chunk j
__device__ __forceinline__ void dyn_chunk(const Sched* sc, int j, long long& start, int& len) {
  const int l = j / sc->neff, i = j - l * sc->neff;
  len = sc->dyn_min << (sc->nlev - 1 - l);
  // levels before l hold neff * dyn_min * (2^nlev - 2^(nlev - l)) tiles
  start = sc->Ts + (long long)sc->neff * sc->dyn_min * ((1 << sc->nlev) - (1 << (sc->nlev - l))) + (long long)i * len;
}
// range id of flat tile x (ranges are numbered in tile order)
__device__ __forceinline__ int range_of_tile(const Sched* sc, long long x) {
  if (x < sc->Ts) return cta_of_tile(x, sc->Ts, sc->neff);
  long long y = x - sc->Ts;
  for (int l = 0; l < sc->nlev; ++l) {
    const int len = sc->dyn_min << (sc->nlev - 1 - l);
    const long long lvl = (long long)sc->neff * len;
    if (y < lvl) return sc->neff * (1 + l) + (int)(y / len);
    y -= lvl;
  }
  return sc->neff * (1 + sc->nlev) - 1;  // not reached for x < T
}
__device__ void compute_sched(const StreamParams& p, Sched* sc, int lane) {
  // tiles per kv head before every 32nd sequence, and the total.  Four blocks of 32 lengths are fetched before any is
  // used: this is the first global access of a CTA that has just started (a miss all the way to HBM, ~1 us), and the
  // kernel cannot issue a single KV byte before the schedule is known.
  long long running = 0;
  for (int b = 0; b < p.S; b += 128) {
    int n4[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int i = b + 32 * k + lane;
      n4[k] = i < p.S ? tiles_of(__ldg(p.seq_lens + i)) : 0;
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int bb = b + 32 * k;
      if (bb < p.S) {
        if (lane == 0 && (bb >> 5) <= kSeqBlocks) sc->blkpre[bb >> 5] = running;
        running += __reduce_add_sync(0xffffffffu, n4[k]);  // 32 sequences of < 2^26 tiles each
      }
    }
  }
  if (lane == 0 && ((p.S + 31) >> 5) <= kSeqBlocks) sc->blkpre[(p.S + 31) >> 5] = running;
  const long long T = running * p.Hkv;
  const int neff = (int)(T < (long long)gridDim.x ? T : (long long)gridDim.x);
  // dynamic region: ~15 % of the tiles, only when every CTA has a static share worth splitting
  int nlev = 0;
  if (p.dyn_min > 0 && neff == (int)gridDim.x && p.S <= 32 * kSeqBlocks) {
    const long long per = T / neff;
    while (nlev < kDynMaxLevels && (long long)p.dyn_min * ((2 << nlev) - 1) * 100 <= per * 15) ++nlev;
    while (nlev > 0 && neff * (1 + nlev) > kFlatMaxRanges) --nlev;
  }
  const long long Ts = T - (long long)neff * p.dyn_min * ((1 << nlev) - 1);
  long long f0 = 0, f1 = 0;
  if ((int)blockIdx.x < neff) {
    f0 = (long long)blockIdx.x * Ts / neff;
    f1 = (long long)(blockIdx.x + 1) * Ts / neff;
  }
  if (lane == 0) {
    sc->f0 = f0; sc->f1 = f1; sc->T = T; sc->Ts = Ts; sc->neff = neff; sc->s0 = 0; sc->base0 = 0;
    sc->nlev = nlev; sc->ndyn = neff * nlev; sc->dyn_min = p.dyn_min;
  }
  if (nlev > 0)
    for (int i = lane; i < kDynMaxClaims; i += 32) sc->claims[i] = -1;
  if (f0 >= f1) return;
  running = 0;  // tiles per kv head before sequence `b`
  for (int b = 0; b < p.S; b += 32) {
    const int n = b + lane < p.S ? tiles_of(__ldg(p.seq_lens + b + lane)) : 0;
    int incl = n;
#pragma unroll
    for (int m = 1; m < 32; m <<= 1) {
      const int o = __shfl_up_sync(0xffffffffu, incl, m);
      if (lane >= m) incl += o;
    }
    const long long lo = (running + incl - n) * p.Hkv, hi = (running + incl) * p.Hkv;
    const bool hit = n > 0 && lo <= f0 && f0 < hi;
    const unsigned ball = __ballot_sync(0xffffffffu, hit);
    if (ball) {
      if (hit) {
        sc->s0 = b + lane;
        sc->base0 = lo;
      }
      return;
    }
    running += __shfl_sync(0xffffffffu, incl, 31);
  }
}

// SOLO: the caller is a single live lane of its warp (producer lanes); otherwise all 32 lanes call in lockstep.
template <bool FLAT, bool SOLO = false>
struct WorkIter {
  long long f, f1, base;
  int s, n, L, item, step, rid, kclaim;
  Sched* sc;
  __device__ __forceinline__ void init(const StreamParams& p, Sched* sched) {
    if constexpr (FLAT) {
      sc = sched;
      f = sc->f0; f1 = sc->f1; base = sc->base0; s = sc->s0;
      rid = (int)blockIdx.x; kclaim = 0;
      L = 0; n = 0;
      if (f < f1) {
        L = __ldg(p.seq_lens + s);
        n = tiles_of(L);
      }
    } else {
      item = (int)blockIdx.x - (int)gridDim.x;
      step = (int)gridDim.x;
    }
  }
  // item schedule, per-warp items: of the CTA's item list blockIdx + k * gridDim, warp w takes k == w (mod 8)
  __device__ __forceinline__ void init_warp_items(int w) {
    step = (int)gridDim.x * kConsumerWarps;
    item = (int)blockIdx.x + w * (int)gridDim.x - step;
  }
  // Claim the CTA's next dynamic chunk.  The first of the CTA's iterators (producer lanes, consumer warps) to get
  // here takes it from the global counter and publishes it in shared memory; the others read it there.
  __device__ __forceinline__ bool next_range(const StreamParams& p) {
    if (sc->nlev == 0 || kclaim >= kDynMaxClaims) return false;
    int j = 0;
    if (SOLO || (threadIdx.x & 31) == 0) {
      volatile int* slot = sc->claims + kclaim;
      j = atomicCAS(sc->claims + kclaim, -1, -2);
      if (j == -1) {
        j = atomicAdd(p.claim_counter, 1);
        // every CTA makes exactly one failing claim: the last of them leaves the counter at zero for the next launch
        if (j == sc->ndyn + (int)gridDim.x - 1) *p.claim_counter = 0;
        if (j > sc->ndyn) j = sc->ndyn;
        __threadfence_block();
        *slot = j;
      } else {
        while (j < 0) j = *slot;
      }
    }
    if constexpr (!SOLO) j = __shfl_sync(0xffffffffu, j, 0);
    ++kclaim;
    if (j >= sc->ndyn) {
      kclaim = kDynMaxClaims;  // exhausted
      return false;
    }
    int len;
    dyn_chunk(sc, j, f, len);
    f1 = f + len;
    rid = sc->neff + j;
    return true;
  }
  __device__ __forceinline__ bool next(const StreamParams& p, Seg& g) {
    if constexpr (FLAT) {
      if (f >= f1 && !next_range(p)) return false;
      while (f >= base + (long long)p.Hkv * n) {  // f < f1 <= T: a later sequence holds tile f
        base += (long long)p.Hkv * n;
        ++s;
        if ((s & 31) == 0 && sc->nlev > 0) {  // whole 32-sequence blocks before tile f are skipped through the prefix table
          int b = s >> 5;
          while (32 * (b + 1) < p.S && sc->blkpre[b + 1] * p.Hkv <= f) ++b;
          s = 32 * b;
          base = sc->blkpre[b] * p.Hkv;
        }
        L = __ldg(p.seq_lens + s);
        n = tiles_of(L);
      }
      const long long r = f - base;
      const int kvh = (int)(r / n);
      const int tile0 = (int)(r - (long long)kvh * n);
      const long long pb = base + (long long)kvh * n, pe = pb + n;
      const long long se = pe < f1 ? pe : f1;
      const int tile1 = tile0 + (int)(se - f);
      g.s = s; g.kvh = kvh; g.L = L; g.part = 0;
      g.t0 = tile0 * kTileTokens;
      g.t1 = min(tile1 * kTileTokens, L);
      g.starts = tile0 == 0; g.ends = se == pe; g.pb = pb; g.n = n; g.rid = rid;
      f = se;
      return true;
    } else {
      const int SH = p.S * p.Hkv;
      for (;;) {
        item += step;
        if (item >= p.total_items) return false;
        g.part = item / SH;
        const int r = item - g.part * SH;
        g.s = r / p.Hkv;
        g.kvh = r - g.s * p.Hkv;
        g.L = __ldg(p.seq_lens + g.s);
        g.t0 = g.part * p.so.chunk;
        if (g.t0 >= g.L) continue;
        g.t1 = min(g.t0 + p.so.chunk, g.L);
        g.starts = g.ends = false; g.pb = 0; g.n = 0; g.rid = 0;
        return true;
      }
    }
  }
};

template <int B, bool FP8>
struct Cfg {
  static constexpr int kElemBytes = FP8 ? 1 : 2;
  static constexpr int kBlkBytes = kD * B * kElemBytes;  // one K (or V) chunk of a block
  static constexpr int kBlocksPerTile = kTileTokens / B;
  static constexpr int kStageBytes = 2 * kBlocksPerTile * kBlkBytes;
  static constexpr int kVOffset = kBlocksPerTile * kBlkBytes;  // V chunks follow the K chunks
};

// One issuing lane of producer warp `pw` (of NP): walks the block table of the CTA's tile stream and streams the
// K chunk and V chunk of every block of its tiles (gt == pw mod NP) into the ring with 1-D TMA bulk copies.
template <int B, bool FP8, bool FLAT, int NP>
__device__ __forceinline__ void producer_loop(const StreamParams& p, Sched* sched, int pw, uint32_t ring_u32,
                                              uint32_t full_u32, uint32_t empty_u32, int nst) {
  using C = Cfg<B, FP8>;
  constexpr int kProducerWarps = NP;
  uint32_t g_next = (uint32_t)pw;   // next tile of the stream this warp issues
  int stage = pw;                   // g_next % nst (nst >= kProducerWarps)
  uint32_t phase = 0;               // (g_next / nst) & 1
  uint32_t tile_base = 0;
  WorkIter<FLAT, true> wi;
  wi.init(p, sched);
  Seg it;
  int item = -1;  // segment ordinal (debug messages)
  while (wi.next(p, it)) {
    ++item;
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

}

// Per-warp items (item schedule): producer lane pw feeds consumer warp pw alone.  It walks the CTA's item list, takes
// the items k == pw (mod 8) — the same rule the consumer applies — and streams ALL tiles of each through the
// warp's ring slice [pw * spw, (pw + 1) * spw).
template <int B, bool FP8>
__device__ __forceinline__ void producer_loop_pw(const StreamParams& p, int pw, uint32_t ring_u32, uint32_t full_u32,
                                                 uint32_t empty_u32, int spw) {
  using C = Cfg<B, FP8>;
  WorkIter<false, true> wi;
  wi.init(p, nullptr);
  wi.init_warp_items(pw);
  int ls = 0;
  [[maybe_unused]] int k = 0;
  uint32_t phase = 0;
  auto next_mine = [&](Seg& sg) -> bool { return wi.next(p, sg); };
  // The first block ids of the NEXT item are fetched while the current item streams: at 16 tiles per item an exposed
  // block-table load per item (~1 us under load) left the warp's three-stage slice empty for ~10 % of the time.
  Seg it, nx;
  bool has = next_mine(nx);
  int fid0 = 0, fid1 = 0;
  auto first_ids = [&](const Seg& sg) {
    const int32_t* b = p.bt + (int64_t)sg.s * p.max_blocks + sg.t0 / B;
    fid0 = __ldg(b);
    fid1 = (C::kBlocksPerTile == 2 && sg.t1 - sg.t0 > B) ? __ldg(b + 1) : 0;
  };
  if (has) first_ids(nx);
  while (has) {
    it = nx;
    const int nblk = (it.t1 - it.t0 + B - 1) / B;
    const int ntiles = (it.t1 - it.t0 + kTileTokens - 1) / kTileTokens;
    const int32_t* bt = p.bt + (int64_t)it.s * p.max_blocks + it.t0 / B;
    const uint8_t* kbase = p.kc + (int64_t)it.kvh * p.kv_head_stride;
    const uint8_t* vbase = p.vc + (int64_t)it.kvh * p.kv_head_stride;
    int nid0 = fid0, nid1 = fid1;
    has = next_mine(nx);
    if (has) first_ids(nx);
    for (int t = 0; t < ntiles; ++t) {
      const int b0 = t * C::kBlocksPerTile;
      const int id0 = nid0, id1 = nid1;
      const int nvalid = min(C::kBlocksPerTile, nblk - b0);
      const int bn = b0 + C::kBlocksPerTile;
      if (bn < nblk) {
        nid0 = __ldg(bt + bn);
        if (C::kBlocksPerTile == 2 && bn + 1 < nblk) nid1 = __ldg(bt + bn + 1);
      }
      const int stage = pw * spw + ls;
      MBAR_WAIT(empty_u32 + 8u * stage, phase ^ 1u, 1, k, stage);
      const uint32_t fb = full_u32 + 8u * stage;
      mbar_expect_tx(fb, (uint32_t)(nvalid * 2 * C::kBlkBytes));
      const uint32_t dst = ring_u32 + (uint32_t)stage * C::kStageBytes;
      bulk_g2s(dst, kbase + (int64_t)id0 * p.kv_block_stride, C::kBlkBytes, fb);
      bulk_g2s(dst + C::kVOffset, vbase + (int64_t)id0 * p.kv_block_stride, C::kBlkBytes, fb);
      if (C::kBlocksPerTile == 2 && nvalid == 2) {
        bulk_g2s(dst + C::kBlkBytes, kbase + (int64_t)id1 * p.kv_block_stride, C::kBlkBytes, fb);
        bulk_g2s(dst + C::kVOffset + C::kBlkBytes, vbase + (int64_t)id1 * p.kv_block_stride, C::kBlkBytes, fb);
      }
      if (++ls == spw) {
        ls = 0;
        phase ^= 1u;
      }
    }
  }
}

// ---- flash layout ([num_blocks, block_size, num_kv_heads, D], NHD or HND strides; bf16) ---------------------
// One token row of one kv head is D contiguous elements (256 B); the rows of a block are page_stride apart
// (NHD: Hkv*D elements, HND: D).  Separate 256-byte bulk copies top out at ~4 TB/s (one TMA operation per ~18
// cycles per SM; profiles/), so the rows travel as TILED TMA boxes through a 4-D tensor map over
// (dim, kv head, token in block, block): one box = 16 tokens x 64 dims of one kv head = 2 KiB, written to shared
// memory with the 128-byte swizzle so that the 8 rows an ldmatrix phase touches fall into distinct bank groups.
// A ring stage (32 tokens) = [K: half 0 dims 0-63 | half 0 dims 64-127 | half 1 ... ][V: same], 8 boxes, 16 KiB.
constexpr int kFlashBoxBytes = 16 * 128;                 // 16 token rows of 128 B (64 bf16 dims)
constexpr int kFlashVOffset = 4 * kFlashBoxBytes;        // V boxes follow the K boxes
constexpr int kFlashStageBytes = 8 * kFlashBoxBytes;     // 16384

__device__ __forceinline__ void tma_load_4d(uint32_t dst, const void* tmap, int c0, int c1, int c2, int c3, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(dst), "l"(tmap), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}

// One issuing lane of producer warp `pw`: per tile, up to 8 boxes (K and V, two 16-token halves, two 64-dim halves).
template <bool FLAT, int NP>
__device__ __forceinline__ void producer_loop_flash(const StreamParams& p, const void* tm_k, const void* tm_v,
                                                    Sched* sched, int pw, uint32_t ring_u32, uint32_t full_u32,
                                                    uint32_t empty_u32, int nst) {
  constexpr int kProducerWarps = NP;
  uint32_t g_next = (uint32_t)pw;
  int stage = pw;
  uint32_t phase = 0;
  uint32_t tile_base = 0;
  const int B = p.block_size;
  WorkIter<FLAT, true> wi;
  wi.init(p, sched);
  Seg it;
  int item = -1;
  while (wi.next(p, it)) {
    ++item;
    const uint32_t ntiles = (uint32_t)((it.t1 - it.t0 + kTileTokens - 1) / kTileTokens);
    const int32_t* bt = p.bt + (int64_t)it.s * p.max_blocks;
    const uint32_t tile_end = tile_base + ntiles;
    // block ids of the two 16-token halves of the next tile, fetched one step ahead
    int ntok = it.t0 + (int)(g_next - tile_base) * kTileTokens;
    int nid0 = 0, nid1 = 0;
    if (g_next < tile_end) {
      nid0 = __ldg(bt + ntok / B);
      if (ntok + 16 < it.t1) nid1 = __ldg(bt + (ntok + 16) / B);
    }
    while (g_next < tile_end) {
      const int tok = ntok, id0 = nid0, id1 = nid1;
      const bool two = tok + 16 < it.t1;  // the second half holds at least one valid token
      ntok = tok + kProducerWarps * kTileTokens;
      if (g_next + kProducerWarps < tile_end) {
        nid0 = __ldg(bt + ntok / B);
        if (ntok + 16 < it.t1) nid1 = __ldg(bt + (ntok + 16) / B);
      }
      MBAR_WAIT(empty_u32 + 8u * stage, phase ^ 1u, 1, item, stage);
      const uint32_t fb = full_u32 + 8u * stage;
      mbar_expect_tx(fb, (uint32_t)((two ? 8 : 4) * kFlashBoxBytes));
      const uint32_t dst = ring_u32 + (uint32_t)stage * kFlashStageBytes;
      const int off0 = tok % B;
      tma_load_4d(dst, tm_k, 0, it.kvh, off0, id0, fb);
      tma_load_4d(dst + kFlashBoxBytes, tm_k, 64, it.kvh, off0, id0, fb);
      tma_load_4d(dst + kFlashVOffset, tm_v, 0, it.kvh, off0, id0, fb);
      tma_load_4d(dst + kFlashVOffset + kFlashBoxBytes, tm_v, 64, it.kvh, off0, id0, fb);
      if (two) {
        const int off1 = (tok + 16) % B;
        tma_load_4d(dst + 2 * kFlashBoxBytes, tm_k, 0, it.kvh, off1, id1, fb);
        tma_load_4d(dst + 3 * kFlashBoxBytes, tm_k, 64, it.kvh, off1, id1, fb);
        tma_load_4d(dst + kFlashVOffset + 2 * kFlashBoxBytes, tm_v, 0, it.kvh, off1, id1, fb);
        tma_load_4d(dst + kFlashVOffset + 3 * kFlashBoxBytes, tm_v, 64, it.kvh, off1, id1, fb);
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
}

// Developer knobs (B200_STREAM_NSTAGES, B200_STREAM_NPROD, B200_DEBUG_NUM_BLOCKS) are read from the environment ONCE
// per process and name (the launch path never calls getenv); 0 = unset.
inline int env_int(const char* name) {
  struct Slot { const char* name; int value; };
  static Slot slots[8];
  static int nslots = 0;
  static std::mutex mu;
  std::lock_guard<std::mutex> lock(mu);
  for (int i = 0; i < nslots; ++i)
    if (strcmp(slots[i].name, name) == 0) return slots[i].value;
  const char* e = getenv(name);
  const int v = e ? atoi(e) : 0;
  if (nslots < 8) slots[nslots++] = Slot{name, v};
  return v;
}

// host: C-ABI parameters -> kernel parameters (strides in bytes)
StreamParams fill_stream_params(const b200_paged_attn_params_t& p) {
  const int eb = p.kv_dtype == B200_KV_FP8_E4M3 ? 1 : 2;
  StreamParams sp;
  memset(&sp, 0, sizeof(sp));
  sp.q = static_cast<const __nv_bfloat16*>(p.query);
  sp.kc = static_cast<const uint8_t*>(p.key_cache);
  sp.vc = static_cast<const uint8_t*>(p.value_cache);
  sp.bt = p.block_tables; sp.seq_lens = p.seq_lens; sp.alibi = p.alibi_slopes;
  sp.k_scale = p.k_scale; sp.v_scale = p.v_scale;
  sp.out = static_cast<__nv_bfloat16*>(p.out);
  sp.lse = p.out_lse;
  sp.q_stride = p.q_stride;
  sp.kv_block_stride = p.kv_block_stride * eb;
  sp.kv_head_stride = p.kv_head_stride * eb;
  sp.S = p.num_seqs; sp.H = p.num_heads; sp.Hkv = p.num_kv_heads; sp.G = p.num_heads / p.num_kv_heads;
  sp.max_blocks = p.max_num_blocks_per_seq;
  sp.scale = p.scale;
  sp.debug_num_blocks = env_int("B200_DEBUG_NUM_BLOCKS");
  return sp;
}

}  // namespace
}  // namespace b200
