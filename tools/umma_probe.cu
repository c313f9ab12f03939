// Developer probe (not part of the product): checks the tcgen05 building blocks the fp8 decode kernel relies on,
// on the paged-cache layouts exactly as they land in shared memory:
//   1. S[64 x 16]  = Q[64 x 128] . K^T   kind::f8f6f4, A = Q (e4m3, K-major, no swizzle, written by us),
//                                         B = one K chunk of the fp8 cache [D/16][16 tok][16 B] used IN PLACE
//                                         (core matrix = 8 tokens x 16 dims; LBO = 256 B, SBO = 128 B)
//   2. O[64 x 128] = P[64 x 32] . V       B = two V chunks [128 dims][16 tok] 2 KiB apart used IN PLACE
//                                         (core matrix = 8 dims x 16 tokens; LBO = 2048 B, SBO = 128 B)
//   3. where M = 64 accumulator rows live in TMEM (rows 16q..16q+15 -> lanes 32q..32q+15)
//   4. issue cost of back-to-back small MMAs.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/umma_probe tools/umma_probe.cu && ./build/umma_probe
#include <cuda_fp8.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include <vector>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n.reg .pred p;\nWAIT_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra DONE_%=;\nbra WAIT_%=;\nDONE_%=:\n}\n" ::"r"(bar),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ uint64_t make_desc(uint32_t addr, uint32_t lbo, uint32_t sbo) {
  uint64_t d = 0;
  d |= (uint64_t)((addr >> 4) & 0x3fff);
  d |= (uint64_t)((lbo >> 4) & 0x3fff) << 16;
  d |= (uint64_t)((sbo >> 4) & 0x3fff) << 32;
  d |= (uint64_t)1 << 46;  // descriptor version (Blackwell)
  return d;                // layout type 0 = no swizzle, base offset 0
}
__device__ __forceinline__ uint32_t make_idesc_e4m3(int M, int N) {
  return (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);  // f32 accum, e4m3 x e4m3, K-major both
}
__device__ __forceinline__ void umma_f8(uint32_t d_tmem, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::f8f6f4 [%0], %1, %2, %3, p;\n}\n" ::"r"(d_tmem),
      "l"(a), "l"(b), "r"(idesc), "r"(acc)
      : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

constexpr int kQBytes = 8 * 1024;   // A (Q): 8 chunks of 16 dims x 64 rows x 16 B
constexpr int kPBytes = 2 * 1024;   // A (P): 2 chunks of 16 tokens x 64 rows x 16 B
constexpr int kKBytes = 2048;       // one K chunk
constexpr int kVBytes = 4096;       // two V chunks

// out_s [128 lanes][16], out_o [128 lanes][128]; cycles[0] = cycles for `reps` batches of 16 small + 2 big MMAs
__global__ void __launch_bounds__(128, 1)
probe(const uint8_t* q, const uint8_t* p, const uint8_t* k, const uint8_t* v, float* out_s, float* out_o,
      long long* cycles, int reps, int sbo_a, float* out_s2) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sq = smem;
  uint8_t* sp = sq + kQBytes;
  uint8_t* sk = sp + kPBytes;
  uint8_t* sv = sk + kKBytes;
  uint64_t* bar = reinterpret_cast<uint64_t*>(sv + kVBytes);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar + 2);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int i = tid; i < kQBytes / 16; i += 128) reinterpret_cast<uint4*>(sq)[i] = reinterpret_cast<const uint4*>(q)[i];
  for (int i = tid; i < kPBytes / 16; i += 128) reinterpret_cast<uint4*>(sp)[i] = reinterpret_cast<const uint4*>(p)[i];
  for (int i = tid; i < kKBytes / 16; i += 128) reinterpret_cast<uint4*>(sk)[i] = reinterpret_cast<const uint4*>(k)[i];
  for (int i = tid; i < kVBytes / 16; i += 128) reinterpret_cast<uint4*>(sv)[i] = reinterpret_cast<const uint4*>(v)[i];
  if (tid == 0) {
    mbar_init(smem_u32(bar), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy smem writes -> visible to the MMA (async proxy)
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 256;" ::"r"(smem_u32(tmem_slot)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = *tmem_slot;
  const uint32_t bar_u32 = smem_u32(bar);
  const uint32_t id_s = make_idesc_e4m3(64, 16), id_o = make_idesc_e4m3(64, 128);
  uint32_t phase = 0;
  if (tid == 0) {
    // S: 4 k-steps of 32 dims; Q chunk stride (LBO) 1024, row-group stride (SBO) 128 (or 0: all groups alias rows 0-7)
    for (int kk = 0; kk < 4; ++kk)
      umma_f8(tmem + 0, make_desc(smem_u32(sq) + kk * 2048, 1024, sbo_a), make_desc(smem_u32(sk) + kk * 512, 256, 128), id_s, kk > 0);
    // O: one K = 32 step: P chunk stride 1024; V: second 16 tokens = next block, 2048 B further
    umma_f8(tmem + 128, make_desc(smem_u32(sp), 1024, sbo_a), make_desc(smem_u32(sv), 2048, 128), id_o, 0);
    umma_commit(bar_u32);
  }
  mbar_wait(bar_u32, phase);
  phase ^= 1;
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  {
    uint32_t r[16];
    const uint32_t lane_base = (uint32_t)(warp * 32) << 16;
    tmem_ld16(tmem + lane_base + 0, r);
    for (int i = 0; i < 16; ++i) out_s[(warp * 32 + lane) * 16 + i] = __uint_as_float(r[i]);
    for (int c = 0; c < 128; c += 16) {
      tmem_ld16(tmem + lane_base + 128 + c, r);
      for (int i = 0; i < 16; ++i) out_o[(warp * 32 + lane) * 128 + c + i] = __uint_as_float(r[i]);
    }
  }
  {
    // 16x256b.x2: 16 TMEM lanes x 16 columns spread over the 32 threads like an m16n8 accumulator fragment
    uint32_t r[8];
    const uint32_t lane_base = (uint32_t)(warp * 32) << 16;
    asm volatile("tcgen05.ld.sync.aligned.16x256b.x2.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(tmem + lane_base + 0));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    for (int i = 0; i < 8; ++i) out_s2[(warp * 32 + lane) * 8 + i] = __uint_as_float(r[i]);
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  // ---- issue / execution cost ------------------------------------------------------------------
  if (tid == 0) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const long long t0 = clock64();
    for (int it = 0; it < reps; ++it) {
      for (int b = 0; b < 4; ++b)
        for (int kk = 0; kk < 4; ++kk)
          umma_f8(tmem + 16 * b, make_desc(smem_u32(sq) + kk * 2048, 1024, sbo_a), make_desc(smem_u32(sk) + kk * 512, 256, 128), id_s, kk > 0);
      umma_f8(tmem + 128, make_desc(smem_u32(sp), 1024, sbo_a), make_desc(smem_u32(sv), 2048, 128), id_o, 1);
      umma_f8(tmem + 128, make_desc(smem_u32(sp), 1024, sbo_a), make_desc(smem_u32(sv), 2048, 128), id_o, 1);
    }
    const long long t1 = clock64();
    umma_commit(bar_u32);
    mbar_wait(bar_u32, phase);
    const long long t2 = clock64();
    cycles[0] = t1 - t0;
    cycles[1] = t2 - t0;
  }
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 256;" ::"r"(tmem) : "memory");
}

// ---- issue / execution cost of one MMA shape: `n` MMAs per batch, chained on one accumulator or independent ----
// The whole warp runs the loop converged (operands stay in uniform registers); one elected lane issues.
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n.reg .pred p;\nelect.sync _|p, 0xffffffff;\nselp.u32 %0, 1, 0, p;\n}" : "=r"(pred));
  return pred != 0;
}
template <int M, int N, int CHAINED>
__global__ void __launch_bounds__(128, 1) time_shape(long long* cycles, int reps) {
  extern __shared__ __align__(1024) uint8_t smem[];   // operands: garbage bytes are fine for timing
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 64 * 1024);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar + 2);
  const int tid = threadIdx.x, warp = tid >> 5;
  for (int i = tid; i < 16 * 1024; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0x38383838u;  // e4m3 1.0
  if (tid == 0) {
    mbar_init(smem_u32(bar), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(tmem_slot)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = *tmem_slot;
  if (warp == 0) {
    constexpr uint32_t idesc = (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
    const uint64_t a = make_desc(smem_u32(smem), 2048, 128), b = make_desc(smem_u32(smem) + 16384, 4096, 128);
    constexpr int kPerBatch = 8;
    constexpr int nacc = CHAINED ? 1 : (512 / N < kPerBatch ? 512 / N : kPerBatch);
    const long long t0 = clock64();
    for (int it = 0; it < reps; ++it) {
#pragma unroll
      for (int i = 0; i < kPerBatch; ++i)
        if (elect_one()) umma_f8(tmem + (uint32_t)((i % nacc) * N), a, b, idesc, 1);
    }
    const long long t1 = clock64();
    if (elect_one()) umma_commit(smem_u32(bar));
    mbar_wait(smem_u32(bar), 0);
    const long long t2 = clock64();
    if (tid == 0) {
      cycles[0] = t1 - t0;
      cycles[1] = t2 - t0;
    }
  }
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
}

// ---- cost of the Q.K^T pattern: 8 MMAs (M64 N16 K32) per tile over 4 A operands (k-steps) and 2 B blocks ----
//   MODE 0: block-major  (A changes on every MMA)          MODE 1: k-step-major (each A used twice in a row)
//   MODE 2: k-step-major with collector::a::fill / lastuse MODE 3: two tiles batched, each A used 4x in a row (fill/use/lastuse)
#define UMMA_F8_COLL(QUAL)                                                                                           \
  asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::f8f6f4" QUAL                \
               " [%0], %1, %2, %3, p;\n}\n" ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory")
__device__ __forceinline__ void umma_fill(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) { UMMA_F8_COLL(".collector::a::fill"); }
__device__ __forceinline__ void umma_use(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) { UMMA_F8_COLL(".collector::a::use"); }
__device__ __forceinline__ void umma_last(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) { UMMA_F8_COLL(".collector::a::lastuse"); }
template <int MODE>
__global__ void __launch_bounds__(128, 1) time_qk(long long* cycles, int reps) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 64 * 1024);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar + 2);
  const int tid = threadIdx.x, warp = tid >> 5;
  for (int i = tid; i < 16 * 1024; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0x38383838u;
  if (tid == 0) {
    mbar_init(smem_u32(bar), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(tmem_slot)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = *tmem_slot;
  if (warp == 0) {
    constexpr uint32_t idesc = (1u << 4) | ((uint32_t)(16 >> 3) << 17) | ((uint32_t)(64 >> 4) << 24);
    const uint32_t qa = smem_u32(smem), ka = smem_u32(smem) + 8192;  // Q operand 8 KiB, then K blocks of 2 KiB
    const long long t0 = clock64();
    for (int it = 0; it < reps; ++it) {
      const uint32_t kt = ka + (uint32_t)(it & 7) * 4096;  // a different ring stage every tile
      if (elect_one()) {
        if (MODE == 0) {
          for (int j = 0; j < 2; ++j)
            for (int kk = 0; kk < 4; ++kk)
              umma_f8(tmem + 16 * j, make_desc(qa + kk * 2048, 1024, 128), make_desc(kt + j * 2048 + kk * 512, 256, 128), idesc, kk > 0);
        } else if (MODE == 1) {
          for (int kk = 0; kk < 4; ++kk)
            for (int j = 0; j < 2; ++j)
              umma_f8(tmem + 16 * j, make_desc(qa + kk * 2048, 1024, 128), make_desc(kt + j * 2048 + kk * 512, 256, 128), idesc, kk > 0);
        } else if (MODE == 2) {
          for (int kk = 0; kk < 4; ++kk) {
            umma_fill(tmem, make_desc(qa + kk * 2048, 1024, 128), make_desc(kt + kk * 512, 256, 128), idesc, kk > 0);
            umma_last(tmem + 16, make_desc(qa + kk * 2048, 1024, 128), make_desc(kt + 2048 + kk * 512, 256, 128), idesc, kk > 0);
          }
        } else {
          // two tiles (4 blocks) per batch: counted as 2 tiles below
          for (int kk = 0; kk < 4; ++kk) {
            umma_fill(tmem, make_desc(qa + kk * 2048, 1024, 128), make_desc(kt + kk * 512, 256, 128), idesc, kk > 0);
            umma_use(tmem + 16, make_desc(qa + kk * 2048, 1024, 128), make_desc(kt + 2048 + kk * 512, 256, 128), idesc, kk > 0);
            umma_use(tmem + 32, make_desc(qa + kk * 2048, 1024, 128), make_desc(kt + 4096 + kk * 512, 256, 128), idesc, kk > 0);
            umma_last(tmem + 48, make_desc(qa + kk * 2048, 1024, 128), make_desc(kt + 6144 + kk * 512, 256, 128), idesc, kk > 0);
          }
        }
      }
      __syncwarp();
    }
    const long long t1 = clock64();
    if (elect_one()) umma_commit(smem_u32(bar));
    mbar_wait(smem_u32(bar), 0);
    const long long t2 = clock64();
    if (tid == 0) {
      cycles[0] = t1 - t0;
      cycles[1] = t2 - t0;
    }
  }
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
}
template <int MODE>
void run_qk(long long* dc) {
  const int reps = 2000;
  cudaFuncSetAttribute(time_qk<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024 + 64);
  time_qk<MODE><<<1, 128, 64 * 1024 + 64>>>(dc, reps);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("time_qk<%d>: %s\n", MODE, cudaGetErrorString(e)); exit(1); }
  long long cyc[2];
  cudaMemcpy(cyc, dc, 16, cudaMemcpyDeviceToHost);
  const int tiles = MODE == 3 ? 2 * reps : reps;
  printf("Q.K^T pattern mode %d: issue %.1f cycles/tile, complete %.1f cycles/tile\n", MODE, (double)cyc[0] / tiles, (double)cyc[1] / tiles);
}

template <int M, int N, int CHAINED>
void run_shape(long long* dc) {
  const int reps = 1000;
  cudaFuncSetAttribute(time_shape<M, N, CHAINED>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024 + 64);
  time_shape<M, N, CHAINED><<<1, 128, 64 * 1024 + 64>>>(dc, reps);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("time_shape: %s\n", cudaGetErrorString(e)); exit(1); }
  long long cyc[2];
  cudaMemcpy(cyc, dc, 16, cudaMemcpyDeviceToHost);
  printf("%5d %5d %8d %10.1f %12.1f\n", M, N, CHAINED, (double)cyc[0] / (8 * reps), (double)cyc[1] / (8 * reps));
}

static float e4m3_to_float(uint8_t b) {
  const int s = b >> 7, e = (b >> 3) & 15, m = b & 7;
  float v = e == 0 ? ldexpf((float)m, -9) : ldexpf((float)(8 + m), e - 10);
  return s ? -v : v;
}

int main() {
  srand(1);
  auto rnd_e4m3 = [] {
    uint8_t b;
    do b = (uint8_t)(rand() & 0xff); while ((b & 0x7f) == 0x7f || (b & 0x78) > 0x40);  // no NaN, |x| < 4
    return b;
  };
  // logical matrices
  std::vector<uint8_t> Q(64 * 128), P(64 * 32), K(16 * 128), V(32 * 128);
  for (auto& x : Q) x = rnd_e4m3();
  for (auto& x : P) x = rnd_e4m3();
  for (auto& x : K) x = rnd_e4m3();
  for (auto& x : V) x = rnd_e4m3();
  // shared-memory images
  std::vector<uint8_t> sq(kQBytes), sp(kPBytes), sk(kKBytes), sv(kVBytes);
  for (int m = 0; m < 64; ++m)
    for (int d = 0; d < 128; ++d) sq[(d / 16) * 1024 + m * 16 + d % 16] = Q[m * 128 + d];
  for (int m = 0; m < 64; ++m)
    for (int t = 0; t < 32; ++t) sp[(t / 16) * 1024 + m * 16 + t % 16] = P[m * 32 + t];
  for (int t = 0; t < 16; ++t)   // fp8 key cache chunk [D/16][B=16][16]
    for (int d = 0; d < 128; ++d) sk[(d / 16) * 256 + t * 16 + d % 16] = K[t * 128 + d];
  for (int t = 0; t < 32; ++t)   // two fp8 value cache chunks [D][B=16]
    for (int d = 0; d < 128; ++d) sv[(t / 16) * 2048 + d * 16 + t % 16] = V[t * 128 + d];
  uint8_t *dq, *dp, *dk, *dv;
  float *ds, *dout;
  long long* dc;
  cudaMalloc(&dq, kQBytes); cudaMalloc(&dp, kPBytes); cudaMalloc(&dk, kKBytes); cudaMalloc(&dv, kVBytes);
  cudaMalloc(&ds, 128 * 16 * 4); cudaMalloc(&dout, 128 * 128 * 4); cudaMalloc(&dc, 16);
  float* ds2;
  cudaMalloc(&ds2, 128 * 8 * 4);
  cudaMemcpy(dq, sq.data(), kQBytes, cudaMemcpyHostToDevice);
  cudaMemcpy(dp, sp.data(), kPBytes, cudaMemcpyHostToDevice);
  cudaMemcpy(dk, sk.data(), kKBytes, cudaMemcpyHostToDevice);
  cudaMemcpy(dv, sv.data(), kVBytes, cudaMemcpyHostToDevice);
  const size_t smem = kQBytes + kPBytes + kKBytes + kVBytes + 64;
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  for (int sbo_a : {128, 0}) {
    cudaMemset(ds, 0xff, 128 * 16 * 4);
    cudaMemset(dout, 0xff, 128 * 128 * 4);
    const int reps = 2000;
    probe<<<1, 128, smem>>>(dq, dp, dk, dv, ds, dout, dc, reps, sbo_a, ds2);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) {
      printf("sbo_a=%d: CUDA error %s\n", sbo_a, cudaGetErrorString(e));
      return 1;
    }
    std::vector<float> S(128 * 16), O(128 * 128);
    long long cyc[2];
    cudaMemcpy(S.data(), ds, S.size() * 4, cudaMemcpyDeviceToHost);
    cudaMemcpy(O.data(), dout, O.size() * 4, cudaMemcpyDeviceToHost);
    cudaMemcpy(cyc, dc, 16, cudaMemcpyDeviceToHost);
    double max_s = 0, max_o = 0;
    int bad_lane = 0;
    for (int m = 0; m < 64; ++m) {
      const int lane = (m / 16) * 32 + m % 16;          // expected TMEM lane of accumulator row m
      const int msrc = sbo_a == 0 ? m % 8 : m;          // SBO = 0: every 8-row group aliases rows 0-7
      for (int n = 0; n < 16; ++n) {
        double ref = 0;
        for (int d = 0; d < 128; ++d) ref += (double)e4m3_to_float(Q[msrc * 128 + d]) * e4m3_to_float(K[n * 128 + d]);
        max_s = fmax(max_s, fabs(ref - S[lane * 16 + n]));
      }
      for (int d = 0; d < 128; ++d) {
        double ref = 0;
        for (int t = 0; t < 32; ++t) ref += (double)e4m3_to_float(P[msrc * 32 + t]) * e4m3_to_float(V[t * 128 + d]);
        max_o = fmax(max_o, fabs(ref - O[lane * 128 + d]));
      }
    }
    {
      // expected 16x256b.x2 layout: thread t of warp w, register 4j+2a+b = S[lane 32w + t/4 + 8a][col 8j + 2(t%4) + b]
      std::vector<float> S2(128 * 8);
      cudaMemcpy(S2.data(), ds2, S2.size() * 4, cudaMemcpyDeviceToHost);
      int mism = 0;
      for (int w = 0; w < 4; ++w)
        for (int t = 0; t < 32; ++t)
          for (int j = 0; j < 2; ++j)
            for (int a = 0; a < 2; ++a)
              for (int b = 0; b < 2; ++b)
                if (S2[(w * 32 + t) * 8 + 4 * j + 2 * a + b] != S[(w * 32 + t / 4 + 8 * a) * 16 + 8 * j + 2 * (t % 4) + b]) ++mism;
      printf("16x256b.x2 fragment layout mismatches: %d of 1024\n", mism);
    }
    for (int l = 0; l < 128; ++l)
      if (l % 32 >= 16 && !isnan(S[l * 16])) ++bad_lane;  // lanes 16-31 of each quadrant must be untouched (0xff.. = NaN)
    printf("sbo_a=%3d: max|S-ref| %.3e  max|O-ref| %.3e  touched upper-half lanes %d  | %d x (16 small + 2 big MMA): issue %.1f cyc/batch, complete %.1f cyc/batch\n",
           sbo_a, max_s, max_o, bad_lane, reps, (double)cyc[0] / reps, (double)cyc[1] / reps);
  }
  run_qk<0>(dc); run_qk<1>(dc); run_qk<2>(dc); run_qk<3>(dc);
  // ---- shape sweep ----
  printf("%5s %5s %8s %10s %12s\n", "M", "N", "chained", "issue/MMA", "complete/MMA");
  run_shape<64, 16, 0>(dc); run_shape<64, 16, 1>(dc);
  run_shape<64, 32, 0>(dc); run_shape<64, 64, 0>(dc);
  run_shape<64, 128, 0>(dc); run_shape<64, 128, 1>(dc); run_shape<64, 256, 0>(dc);
  run_shape<128, 16, 0>(dc); run_shape<128, 16, 1>(dc);
  run_shape<128, 32, 0>(dc); run_shape<128, 64, 0>(dc);
  run_shape<128, 128, 0>(dc); run_shape<128, 256, 0>(dc);
  return 0;
}
