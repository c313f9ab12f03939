// How fast can one SM pull rows of a [tokens x 576] bf16 latent cache into shared memory?  (development probe)
// One issuing thread per CTA keeps a ring of shared-memory slots full; nothing consumes the data.  Modes:
//   0  2-D tiled box [64 dims x 128 rows], 128-byte swizzle, source row pitch 1152 B   (the MLA kernel's K/V chunk)
//   1  same bytes as two boxes of 64 rows
//   2  3-D tiled box [64 dims x 64 rows x 9 chunks] (one instruction = 64 whole rows, lands as [chunk][row][128 B])
//   3  1-D bulk copy of 16 KiB contiguous                                                (ceiling of the copy engine)
//   4  1-D bulk copies of whole 1152-byte rows, 14 rows (16128 B) per slot
//   5  mode 0 issued by 4 threads (one per warp), each owning a quarter of the slots
//   6  2-D tiled box [64 dims x 128 rows] from a DENSE source (row pitch 128 B)
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/tma_ingest tools/tma_ingest_bench.cu -lcuda
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e_), __LINE__); exit(1); } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(c)); }
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile("{\n.reg .pred p;\nW_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D_%=;\nbra W_%=;\nD_%=:\n}\n" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void mbar_spin(uint32_t bar, uint32_t parity) {
  uint32_t ok = 0;
  while (!ok)
    asm volatile("{\n.reg .pred p;\nmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n" : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
}
__device__ long long g_cyc[4];
__device__ __forceinline__ void tma2(uint32_t dst, const void* m, int c0, int c1, uint32_t bar) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
               ::"r"(dst), "l"(m), "r"(bar), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tma3(uint32_t dst, const void* m, int c0, int c1, int c2, uint32_t bar) {
  asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
               ::"r"(dst), "l"(m), "r"(bar), "r"(c0), "r"(c1), "r"(c2) : "memory");
}
__device__ __forceinline__ void bulk(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}

struct alignas(64) Map { uint64_t w[16]; };

// rows_per_cta rows of 1152 B per CTA region; the CTA walks its region `passes` times
template <int MODE, bool SPIN = false, bool TRANSPOSED = false>
__global__ void __launch_bounds__(128, 1) ingest(const __grid_constant__ Map m2, const __grid_constant__ Map m3,
                                                 const __grid_constant__ Map mdense, const uint8_t* base, int rows_per_cta,
                                                 int passes, int nslots, int slot_bytes) {
  extern __shared__ uint8_t raw[];
  const uint32_t s0 = (smem_u32(raw) + 1023u) & ~1023u;
  uint64_t* bars = reinterpret_cast<uint64_t*>(raw + (s0 - smem_u32(raw)) + (size_t)nslots * slot_bytes);
  if (threadIdx.x == 0) {
    for (int i = 0; i < nslots; ++i) mbar_init(smem_u32(&bars[i]), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const int row0 = blockIdx.x * rows_per_cta;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nissuers = MODE == 5 ? 4 : 1;
  if (lane != 0 || warp >= nissuers) return;
  // unit of work = one slot fill.  rows consumed per fill depend on the mode
  int rows_per_fill, fills_per_group;   // a "group" = 128 rows (or 64 for mode 2, 14 for mode 4)
  if (MODE == 2) { rows_per_fill = 64; fills_per_group = 1; }
  else if (MODE == 4) { rows_per_fill = 14; fills_per_group = 1; }
  else if (MODE == 3 || MODE == 6) { rows_per_fill = 0; fills_per_group = 1; }
  else { rows_per_fill = 128; fills_per_group = 9; }   // 9 chunks of 64 dims over the same 128 rows
  const long long region_bytes = (long long)rows_per_cta * 1152;
  int fills;
  if (MODE == 3 || MODE == 6) fills = (int)(region_bytes / 16384);
  else if (MODE == 2) fills = rows_per_cta / 64;
  else if (MODE == 4) fills = rows_per_cta / 14;
  else fills = (rows_per_cta / 128) * 9;
  (void)rows_per_fill; (void)fills_per_group;
  long long issued = 0, waited = 0, t_wait = 0, t_issue = 0;
  const int total = fills * passes;
  const int my_first = warp, my_step = nissuers;   // mode 5: fill f handled by issuer f % 4, slot = f % nslots
  for (int f = my_first; f < total + nslots; f += my_step) {
    // retire the fill that used this slot nslots fills ago
    if (f >= nslots) {
      const int g = f - nslots;
      const long long c0 = clock64();
      if (SPIN) mbar_spin(smem_u32(&bars[g % nslots]), (uint32_t)((g / nslots) & 1));
      else mbar_wait(smem_u32(&bars[g % nslots]), (uint32_t)((g / nslots) & 1));
      t_wait += clock64() - c0;
      ++waited;
    }
    if (f < total) {
      const int ff = f % fills;
      const int slot = (int)(f % nslots);
      const uint32_t dst = s0 + (uint32_t)slot * slot_bytes, bar = smem_u32(&bars[slot]);
      const long long c1 = clock64();
      if (MODE == 0 || MODE == 5) {
        const int ngrp = rows_per_cta / 128;
        const int grp = TRANSPOSED ? ff % ngrp : ff / 9, c = TRANSPOSED ? ff / ngrp : ff % 9;
        mbar_expect_tx(bar, 16384);
        tma2(dst, &m2, c * 64, row0 + grp * 128, bar);
      } else if (MODE == 1) {
        const int grp = (int)(ff / 9), c = (int)(ff % 9);
        mbar_expect_tx(bar, 16384);
        tma2(dst, &m2, c * 64, row0 + grp * 128, bar);         // m2 encoded with a 64-row box for this mode
        tma2(dst + 8192, &m2, c * 64, row0 + grp * 128 + 64, bar);
      } else if (MODE == 2) {
        mbar_expect_tx(bar, 64 * 1152);
        tma3(dst, &m3, 0, row0 + (int)ff * 64, 0, bar);
      } else if (MODE == 3) {
        mbar_expect_tx(bar, 16384);
        bulk(dst, base + (long long)row0 * 1152 + ff * 16384, 16384, bar);
      } else if (MODE == 4) {
        mbar_expect_tx(bar, 14 * 1152);
        for (int r = 0; r < 14; ++r) bulk(dst + r * 1152, base + ((long long)row0 + ff * 14 + r) * 1152, 1152, bar);
      } else if (MODE == 6) {
        mbar_expect_tx(bar, 16384);
        tma2(dst, &mdense, 0, (int)(((long long)row0 * 1152 + ff * 16384) / 128), bar);
      }
      t_issue += clock64() - c1;
      ++issued;
    }
  }
  if (blockIdx.x == 0 && warp == 0) { g_cyc[0] = t_wait; g_cyc[1] = t_issue; g_cyc[2] = issued; }
}

__global__ void spin_kernel(long long cycles, long long* out) {
  const long long t0 = clock64();
  while (clock64() - t0 < cycles) {}
  if (threadIdx.x == 0 && blockIdx.x == 0) *out = clock64() - t0;
}

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                             const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main(int argc, char** argv) {
  const int sms = 148;
  EncodeFn enc = nullptr;
  cudaDriverEntryPointQueryResult q;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void**)&enc, cudaEnableDefault, &q));
  const long long rows_total = 1 << 20;   // 1.2 GB
  uint8_t* buf;
  CK(cudaMalloc(&buf, rows_total * 1152));
  CK(cudaMemset(buf, 1, rows_total * 1152));
  auto make2 = [&](int box_rows, CUtensorMapL2promotion prom) {
    Map m; CUtensorMap t;
    cuuint64_t dims[2] = {576, (cuuint64_t)rows_total}; cuuint64_t str[1] = {1152};
    cuuint32_t box[2] = {64, (cuuint32_t)box_rows}; cuuint32_t es[2] = {1, 1};
    CUresult r = enc(&t, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, buf, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, prom, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("encode2 failed %d\n", (int)r); exit(1); }
    memcpy(&m, &t, 128); return m;
  };
  auto make3 = [&]() {
    Map m; CUtensorMap t;
    cuuint64_t dims[3] = {64, (cuuint64_t)rows_total, 9}; cuuint64_t str[2] = {1152, 128};
    cuuint32_t box[3] = {64, 64, 9}; cuuint32_t es[3] = {1, 1, 1};
    CUresult r = enc(&t, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, buf, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("encode3 failed %d (3-D overlapped-stride map not accepted)\n", (int)r); memset(&m, 0, 128); return m; }
    memcpy(&m, &t, 128); return m;
  };
  auto makedense = [&]() {
    Map m; CUtensorMap t;
    cuuint64_t dims[2] = {64, (cuuint64_t)(rows_total * 9)}; cuuint64_t str[1] = {128};
    cuuint32_t box[2] = {64, 128}; cuuint32_t es[2] = {1, 1};
    CUresult r = enc(&t, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, buf, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("encode dense failed %d\n", (int)r); exit(1); }
    memcpy(&m, &t, 128); return m;
  };
  const Map m128 = make2(128, CU_TENSOR_MAP_L2_PROMOTION_L2_256B), m64 = make2(64, CU_TENSOR_MAP_L2_PROMOTION_L2_256B);
  const Map m128np = make2(128, CU_TENSOR_MAP_L2_PROMOTION_NONE);
  const Map m3 = make3(), md = makedense();
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  {  // warm the clocks up and measure them: 148 CTAs spinning for 1e9 cycles
    long long* d; CK(cudaMalloc(&d, 8));
    for (int k = 0; k < 2; ++k) {
      CK(cudaEventRecord(e0));
      spin_kernel<<<sms, 128>>>(1000000000LL, d);
      CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
      float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
      printf("spin: 1e9 cycles in %.1f ms -> SM clock %.0f MHz\n", ms, 1e6 / ms);
    }
  }
  struct Case { int mode; const char* name; int slot_bytes; };
  const Case cases[] = {{0, "2D box 64x128 pitch1152", 16384}, {1, "2x 2D box 64x64 pitch1152", 16384},
                        {2, "3D box 64x64x9 (whole rows)", 73728}, {3, "1D bulk 16 KiB", 16384},
                        {4, "1D bulk 14 rows x 1152 B", 16384}, {5, "2D box 64x128, 4 issuers", 16384},
                        {6, "2D box 64x128 dense pitch128", 16384}, {7, "2D box 64x128 pitch1152 no L2 promotion", 16384}, {8, "2D box 64x128 pitch1152, test_wait spin", 16384},
                        {9, "1D bulk 16 KiB, test_wait spin", 16384},
                        {10, "2D box 64x128 pitch1152, chunk-major order, no L2 promotion", 16384},
                        {11, "2D box 64x128 pitch1152, chunk-major order, 256 B promotion", 16384}};
  for (int hbm = 1; hbm >= 0; --hbm) {
    // HBM: every CTA walks 6272 rows (7.2 MB; 1.07 GB in total) once; L2: 224 rows (258 KB; 38 MB in total) 32 times
    const int rows_per_cta = hbm ? 7040 - 7040 % (128 * 7) : 256;   // L2: 295 KB per CTA, 44 MB in total
    const int passes = hbm ? 4 : 128;
    for (const Case& c : cases) {
      for (int ring_kb : {128, 192}) {
        const int nslots = ring_kb * 1024 / c.slot_bytes;
        if (nslots < 2) continue;
        const size_t smem = 1024 + (size_t)nslots * c.slot_bytes + 8 * nslots + 64;
        if (smem > 232448) continue;
        float best = 1e30f;
        for (int rep = 0; rep < 3; ++rep) {
          CK(cudaEventRecord(e0));
#define LAUNCH(M, MAP2) do { CK(cudaFuncSetAttribute(ingest<M>, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448)); \
          ingest<M><<<sms, 128, smem>>>(MAP2, m3, md, buf, rows_per_cta, passes, nslots, c.slot_bytes); } while (0)
          switch (c.mode) {
            case 0: LAUNCH(0, m128); break;
            case 1: LAUNCH(1, m64); break;
            case 2: if (m3.w[0] == 0 && m3.w[1] == 0) { best = -1; } else LAUNCH(2, m128); break;
            case 3: LAUNCH(3, m128); break;
            case 4: LAUNCH(4, m128); break;
            case 5: LAUNCH(5, m128); break;
            case 6: LAUNCH(6, m128); break;
            case 7: LAUNCH(0, m128np); break;
            case 8: CK(cudaFuncSetAttribute(ingest<0, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448));
                    ingest<0, true><<<sms, 128, smem>>>(m128, m3, md, buf, rows_per_cta, passes, nslots, c.slot_bytes); break;
            case 10: CK(cudaFuncSetAttribute(ingest<0, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448));
                    ingest<0, false, true><<<sms, 128, smem>>>(m128np, m3, md, buf, rows_per_cta, passes, nslots, c.slot_bytes); break;
            case 11: CK(cudaFuncSetAttribute(ingest<0, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448));
                    ingest<0, false, true><<<sms, 128, smem>>>(m128, m3, md, buf, rows_per_cta, passes, nslots, c.slot_bytes); break;
            case 9: CK(cudaFuncSetAttribute(ingest<3, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448));
                    ingest<3, true><<<sms, 128, smem>>>(m128, m3, md, buf, rows_per_cta, passes, nslots, c.slot_bytes); break;
          }
          CK(cudaEventRecord(e1));
          CK(cudaEventSynchronize(e1));
          CK(cudaGetLastError());
          float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
          if (best >= 0 && ms < best) best = ms;
        }
        if (best < 0) { printf("%s mode %d (%s): unavailable\n", hbm ? "HBM" : "L2 ", c.mode, c.name); continue; }
        long long bytes_per_cta;
        if (c.mode == 2) bytes_per_cta = (long long)(rows_per_cta / 64) * 73728;
        else if (c.mode == 4) bytes_per_cta = (long long)(rows_per_cta / 14) * 14 * 1152;
        else if (c.mode == 3 || c.mode == 6) bytes_per_cta = (long long)rows_per_cta * 1152 / 16384 * 16384;
        else bytes_per_cta = (long long)(rows_per_cta / 128) * 9 * 16384;
        const double tot = (double)bytes_per_cta * passes * sms;
        long long cyc[4];
        CK(cudaMemcpyFromSymbol(cyc, g_cyc, sizeof(cyc)));
        printf("%s ring %3d KiB  mode %d %-42s %8.1f us  %7.1f GB/s total  %6.1f GB/s per SM   (CTA 0 issuer 0: %lld cyc/wait, %lld cyc/issue; CTA 0 busy %.0f us at 1.9 GHz)\n",
               hbm ? "HBM" : "L2 ", ring_kb, c.mode, c.name, best * 1e3, tot / best / 1e6, tot / best / 1e6 / sms,
               cyc[2] ? cyc[0] / cyc[2] : 0, cyc[2] ? cyc[1] / cyc[2] : 0, (cyc[0] + cyc[1]) / 1.9e3);
        fflush(stdout);
      }
    }
  }
  return 0;
}
