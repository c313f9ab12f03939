// Micro-benchmark (developer tool): what read bandwidth does B200 give a persistent kernel that only
// streams randomly placed chunks into a shared-memory ring with 1-D TMA bulk copies?  This is the
// ceiling of the paged KV gather: the attention kernel cannot be faster.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gpurun_out/gather_read_bench tools/gather_read_bench.cu
//   ./gather_read_bench
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include <algorithm>
#include <numeric>
#include <random>
#include <vector>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n.reg .pred p;\nWAIT_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra DONE_%=;\nbra WAIT_%=;\nDONE_%=:\n}\n" ::"r"(bar),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}

// Each CTA processes `per_cta` stage-loads; a stage = `chunks_per_stage` chunks of `chunk` bytes whose
// source offsets come from `table`.  `nprod` producer lanes (one per warp) deal stages round-robin;
// one consumer warp per stage reads 16 B per lane (touches the data) and releases it.
__global__ void __launch_bounds__(384, 1)
gather_kernel(const uint8_t* __restrict__ src, const int32_t* __restrict__ table, int per_cta, int chunk,
              int chunks_per_stage, int nst, int nprod, int ncons, int lanes_p, unsigned long long* sink,
              unsigned int* bad, size_t base_off, int delay_ns) {
  extern __shared__ __align__(128) uint8_t smem[];
  const int stage_bytes = chunk * chunks_per_stage;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (size_t)nst * stage_bytes);
  const uint32_t ring = smem_u32(smem), full = smem_u32(bars), empty = full + 8u * nst;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) {
    for (int i = 0; i < nst; ++i) {
      mbar_init(full + 8u * i, 1);
      mbar_init(empty + 8u * i, 1);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const int32_t* tb = table + (size_t)blockIdx.x * per_cta * chunks_per_stage;
  if (warp >= ncons) {
    const int pw = warp - ncons;
    if (pw >= nprod || lane >= lanes_p) return;
    // lanes_p lanes share the chunks of a stage; lane 0 posts the expected byte count
    int32_t nxt = -1;
    if (pw < per_cta && lane < chunks_per_stage) nxt = __ldg(tb + pw * chunks_per_stage + lane);
    for (int t = pw; t < per_cta; t += nprod) {
      const int stage = t % nst;
      const uint32_t ph = (t / nst) & 1;
      const int32_t cur = nxt;
      if (t + nprod < per_cta && lane < chunks_per_stage) nxt = __ldg(tb + (t + nprod) * chunks_per_stage + lane);
      mbar_wait(empty + 8u * stage, ph ^ 1u);
      if (lane == 0) mbar_expect_tx(full + 8u * stage, stage_bytes);
      if (lanes_p == 1) {
        bulk_g2s(ring + stage * stage_bytes, src + (size_t)cur * chunk, chunk, full + 8u * stage);
        for (int c = 1; c < chunks_per_stage; ++c)
          bulk_g2s(ring + stage * stage_bytes + c * chunk, src + (size_t)__ldg(tb + t * chunks_per_stage + c) * chunk,
                   chunk, full + 8u * stage);
      } else {
        for (int c = lane; c < chunks_per_stage; c += lanes_p)
          bulk_g2s(ring + stage * stage_bytes + c * chunk,
                   src + (size_t)(c == lane ? cur : __ldg(tb + t * chunks_per_stage + c)) * chunk, chunk,
                   full + 8u * stage);
      }
    }
    return;
  }
  unsigned long long acc = 0;
  for (int t = warp; t < per_cta; t += ncons) {
    const int stage = t % nst;
    const uint32_t ph = (t / nst) & 1;
    mbar_wait(full + 8u * stage, ph);
    if (delay_ns > 0) __nanosleep(delay_ns);  // slow consumer: the ring stays full, producers block on `empty`
    const uint4 v = *reinterpret_cast<const uint4*>(smem + (size_t)stage * stage_bytes + lane * 16);
    acc += v.x ^ v.y ^ v.z ^ v.w;
    if (bad != nullptr) {
      // verify the data: every chunk of the stage, 16 B at 32 evenly spread offsets incl. the last 16 B
      for (int c = 0; c < chunks_per_stage; ++c) {
        const size_t so = (size_t)__ldg(tb + t * chunks_per_stage + c) * chunk;
        const int off = lane == 31 ? chunk - 16 : lane * (chunk / 32);
        const uint4 w = *reinterpret_cast<const uint4*>(smem + (size_t)stage * stage_bytes + c * chunk + off);
        const unsigned e0 = (unsigned)((so + off + base_off) >> 2) * 2654435761u;
        if (w.x != e0 || w.y != e0 + 2654435761u || w.z != e0 + 2u * 2654435761u || w.w != e0 + 3u * 2654435761u)
          atomicAdd(bad + (lane == 31 ? 1 : 0), 1u);
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(empty + 8u * stage);
  }
  if (acc == 0x123456789abcdefULL) *sink = acc;
}

__global__ void fill_pattern(unsigned* p, size_t nwords) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < nwords; i += (size_t)gridDim.x * blockDim.x)
    p[i] = (unsigned)i * 2654435761u;
}

int main(int argc, char** argv) {
  const bool verify = argc > 1 && argv[1][0] == 'v';
  const size_t base_off = argc > 2 ? (size_t)atoi(argv[2]) : 0;  // misalign the chunk grid (bytes, multiple of 16)
  const int delay_ns = argc > 3 ? atoi(argv[3]) : 0;
  const size_t total = (size_t)2 << 30;  // 2 GiB pool
  uint8_t* src;
  cudaMalloc(&src, total);
  fill_pattern<<<1184, 256>>>(reinterpret_cast<unsigned*>(src), total / 4);
  unsigned int* bad;
  cudaMalloc(&bad, 8);
  cudaMemset(bad, 0, 8);
  unsigned long long* sink;
  cudaMalloc(&sink, 8);
  int nsm = 0;
  cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, 0);
  cudaFuncSetAttribute(gather_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
  printf("SMs %d\n", nsm);
  printf("%8s %6s %8s %6s %6s %6s %6s %6s %10s %10s\n", "chunk", "cps", "inflight", "nst", "nprod", "lanes", "ncons", "random", "us", "GB/s");
  const size_t read_bytes = (size_t)1 << 30;  // 1 GiB per launch
  std::mt19937 rng(0);
  struct PC { int nprod, lanes, ncons; };
  const PC pcs[] = {{4, 1, 8}, {4, 2, 8}, {4, 4, 8}, {2, 4, 8}, {1, 4, 8}, {8, 1, 4}, {8, 4, 4}, {8, 2, 4}};
  for (int random = 1; random >= 0; --random)
    for (int chunk : {2048, 4096, 8192, 16384})
      for (int cps : {1, 2, 4})
        for (int inflight_kib : {32, 64, 96, 128, 192})
          for (const PC& pc : pcs) {
            const int nprod = pc.nprod, ncons = pc.ncons, lanes_p = pc.lanes;
            if (lanes_p > cps) continue;
            const int stage_bytes = chunk * cps;
            if (stage_bytes > 16384) continue;
            if (verify && (pc.nprod != 4 || pc.lanes != 1)) continue;
            // a stage must always be served by the same producer / consumer warp (mbarrier parity only tells odd from even)
            const int m = std::max(ncons, nprod);
            const int nst = inflight_kib * 1024 / stage_bytes / m * m;
            if (nst < m || nst * stage_bytes != inflight_kib * 1024) continue;
            const size_t nchunks_pool = (total - base_off) / chunk - 1;
            const int per_cta = (int)(read_bytes / nsm / stage_bytes);
            const size_t nt = (size_t)nsm * per_cta * cps;
            std::vector<int32_t> tab(nchunks_pool);
            std::iota(tab.begin(), tab.end(), 0);
            if (random) std::shuffle(tab.begin(), tab.end(), rng);
            tab.resize(nt);
            int32_t* dtab;
            cudaMalloc(&dtab, nt * 4);
            cudaMemcpy(dtab, tab.data(), nt * 4, cudaMemcpyHostToDevice);
            const size_t smem = (size_t)nst * stage_bytes + nst * 16 + 128;
            cudaEvent_t e0, e1;
            cudaEventCreate(&e0);
            cudaEventCreate(&e1);
            float best = 1e9f, sum = 0;
            const int iters = 8;
            for (int it = 0; it < iters + 2; ++it) {
              cudaEventRecord(e0);
              gather_kernel<<<nsm, 384, smem>>>(src + base_off, dtab, per_cta, chunk, cps, nst, nprod, ncons, lanes_p, sink, verify ? bad : nullptr, base_off, delay_ns);
              cudaEventRecord(e1);
              cudaEventSynchronize(e1);
              float ms;
              cudaEventElapsedTime(&ms, e0, e1);
              if (it >= 2) {
                best = std::min(best, ms);
                sum += ms;
              }
            }
            cudaError_t err = cudaGetLastError();
            if (err != cudaSuccess) {
              printf("error: %s\n", cudaGetErrorString(err));
              return 1;
            }
            const double bytes = (double)nt * chunk;
            printf("%8d %6d %7dK %6d %6d %6d %6d %6d %10.1f %10.0f   (best %.0f)\n", chunk, cps, inflight_kib, nst, nprod, lanes_p, ncons, random,
                   sum / iters * 1e3, bytes / (sum / iters * 1e-3) / 1e9, bytes / (best * 1e-3) / 1e9);
            if (verify) {
              unsigned hb[2];
              cudaMemcpy(hb, bad, 8, cudaMemcpyDeviceToHost);
              printf("      verify: %u bad 16-byte samples in chunk bodies, %u in chunk tails (of %zu chunks x %d launches)\n", hb[0], hb[1], nt, iters + 2);
              cudaMemset(bad, 0, 8);
            }
            cudaFree(dtab);
          }
  return 0;
}
