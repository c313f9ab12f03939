// Decode context parallel: the exchange + combine step as ONE kernel over peer memory.
//
// After the local attention launch every rank r of the group holds, in a buffer that its peers have mapped
// (CUDA IPC over NVLink / NVSwitch), the partial result of ALL query heads over ITS token shard:
//     partial_out_r [num_seqs, world * heads_local, head_size]   (dtype)
//     partial_lse_r [world * heads_local, num_seqs]              (float32)
// Rank `rank` needs, for its heads_local heads, the N partials and their combination by the merge_attn_states rule
// (reference: all-gather of the lse + reduce-scatter of the corrected outputs, then nothing else —
// vllm_metax/v1/attention/backends/flash_attn.py:836-841 `cp_lse_ag_out_rs`; combination rule
// csrc/attention/merge_attn_states.cu:14-88).  This kernel
//   1. publishes "my partial of step `epoch` is complete" by storing `epoch` into flag[rank] of every peer
//      (st.release.sys after a system fence; the attention launch that produced the partial precedes this kernel on
//      the stream),
//   2. waits until flag[r] >= epoch for every r (ld.acquire.sys on local memory),
//   3. pulls the head slice [rank*heads_local, (rank+1)*heads_local) of every peer's partial with 16-byte loads over
//      NVLink and folds the N partials in fp32:  m = max_r lse_r, w_r = exp(lse_r - m),
//      out = sum_r w_r o_r / sum_r w_r,  lse = m + log sum_r w_r   (one rounding, at the store).
// Each rank pulls exactly the N-1 remote slices of its own heads — (N-1) * num_seqs * heads_local * (head_size *
// sizeof(T) + 4) bytes over the fabric, the volume of a reduce-scatter — with no staging copy, pack or second launch.
//
// Buffer reuse: the caller alternates two partial buffers by step parity.  A peer can be at most one step ahead
// (it cannot pass the wait of step k+1 before this rank has launched its step-k+1 kernel, which follows this rank's
// step-k kernel on the stream), so the buffer of parity k%2 is not overwritten while step k is still being read.
// Flags only grow, so "a peer is already one step ahead" also satisfies the wait.
#include <stdio.h>
#include <stdlib.h>

#include "common.cuh"

namespace b200 {
namespace {

struct DcpFoldArgs {
  const void* peer_out[B200_DCP_MAX_WORLD];
  const float* peer_lse[B200_DCP_MAX_WORLD];
  uint32_t* peer_flags[B200_DCP_MAX_WORLD];
  void* out;
  float* out_lse;
  int32_t rank, world, num_seqs, heads_local, head_size;
  uint32_t epoch;
  const uint32_t* step_counter;  // device-side step number (nullable): written by the gather of the same step
  uint64_t wait_limit_ns;
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void st_relaxed_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float ld_relaxed_sys_f32(const float* p) {
  float v;
  asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ Vec16 ld_relaxed_sys_16(const void* p) {
  Vec16 v;
  uint32_t* w = reinterpret_cast<uint32_t*>(&v);
  asm volatile("ld.relaxed.sys.global.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3])
               : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ uint64_t globaltimer_ns() {
  uint64_t t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

// A peer that never arrives traps the kernel instead of hanging the GPU.  The limit rides in the kernel arguments
// (B200_DCP_WAIT_SECONDS, default 120 s: a peer can be late for ordinary reasons — first-use module load, an
// allocator stall, a debugger — and NCCL's own watchdog counts minutes); the trap reports who waited for whom.
__device__ __noinline__ void wait_timed_out(const char* what, int rank, int peer, uint32_t epoch, uint32_t seen) {
  printf("[b200 dcp] %s: rank %d waited for rank %d to reach step %u (flag at %u): giving up\n", what, rank, peer, epoch, seen);
  __trap();
}

template <typename T>
__global__ void __launch_bounds__(256) dcp_exchange_fold_kernel(const __grid_constant__ DcpFoldArgs a) {
  constexpr int kPack = 16 / sizeof(T);
  // the step: a kernel argument, or the word the gather launch of this step advanced (nothing writes it during this launch)
  const uint32_t epoch = a.step_counter != nullptr ? *a.step_counter : a.epoch;
  // 1. publish (one CTA; every CTA of this launch is past the producing attention launch)
  // (st.release.sys orders the partial — written by the previous launch on this stream — before the flag; a separate
  // membar.sys per thread on top of it cost ~10 us per launch in the first version)
  if (blockIdx.x == 0 && threadIdx.x < a.world) st_release_sys(a.peer_flags[threadIdx.x] + a.rank, epoch);
  // 2. wait for every rank's partial
  if (threadIdx.x < a.world) {
    const uint32_t* f = a.peer_flags[a.rank] + threadIdx.x;
    const uint64_t t0 = globaltimer_ns();
    // flags only grow; the signed difference keeps the comparison right across a 2^32 wrap
    uint32_t seen;
    while ((int32_t)((seen = ld_acquire_sys(f)) - epoch) < 0) {
      __nanosleep(64);
      if (globaltimer_ns() - t0 > a.wait_limit_ns) wait_timed_out("exchange_fold", a.rank, (int)threadIdx.x, epoch, seen);
    }
  }
  __syncthreads();

  // 3. pull + fold
  const uint32_t packs_per_head = a.head_size / kPack;
  const uint64_t total = (uint64_t)a.num_seqs * a.heads_local * packs_per_head;
  const uint64_t idx = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const uint32_t pack = idx % packs_per_head;
  const uint64_t sh = idx / packs_per_head;
  const uint32_t hl = sh % a.heads_local;
  const uint64_t seq = sh / a.heads_local;
  const uint32_t heads_all = a.world * a.heads_local;
  const uint32_t head = a.rank * a.heads_local + hl;
  const uint64_t lse_off = (uint64_t)head * a.num_seqs + seq;
  const uint64_t out_off = (seq * heads_all + head) * a.head_size + (uint64_t)pack * kPack;

  float lse[B200_DCP_MAX_WORLD];
  Vec16 v[B200_DCP_MAX_WORLD];
  float m = -INFINITY;
#pragma unroll
  for (int r = 0; r < B200_DCP_MAX_WORLD; ++r) {
    if (r < a.world) {
      float l = ld_relaxed_sys_f32(a.peer_lse[r] + lse_off);
      l = isinf(l) ? -INFINITY : l;  // merge_attn_states.cu:45-46
      lse[r] = l;
      v[r] = ld_relaxed_sys_16(reinterpret_cast<const T*>(a.peer_out[r]) + out_off);
      m = fmaxf(m, l);
    }
  }
  float acc[kPack];
#pragma unroll
  for (int i = 0; i < kPack; ++i) acc[i] = 0.f;
  float se = 0.f;
#pragma unroll
  for (int r = 0; r < B200_DCP_MAX_WORLD; ++r) {
    if (r < a.world) {
      const float w = expf(lse[r] - m);
      se += w;
      const T* e = reinterpret_cast<const T*>(&v[r]);
#pragma unroll
      for (int i = 0; i < kPack; ++i) acc[i] = __fmaf_rn(to_f32<T>(e[i]), w, acc[i]);
    }
  }
  const float inv = 1.f / se;
  Vec16 ov;
  T* oe = reinterpret_cast<T*>(&ov);
#pragma unroll
  for (int i = 0; i < kPack; ++i) oe[i] = from_f32<T>(acc[i] * inv);
  *reinterpret_cast<Vec16*>(reinterpret_cast<T*>(a.out) + (seq * a.heads_local + hl) * a.head_size + (uint64_t)pack * kPack) = ov;
  if (a.out_lse != nullptr && pack == 0) a.out_lse[(uint64_t)hl * a.num_seqs + seq] = logf(se) + m;
}

// Query all-gather of the DCP step as pushes over peer memory: rank `rank` stores its [num_seqs, heads_local, head_size]
// query slice into the gathered query [num_seqs, world*heads_local, head_size] of EVERY rank (its own included) at heads
// [rank*heads_local, ...), then the last CTA to finish publishes `epoch` to the peers' flags and waits for theirs; when
// the launch completes, this rank's gathered query is whole and the attention launch that follows on the stream may
// read it.  Replaces `get_dcp_group().all_gather(query, dim=1)` (flash_attn.py:812) and the layout permute after it.
struct DcpGatherArgs {
  void* peer_q[B200_DCP_MAX_WORLD];
  uint32_t* peer_flags[B200_DCP_MAX_WORLD];
  const void* query;
  uint32_t* cta_counter;
  int64_t q_stride;  // elements between sequences of `query`
  int32_t rank, world, num_seqs, heads_local, head_size;
  uint32_t epoch;
  uint32_t* step_counter;  // device-side step number (nullable): this launch is step *step_counter + 1 and stores it back
  uint64_t wait_limit_ns;
};

template <typename T>
__global__ void __launch_bounds__(1024) dcp_gather_query_kernel(const __grid_constant__ DcpGatherArgs a) {
  constexpr int kPack = 16 / sizeof(T);
  const uint32_t packs_per_seq = a.heads_local * a.head_size / kPack;  // the local heads of a sequence are contiguous
  const uint64_t per_dest = (uint64_t)a.num_seqs * packs_per_seq;
  const uint64_t total = per_dest * a.world;
  const uint64_t row_all = (uint64_t)a.world * a.heads_local * a.head_size;
  // Every CTA reads the counter before it is counted in below; the CTA that is counted in last advances it, so no CTA
  // of this launch can see the new value and the fold launch that follows on the stream sees nothing else.
  const uint32_t epoch = a.step_counter != nullptr ? *reinterpret_cast<volatile uint32_t*>(a.step_counter) + 1u : a.epoch;
  for (uint64_t idx = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (uint64_t)gridDim.x * blockDim.x) {
    const uint32_t dest = idx / per_dest;
    const uint64_t rem = idx % per_dest;
    const uint64_t seq = rem / packs_per_seq;
    const uint32_t pk = rem % packs_per_seq;
    const Vec16 v = *reinterpret_cast<const Vec16*>(reinterpret_cast<const T*>(a.query) + seq * a.q_stride + (uint64_t)pk * kPack);
    T* dst = reinterpret_cast<T*>(a.peer_q[dest]) + seq * row_all + (uint64_t)a.rank * a.heads_local * a.head_size +
             (uint64_t)pk * kPack;
    *reinterpret_cast<Vec16*>(dst) = v;
  }
  // One system fence per CTA, by the thread that counts the CTA in: the barrier orders every thread's stores before
  // the fence, the fence before the counter increment (release side).  The CTA that sees the full count fences AGAIN
  // after reading the counter (acquire side), so that the other CTAs' peer stores are ordered before the flag stores
  // below in the PTX memory model and not only in practice.
  __syncthreads();
  __shared__ bool last;
  if (gridDim.x == 1) {
    // One CTA (decode-sized batches): no counter round trip.  The barrier orders every thread's peer stores before
    // the fence of each signalling thread; the flag store after the fence can then be relaxed (fence + store is the
    // release pattern), which saves the second MEMBAR.SYS a st.release would issue.
    if (threadIdx.x == 0 && a.step_counter != nullptr) *a.step_counter = epoch;
    if (threadIdx.x >= a.world) return;
    asm volatile("fence.acq_rel.sys;" ::: "memory");
    st_relaxed_sys(a.peer_flags[threadIdx.x] + a.rank, epoch);
  } else {
    if (threadIdx.x == 0) {
      asm volatile("fence.acq_rel.sys;" ::: "memory");
      const uint32_t prev = atomicAdd(a.cta_counter, 1u);
      last = prev == gridDim.x - 1;
      if (last) {
        asm volatile("fence.acq_rel.sys;" ::: "memory");
        *a.cta_counter = 0;  // the next launch on this stream starts from zero
        if (a.step_counter != nullptr) *a.step_counter = epoch;
      }
    }
    __syncthreads();
    if (!last || threadIdx.x >= a.world) return;
    // ordered after the other CTAs' stores by the fence above (thread 0, then the barrier): the fence of this store's
    // own release is what orders it for the threads other than thread 0
    st_release_sys(a.peer_flags[threadIdx.x] + a.rank, epoch);
  }
  const uint32_t* f = a.peer_flags[a.rank] + threadIdx.x;
  const uint64_t t0 = globaltimer_ns();
  uint32_t seen;
  while ((int32_t)((seen = ld_acquire_sys(f)) - epoch) < 0) {
    __nanosleep(64);
    if (globaltimer_ns() - t0 > a.wait_limit_ns) wait_timed_out("gather_query", a.rank, (int)threadIdx.x, epoch, seen);
  }
}

uint64_t wait_limit_ns() {
  static const uint64_t ns = [] {
    const char* e = getenv("B200_DCP_WAIT_SECONDS");
    const long long sec = e ? atoll(e) : 120;
    return (uint64_t)(sec > 0 ? sec : 120) * 1000000000ull;
  }();
  return ns;
}
// A step number passed as a host-side kernel argument would be frozen by a CUDA graph: every `flag >= epoch` wait of a
// replay would already be satisfied and the ranks would read each other's buffers unsynchronised.  Launches in that
// mode refuse to be captured; with the step number in device memory (step_counter) nothing in the launch changes from
// step to step and capture is fine.
bool stream_is_capturing(cudaStream_t s) {
  cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(s, &st) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return st != cudaStreamCaptureStatusNone;
}
}  // namespace
}  // namespace b200

extern "C" int b200_dcp_exchange_fold(const b200_dcp_fold_params_t* p, void* stream) {
  using namespace b200;
  B200_REQUIRE(p != nullptr, B200_ERR_INVALID_ARGUMENT, "params is NULL");
  B200_REQUIRE(p->world >= 1 && p->world <= B200_DCP_MAX_WORLD, B200_ERR_UNSUPPORTED,
               "dcp_exchange_fold: world size %d outside 1..%d", p->world, B200_DCP_MAX_WORLD);
  B200_REQUIRE(p->rank >= 0 && p->rank < p->world, B200_ERR_INVALID_ARGUMENT, "dcp_exchange_fold: rank %d of %d", p->rank,
               p->world);
  B200_REQUIRE(p->num_seqs >= 0 && p->heads_local > 0 && p->head_size > 0, B200_ERR_INVALID_ARGUMENT,
               "bad sizes: seqs=%d heads_local=%d head_size=%d", p->num_seqs, p->heads_local, p->head_size);
  B200_REQUIRE(p->dtype == B200_DTYPE_F32 || p->dtype == B200_DTYPE_F16 || p->dtype == B200_DTYPE_BF16,
               B200_ERR_UNSUPPORTED, "Unsupported data type of O: %d", p->dtype);
  const int pack = 16 / dtype_bytes(p->dtype);
  B200_REQUIRE(p->head_size % pack == 0, B200_ERR_UNSUPPORTED, "headsize must be multiple of pack_size:%d", pack);
  B200_REQUIRE(p->epoch != 0 || p->step_counter != nullptr, B200_ERR_INVALID_ARGUMENT,
               "dcp_exchange_fold: epoch 0 is the initial flag value");
  B200_REQUIRE(p->peer_out && p->peer_lse && p->peer_flags && p->out, B200_ERR_INVALID_ARGUMENT, "NULL pointer");
  DcpFoldArgs a{};
  for (int r = 0; r < p->world; ++r) {
    B200_REQUIRE(p->peer_out[r] && p->peer_lse[r] && p->peer_flags[r], B200_ERR_INVALID_ARGUMENT,
                 "dcp_exchange_fold: NULL buffer of rank %d", r);
    B200_REQUIRE((reinterpret_cast<uintptr_t>(p->peer_out[r]) & 15) == 0, B200_ERR_INVALID_ARGUMENT,
                 "dcp_exchange_fold: partial of rank %d not 16-byte aligned", r);
    a.peer_out[r] = p->peer_out[r];
    a.peer_lse[r] = p->peer_lse[r];
    a.peer_flags[r] = p->peer_flags[r];
  }
  B200_REQUIRE((reinterpret_cast<uintptr_t>(p->out) & 15) == 0, B200_ERR_INVALID_ARGUMENT, "out not 16-byte aligned");
  a.out = p->out;
  a.out_lse = p->out_lse;
  a.rank = p->rank, a.world = p->world, a.num_seqs = p->num_seqs, a.heads_local = p->heads_local;
  a.head_size = p->head_size, a.epoch = p->epoch, a.step_counter = p->step_counter;
  a.wait_limit_ns = wait_limit_ns();
  B200_REQUIRE(p->step_counter != nullptr || !stream_is_capturing(static_cast<cudaStream_t>(stream)), B200_ERR_UNSUPPORTED,
               "dcp_exchange_fold cannot be captured into a CUDA graph with a host-side step number: pass step_counter");
  const uint64_t total = (uint64_t)p->num_seqs * p->heads_local * (p->head_size / pack);
  const int threads = 256;
  // at least one CTA even for an empty batch: the peers wait for this rank's flag
  const uint64_t blocks = total == 0 ? 1 : (total + threads - 1) / threads;
  B200_REQUIRE(blocks <= 0x7fffffffull, B200_ERR_UNSUPPORTED, "dcp_exchange_fold: too many elements");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  switch (p->dtype) {
    case B200_DTYPE_F32: dcp_exchange_fold_kernel<float><<<(unsigned)blocks, threads, 0, s>>>(a); break;
    case B200_DTYPE_F16: dcp_exchange_fold_kernel<__half><<<(unsigned)blocks, threads, 0, s>>>(a); break;
    default: dcp_exchange_fold_kernel<__nv_bfloat16><<<(unsigned)blocks, threads, 0, s>>>(a); break;
  }
  count_launch();
  return check_launch("dcp_exchange_fold");
}

extern "C" int b200_dcp_gather_query(const b200_dcp_gather_params_t* p, void* stream) {
  using namespace b200;
  B200_REQUIRE(p != nullptr, B200_ERR_INVALID_ARGUMENT, "params is NULL");
  B200_REQUIRE(p->world >= 1 && p->world <= B200_DCP_MAX_WORLD, B200_ERR_UNSUPPORTED,
               "dcp_gather_query: world size %d outside 1..%d", p->world, B200_DCP_MAX_WORLD);
  B200_REQUIRE(p->rank >= 0 && p->rank < p->world, B200_ERR_INVALID_ARGUMENT, "dcp_gather_query: rank %d of %d", p->rank,
               p->world);
  B200_REQUIRE(p->num_seqs >= 0 && p->heads_local > 0 && p->head_size > 0, B200_ERR_INVALID_ARGUMENT,
               "bad sizes: seqs=%d heads_local=%d head_size=%d", p->num_seqs, p->heads_local, p->head_size);
  B200_REQUIRE(p->dtype == B200_DTYPE_F32 || p->dtype == B200_DTYPE_F16 || p->dtype == B200_DTYPE_BF16,
               B200_ERR_UNSUPPORTED, "Unsupported data type of Q: %d", p->dtype);
  const int pack = 16 / dtype_bytes(p->dtype);
  B200_REQUIRE(p->head_size % pack == 0, B200_ERR_UNSUPPORTED, "headsize must be multiple of pack_size:%d", pack);
  B200_REQUIRE(p->epoch != 0 || p->step_counter != nullptr, B200_ERR_INVALID_ARGUMENT,
               "dcp_gather_query: epoch 0 is the initial flag value");
  B200_REQUIRE(p->peer_q && p->peer_flags && p->cta_counter && (p->query || p->num_seqs == 0), B200_ERR_INVALID_ARGUMENT,
               "NULL pointer");
  B200_REQUIRE(p->q_stride >= (int64_t)p->heads_local * p->head_size && p->q_stride % pack == 0 &&
                   (reinterpret_cast<uintptr_t>(p->query) & 15) == 0,
               B200_ERR_INVALID_ARGUMENT, "dcp_gather_query: query rows must be 16-byte aligned, heads contiguous");
  DcpGatherArgs a{};
  for (int r = 0; r < p->world; ++r) {
    B200_REQUIRE(p->peer_q[r] && p->peer_flags[r], B200_ERR_INVALID_ARGUMENT, "dcp_gather_query: NULL buffer of rank %d", r);
    B200_REQUIRE((reinterpret_cast<uintptr_t>(p->peer_q[r]) & 15) == 0, B200_ERR_INVALID_ARGUMENT,
                 "dcp_gather_query: gathered query of rank %d not 16-byte aligned", r);
    a.peer_q[r] = p->peer_q[r];
    a.peer_flags[r] = p->peer_flags[r];
  }
  a.query = p->query, a.cta_counter = p->cta_counter, a.q_stride = p->q_stride;
  a.rank = p->rank, a.world = p->world, a.num_seqs = p->num_seqs, a.heads_local = p->heads_local;
  a.head_size = p->head_size, a.epoch = p->epoch, a.step_counter = p->step_counter;
  a.wait_limit_ns = wait_limit_ns();
  B200_REQUIRE(p->step_counter != nullptr || !stream_is_capturing(static_cast<cudaStream_t>(stream)), B200_ERR_UNSUPPORTED,
               "dcp_gather_query cannot be captured into a CUDA graph with a host-side step number: pass step_counter");
  const uint64_t total = (uint64_t)p->num_seqs * p->heads_local * (p->head_size / pack) * p->world;
  // up to 16 packs per thread in ONE CTA of 1024 threads (decode batches): the single-CTA path has no CTA counter
  // round trip and one system fence fewer; beyond that, 64 CTAs of 256
  int threads = 256;
  uint64_t blocks = (total + threads - 1) / threads;
  blocks = blocks < 1 ? 1 : (blocks > 64 ? 64 : blocks);
  if (total <= 16 * 1024) threads = total <= 256 ? 256 : 1024, blocks = 1;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  switch (p->dtype) {
    case B200_DTYPE_F32: dcp_gather_query_kernel<float><<<(unsigned)blocks, threads, 0, s>>>(a); break;
    case B200_DTYPE_F16: dcp_gather_query_kernel<__half><<<(unsigned)blocks, threads, 0, s>>>(a); break;
    default: dcp_gather_query_kernel<__nv_bfloat16><<<(unsigned)blocks, threads, 0, s>>>(a); break;
  }
  count_launch();
  return check_launch("dcp_gather_query");
}

// Peer-mappable device memory for the partial buffers: plain cudaMalloc (one allocation = one IPC handle, offset 0),
// outside any caching allocator.
extern "C" int b200_peer_alloc(uint64_t bytes, void** ptr, void* handle64) {
  using namespace b200;
  B200_REQUIRE(ptr != nullptr && handle64 != nullptr && bytes > 0, B200_ERR_INVALID_ARGUMENT, "peer_alloc: bad arguments");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle is 64 bytes");
  void* d = nullptr;
  cudaError_t e = cudaMalloc(&d, bytes);
  B200_REQUIRE(e == cudaSuccess, B200_ERR_CUDA, "peer_alloc: cudaMalloc(%llu): %s", (unsigned long long)bytes, cudaGetErrorString(e));
  e = cudaMemset(d, 0, bytes);
  if (e == cudaSuccess) e = cudaDeviceSynchronize();  // the zeroes are in place before any stream or peer touches the flags
  if (e == cudaSuccess) e = cudaIpcGetMemHandle(static_cast<cudaIpcMemHandle_t*>(handle64), d);
  if (e != cudaSuccess) {
    cudaFree(d);
    B200_REQUIRE(false, B200_ERR_CUDA, "peer_alloc: %s", cudaGetErrorString(e));
  }
  *ptr = d;
  return B200_OK;
}

extern "C" int b200_peer_open(const void* handle64, void** ptr) {
  using namespace b200;
  B200_REQUIRE(ptr != nullptr && handle64 != nullptr, B200_ERR_INVALID_ARGUMENT, "peer_open: bad arguments");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, sizeof(h));
  void* d = nullptr;
  cudaError_t e = cudaIpcOpenMemHandle(&d, h, cudaIpcMemLazyEnablePeerAccess);
  B200_REQUIRE(e == cudaSuccess, B200_ERR_CUDA, "peer_open: cudaIpcOpenMemHandle: %s", cudaGetErrorString(e));
  *ptr = d;
  return B200_OK;
}

extern "C" int b200_peer_close(void* ptr) {
  using namespace b200;
  if (ptr == nullptr) return B200_OK;
  cudaError_t e = cudaIpcCloseMemHandle(ptr);
  B200_REQUIRE(e == cudaSuccess, B200_ERR_CUDA, "peer_close: %s", cudaGetErrorString(e));
  return B200_OK;
}

extern "C" int b200_peer_free(void* ptr) {
  using namespace b200;
  if (ptr == nullptr) return B200_OK;
  cudaError_t e = cudaFree(ptr);
  B200_REQUIRE(e == cudaSuccess, B200_ERR_CUDA, "peer_free: %s", cudaGetErrorString(e));
  return B200_OK;
}
