// b200_paged_attention_v1_static: paged_attention_v1 for callers that cannot allocate device memory (CUDA-graph capture,
// tight serving loops, harnesses that forbid cudaMalloc).  Same kernels as b200_paged_attention_v1; the only difference
// is where the split-KV scratch lives: not in a caller-provided workspace but in the file-scope device array below, which
// every GPU that loads this library holds until the process exits.
//
// No counterpart in the reference.
//
// Rules of the shared array (include/b200_paged_attn.h repeats them):
//   - one array per device, shared by every call on that device: calls that may overlap on the GPU (two streams, or two
//     host threads on different streams) race on it.  Serialise them, or use b200_paged_attention_v1 with one workspace
//     per stream.
//   - the array keeps whatever the last call left.  The arrival counters in its first kFlatCounterBytes must be zero
//     when a launch starts, so every call re-zeroes them on `stream` ahead of the kernel (cudaMemsetAsync: no
//     allocation, capturable); the partial-result slots behind them are written before they are read.
//   - shapes whose scratch exceeds B200_STATIC_SCRATCH_BYTES are refused with B200_ERR_WORKSPACE.
// This entry point calls no cudaMalloc / cudaFree and no allocating library routine.
#include <mutex>

#include "common.cuh"

namespace {

constexpr uint64_t kStaticBytes = B200_STATIC_SCRATCH_BYTES;
static_assert(kStaticBytes >= b200::kFlatWorkspaceBytes, "the static scratch must hold the flat schedule's workspace");
static_assert(kStaticBytes % 256 == 0, "keep the array a whole number of 256-byte lines");

// In host code this name is only a handle for cudaGetSymbolAddress: it is NOT a device address.
static __device__ __align__(256) unsigned char g_static_scratch[kStaticBytes];

constexpr int kMaxDevices = 64;
std::mutex g_mu;
void* g_scratch_ptr[kMaxDevices] = {};

// Device address of the array on the CURRENT device, looked up once per device.
int static_scratch_ptr(void** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  B200_REQUIRE(e == cudaSuccess, B200_ERR_CUDA, "paged_attention_v1_static: cudaGetDevice: %s", cudaGetErrorString(e));
  B200_REQUIRE(dev >= 0 && dev < kMaxDevices, B200_ERR_UNSUPPORTED, "paged_attention_v1_static: device ordinal %d >= %d",
               dev, kMaxDevices);
  std::lock_guard<std::mutex> lock(g_mu);
  if (g_scratch_ptr[dev] == nullptr) {
    void* ptr = nullptr;
    e = cudaGetSymbolAddress(&ptr, g_static_scratch);
    B200_REQUIRE(e == cudaSuccess && ptr != nullptr, B200_ERR_CUDA,
                 "paged_attention_v1_static: cudaGetSymbolAddress on device %d: %s", dev, cudaGetErrorString(e));
    g_scratch_ptr[dev] = ptr;
  }
  *out = g_scratch_ptr[dev];
  return B200_OK;
}

}  // namespace

extern "C" uint64_t b200_paged_attention_static_scratch_bytes(void) { return kStaticBytes; }

extern "C" int b200_paged_attention_v1_static(const b200_paged_attn_params_t* p, void* stream) {
  B200_REQUIRE(p != nullptr, B200_ERR_INVALID_ARGUMENT, "params is NULL");
  const uint64_t need = b200_paged_attention_workspace_bytes(p);
  B200_REQUIRE(need <= kStaticBytes, B200_ERR_WORKSPACE,
               "paged_attention_v1_static: this shape needs %llu bytes of scratch, the static array holds %llu "
               "(B200_STATIC_SCRATCH_BYTES); use b200_paged_attention_v1 with a caller-allocated workspace",
               (unsigned long long)need, (unsigned long long)kStaticBytes);
  void* scratch = nullptr;
  int rc = static_scratch_ptr(&scratch);
  if (rc != B200_OK) return rc;
  if (need != 0 && p->num_seqs > 0 && p->max_seq_len > 0) {
    rc = b200_paged_attention_workspace_reset(scratch, kStaticBytes, stream);
    if (rc != B200_OK) return rc;
  }
  b200_paged_attn_params_t q = *p;
  q.workspace = scratch;
  q.workspace_bytes = kStaticBytes;
  return b200_paged_attention_v1(&q, stream);
}
