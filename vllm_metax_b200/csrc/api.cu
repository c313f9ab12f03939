// C ABI of the paged-attention decode path: validation, kernel-family
// dispatch, error plumbing.  See include/b200_paged_attn.h for the contract
// and the reference entry points each function replaces
// (csrc/attention/paged_attention_v1.cu:160-182, paged_attention_v2.cu:167-192).
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <atomic>

#include "common.cuh"

namespace b200 {

namespace {
thread_local char g_error[512] = "";
thread_local int g_forced_variant = B200_VARIANT_AUTO;
thread_local int g_last_variant = B200_VARIANT_AUTO;
std::atomic<uint64_t> g_launches{0};
}  // namespace

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_error, sizeof(g_error), fmt, ap);
  va_end(ap);
}

void count_launch(int n) { g_launches.fetch_add((uint64_t)n, std::memory_order_relaxed); }

int check_launch(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("%s: kernel launch failed: %s", what, cudaGetErrorString(e));
    return B200_ERR_CUDA;
  }
  return B200_OK;
}

int device_sm_count() {
  static std::atomic<int> cache[64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) {
    cudaGetLastError();
    return 148;
  }
  int v = cache[dev].load(std::memory_order_relaxed);
  if (v > 0) return v;
  if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) {
    cudaGetLastError();
    v = 148;
  }
  cache[dev].store(v, std::memory_order_relaxed);
  return v;
}

namespace {

bool head_size_supported(int d) {
  // csrc/attention/paged_attention_v1.cu:90-124
  switch (d) {
    case 32: case 64: case 80: case 96: case 112: case 120: case 128: case 192: case 256: return true;
    default: return false;
  }
}

int validate_attn(const b200_paged_attn_params_t* p, bool v2) {
  B200_REQUIRE(p != nullptr, B200_ERR_INVALID_ARGUMENT, "params is NULL");
  B200_REQUIRE(p->num_seqs >= 0 && p->num_heads > 0 && p->num_kv_heads > 0, B200_ERR_INVALID_ARGUMENT,
               "bad sizes: seqs=%d heads=%d kv_heads=%d", p->num_seqs, p->num_heads, p->num_kv_heads);
  B200_REQUIRE(p->num_heads % p->num_kv_heads == 0, B200_ERR_INVALID_ARGUMENT,
               "num_heads %d is not a multiple of num_kv_heads %d", p->num_heads, p->num_kv_heads);
  B200_REQUIRE(p->dtype == B200_DTYPE_F32 || p->dtype == B200_DTYPE_F16 || p->dtype == B200_DTYPE_BF16,
               B200_ERR_UNSUPPORTED, "Unsupported data type: %d", p->dtype);
  B200_REQUIRE(p->kv_dtype == B200_KV_AUTO || p->kv_dtype == B200_KV_FP8_E4M3, B200_ERR_UNSUPPORTED,
               "Unsupported data type of kv cache: %d", p->kv_dtype);
  B200_REQUIRE(head_size_supported(p->head_size), B200_ERR_UNSUPPORTED, "Unsupported head size: %d", p->head_size);
  B200_REQUIRE(p->block_size == 8 || p->block_size == 16 || p->block_size == 32, B200_ERR_UNSUPPORTED,
               "Unsupported block size: %d", p->block_size);
  B200_REQUIRE(p->blocksparse_vert_stride <= 1, B200_ERR_UNSUPPORTED,
               "block-sparse attention (blocksparse_vert_stride=%d) is not supported", p->blocksparse_vert_stride);
  if (p->num_seqs == 0) return B200_OK;
  B200_REQUIRE(p->out && p->query && p->key_cache && p->value_cache && p->block_tables && p->seq_lens,
               B200_ERR_INVALID_ARGUMENT, "NULL tensor pointer");
  B200_REQUIRE(p->max_seq_len >= 0 &&
                   (int64_t)p->max_seq_len <= (int64_t)p->max_num_blocks_per_seq * p->block_size,
               B200_ERR_INVALID_ARGUMENT, "max_seq_len %d exceeds block table capacity %d x %d", p->max_seq_len,
               p->max_num_blocks_per_seq, p->block_size);
  const int eb = p->kv_dtype == B200_KV_FP8_E4M3 ? 1 : dtype_bytes(p->dtype);
  const int x = 16 / eb;
  B200_REQUIRE(p->head_size % x == 0, B200_ERR_UNSUPPORTED, "head size %d is not a multiple of x=%d", p->head_size, x);
  if (p->kv_dtype != B200_KV_AUTO)
    B200_REQUIRE(p->k_scale && p->v_scale, B200_ERR_INVALID_ARGUMENT, "fp8 cache needs k_scale/v_scale");
  B200_REQUIRE((p->kv_block_stride * eb) % 16 == 0 && (p->kv_head_stride * eb) % 16 == 0 &&
                   reinterpret_cast<uintptr_t>(p->key_cache) % 16 == 0 &&
                   reinterpret_cast<uintptr_t>(p->value_cache) % 16 == 0,
               B200_ERR_UNSUPPORTED, "KV cache must be 16-byte aligned with 16-byte-multiple block/head strides");
  if (v2)
    B200_REQUIRE(p->exp_sums && p->max_logits && p->tmp_out, B200_ERR_INVALID_ARGUMENT,
                 "paged_attention_v2 needs exp_sums / max_logits / tmp_out");
  return B200_OK;
}

// returns B200_VARIANT_* to run, or a negative status on failure
int pick_variant(const b200_paged_attn_params_t& p) {
  const char* why = nullptr;
  const bool ok = stream_variant_eligible(p, &why);
  const char* why_u = nullptr;
  bool ok_u = umma_fp8_variant_eligible(p, &why_u);
  if (ok_u && p.out_lse != nullptr) {
    ok_u = false;
    why_u = "the tcgen05 fp8 kernel has no LSE output";
  }
  if (g_forced_variant == B200_VARIANT_UMMA_FP8) {
    if (!ok_u) {
      set_error("tcgen05 fp8 variant forced but not eligible: %s", why_u);
      return -B200_ERR_UNSUPPORTED;
    }
    return B200_VARIANT_UMMA_FP8;
  }
  if (g_forced_variant == B200_VARIANT_STREAM_MMA) {
    if (!ok) {
      set_error("stream variant forced but not eligible: %s", why);
      return -B200_ERR_UNSUPPORTED;
    }
    return B200_VARIANT_STREAM_MMA;
  }
  if (g_forced_variant == B200_VARIANT_GENERIC) return B200_VARIANT_GENERIC;
  // The tcgen05 fp8 kernel is parity-green but its tensor pipe is the bottleneck (8 M64 N16 K32 MMAs per 32-token tile,
  // 78 % tensor-active at 3.2 TB/s; profiles/README.md), so the mma.sync kernel (5.8 TB/s) stays the default for fp8.
  // Opt in with B200_ENABLE_UMMA=1 or b200_set_attention_variant(B200_VARIANT_UMMA_FP8).
  static const bool use_umma = getenv("B200_ENABLE_UMMA") != nullptr;
  if (ok_u && use_umma) return B200_VARIANT_UMMA_FP8;
  return ok ? B200_VARIANT_STREAM_MMA : B200_VARIANT_GENERIC;
}

// Workspace of the v1 op: [flat-schedule region of the streaming kernel: kFlatWorkspaceBytes, see common.cuh]
// [generic kernel's split-KV partials: o | m | l per (seq, head, 512-token partition)].  The first region has a
// fixed size and position, so its arrival counters stay all-zero between launches whatever shapes share it.
uint64_t generic_split_bytes(const b200_paged_attn_params_t& p) {
  const int64_t P = (p.max_seq_len + B200_PARTITION_SIZE - 1) / B200_PARTITION_SIZE;
  if (P <= 1) return 0;
  return (uint64_t)p.num_seqs * p.num_heads * P * (p.head_size + 2) * sizeof(float);
}

// Sequences of length 0 own no tile in the streaming schedules, so nothing would write their rows: out = 0, and
// lse = -inf when the caller wants the LSE (to merge this result with another one).
__global__ void init_empty_rows_kernel(const int32_t* seq_lens, void* out, float* lse, int S, int H, int row_bytes) {
  const int s = blockIdx.x;
  if (seq_lens[s] > 0) return;
  uint32_t* o = reinterpret_cast<uint32_t*>(static_cast<char*>(out) + (int64_t)s * row_bytes);
  for (int i = threadIdx.x; i < row_bytes / 4; i += blockDim.x) o[i] = 0u;
  if (lse != nullptr)
    for (int h = threadIdx.x; h < H; h += blockDim.x) lse[(int64_t)h * S + s] = -INFINITY;
}

// The out rows are zeroed whether or not an LSE is asked for: the reference kernels write zeros for an empty
// sequence (acc = 0 times 1/(0 + 1e-6), csrc/attention/attention_kernels.cuh:331), and CUDA-graph-padded decode batches
// carry such rows.
int init_empty_rows(const b200_paged_attn_params_t& p, cudaStream_t s) {
  const int row_bytes = p.num_heads * p.head_size * dtype_bytes(p.dtype);  // multiple of 4: head sizes are even
  init_empty_rows_kernel<<<p.num_seqs, 128, 0, s>>>(p.seq_lens, p.out, p.out_lse, p.num_seqs, p.num_heads, row_bytes);
  count_launch();
  return check_launch("paged_attention (empty rows)");
}

}  // namespace
}  // namespace b200

using namespace b200;

extern "C" uint64_t b200_paged_attention_workspace_bytes(const b200_paged_attn_params_t* p) {
  if (p == nullptr || p->num_seqs <= 0 || p->num_heads <= 0 || p->num_kv_heads <= 0 || p->max_seq_len <= 0) return 0;
  return kFlatWorkspaceBytes + generic_split_bytes(*p);
}

extern "C" int b200_paged_attention_workspace_reset(void* workspace, uint64_t workspace_bytes, void* stream) {
  B200_REQUIRE(workspace != nullptr && workspace_bytes >= kFlatCounterBytes, B200_ERR_WORKSPACE,
               "workspace_reset: workspace missing or smaller than its counter region (%llu bytes)",
               (unsigned long long)kFlatCounterBytes);
  cudaError_t e = cudaMemsetAsync(workspace, 0, kFlatCounterBytes, static_cast<cudaStream_t>(stream));
  if (e != cudaSuccess) {
    set_error("workspace_reset: cudaMemsetAsync failed: %s", cudaGetErrorString(e));
    return B200_ERR_CUDA;
  }
  return B200_OK;
}

extern "C" int b200_paged_attention_v1(const b200_paged_attn_params_t* p, void* stream) {
  int rc = validate_attn(p, false);
  if (rc != B200_OK) return rc;
  if (p->num_seqs == 0) return B200_OK;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (p->max_seq_len == 0) return init_empty_rows(*p, s);
  const int variant = pick_variant(*p);
  if (variant < 0) return -variant;
  // the streaming and generic kernels write the rows of empty sequences themselves; the tcgen05 kernel does not
  if (variant == B200_VARIANT_UMMA_FP8 && (rc = init_empty_rows(*p, s)) != B200_OK) return rc;
  if (variant == B200_VARIANT_STREAM_MMA || variant == B200_VARIANT_UMMA_FP8) {
    B200_REQUIRE(p->workspace != nullptr && p->workspace_bytes >= kFlatWorkspaceBytes, B200_ERR_WORKSPACE,
                 "paged_attention_v1 needs %llu bytes of workspace, got %llu", (unsigned long long)kFlatWorkspaceBytes,
                 (unsigned long long)(p->workspace ? p->workspace_bytes : 0));
    B200_REQUIRE(reinterpret_cast<uintptr_t>(p->workspace) % 16 == 0, B200_ERR_WORKSPACE,
                 "workspace must be 16-byte aligned");
    // one launch: split pairs are merged in the kernel
    rc = variant == B200_VARIANT_UMMA_FP8 ? launch_attention_umma_fp8_flat(*p, p->workspace, s)
                                          : launch_attention_stream_flat(*p, p->workspace, s);
    if (rc == B200_OK) g_last_variant = variant;
    return rc;
  }
  SplitOut so;
  so.chunk = B200_PARTITION_SIZE;  // the generic kernel's shared-memory logit buffer
  so.num_partitions = (p->max_seq_len + so.chunk - 1) / so.chunk;
  so.normalized = 0;
  so.m = so.l = nullptr;
  so.o = nullptr;
  if (so.num_partitions > 1) {
    const uint64_t need = kFlatWorkspaceBytes + generic_split_bytes(*p);
    B200_REQUIRE(p->workspace != nullptr && p->workspace_bytes >= need, B200_ERR_WORKSPACE,
                 "paged_attention_v1 needs %llu bytes of workspace, got %llu", (unsigned long long)need,
                 (unsigned long long)(p->workspace ? p->workspace_bytes : 0));
    B200_REQUIRE(reinterpret_cast<uintptr_t>(p->workspace) % 16 == 0, B200_ERR_WORKSPACE,
                 "workspace must be 16-byte aligned");
    const int64_t n = (int64_t)p->num_seqs * p->num_heads * so.num_partitions;
    float* base = reinterpret_cast<float*>(static_cast<char*>(p->workspace) + kFlatWorkspaceBytes);
    so.o = base;
    so.m = base + n * p->head_size;
    so.l = so.m + n;
  }
  rc = launch_attention_generic(*p, so, s);
  if (rc != B200_OK) return rc;
  g_last_variant = variant;
  return launch_split_reduce(*p, so, s);
}

extern "C" int b200_paged_attention_flash(const b200_paged_attn_params_t* p, void* stream) {
  B200_REQUIRE(p != nullptr, B200_ERR_INVALID_ARGUMENT, "params is NULL");
  B200_REQUIRE(p->num_seqs >= 0 && p->num_heads > 0 && p->num_kv_heads > 0, B200_ERR_INVALID_ARGUMENT,
               "bad sizes: seqs=%d heads=%d kv_heads=%d", p->num_seqs, p->num_heads, p->num_kv_heads);
  B200_REQUIRE(p->num_heads % p->num_kv_heads == 0, B200_ERR_INVALID_ARGUMENT,
               "num_heads %d is not a multiple of num_kv_heads %d", p->num_heads, p->num_kv_heads);
  B200_REQUIRE(p->blocksparse_vert_stride <= 1, B200_ERR_UNSUPPORTED, "block-sparse attention is not supported");
  const char* why = nullptr;
  B200_REQUIRE(stream_flash_eligible(*p, &why), B200_ERR_UNSUPPORTED, "paged_attention_flash: %s", why);
  if (p->num_seqs == 0) return B200_OK;
  B200_REQUIRE(p->out && p->query && p->key_cache && p->value_cache && p->block_tables && p->seq_lens,
               B200_ERR_INVALID_ARGUMENT, "NULL tensor pointer");
  B200_REQUIRE(p->max_seq_len >= 0 && (int64_t)p->max_num_blocks_per_seq * p->block_size >= p->max_seq_len,
               B200_ERR_INVALID_ARGUMENT, "max_seq_len %d exceeds the block table (%d blocks of %d)", p->max_seq_len,
               p->max_num_blocks_per_seq, p->block_size);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (p->max_seq_len == 0) return init_empty_rows(*p, s);
  int rc;
  B200_REQUIRE(p->workspace != nullptr && p->workspace_bytes >= kFlatWorkspaceBytes, B200_ERR_WORKSPACE,
               "paged_attention_flash needs %llu bytes of workspace, got %llu", (unsigned long long)kFlatWorkspaceBytes,
               (unsigned long long)(p->workspace ? p->workspace_bytes : 0));
  B200_REQUIRE(reinterpret_cast<uintptr_t>(p->workspace) % 16 == 0, B200_ERR_WORKSPACE,
               "workspace must be 16-byte aligned");
  rc = launch_attention_stream_flash(*p, p->workspace, s);
  if (rc == B200_OK) g_last_variant = B200_VARIANT_STREAM_MMA;
  return rc;
}

extern "C" int b200_paged_attention_v2(const b200_paged_attn_params_t* p, void* stream) {
  int rc = validate_attn(p, true);
  if (rc != B200_OK) return rc;
  if (p->num_seqs == 0) return B200_OK;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (p->max_seq_len == 0) return init_empty_rows(*p, s);
  const int variant = pick_variant(*p);
  if (variant < 0) return -variant;
  if (variant == B200_VARIANT_UMMA_FP8 && (rc = init_empty_rows(*p, s)) != B200_OK) return rc;
  // the caller sized exp_sums / max_logits / tmp_out for 512-token partitions
  // (csrc/attention/paged_attention_v2.cu:45,82)
  SplitOut so;
  so.chunk = B200_PARTITION_SIZE;
  so.num_partitions = (p->max_seq_len + so.chunk - 1) / so.chunk;
  so.normalized = 1;
  so.m = p->max_logits;
  so.l = p->exp_sums;
  so.o = p->tmp_out;
  rc = variant == B200_VARIANT_UMMA_FP8     ? launch_attention_umma_fp8(*p, so, s)
       : variant == B200_VARIANT_STREAM_MMA ? launch_attention_stream(*p, so, s)
                                            : launch_attention_generic(*p, so, s);
  if (rc != B200_OK) return rc;
  g_last_variant = variant;
  return launch_split_reduce(*p, so, s);
}

extern "C" int b200_kv_dtype_from_string(const char* s) {
  if (s == nullptr) return -1;
  if (strcmp(s, "auto") == 0) return B200_KV_AUTO;
  if (strcmp(s, "fp8") == 0 || strcmp(s, "fp8_e4m3") == 0) return B200_KV_FP8_E4M3;
  return -1;
}

extern "C" const char* b200_last_error(void) { return g_error; }
extern "C" int b200_abi_version(void) { return B200_PAGED_ATTN_ABI_VERSION; }

extern "C" int b200_set_attention_variant(int variant) {
  if (variant != B200_VARIANT_AUTO && variant != B200_VARIANT_GENERIC && variant != B200_VARIANT_STREAM_MMA &&
      variant != B200_VARIANT_UMMA_FP8) {
    set_error("unknown attention variant %d", variant);
    return B200_ERR_INVALID_ARGUMENT;
  }
  g_forced_variant = variant;
  return B200_OK;
}
extern "C" int b200_last_attention_variant(void) { return g_last_variant; }
extern "C" int b200_set_stream_dynamic(int min_chunk_tiles) {
  if (min_chunk_tiles > 64) {
    set_error("stream_dynamic: minimum chunk of %d tiles (max 64)", min_chunk_tiles);
    return B200_ERR_INVALID_ARGUMENT;
  }
  set_stream_dynamic(min_chunk_tiles);
  return B200_OK;
}
extern "C" uint64_t b200_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }
