// KV-cache block movers and gathers around the decode path:
//   copy_blocks                      (reference csrc/cache_kernels.cu:64-92, :113-163)
//   swap_blocks                      (:18-60)
//   gather_and_maybe_dequant_cache   (:618-773)
//   cp_gather_cache                  (:775-839, :1075-1134)
// All of it is byte movement bound by HBM (or PCIe/NVLink-C2C for swaps), so the
// kernels are 16-byte vectorised copies with the grid spread over every SM; the
// FP8 gather fuses the e4m3 -> dtype dequantisation (float(e4m3) * scale).
#include <string.h>

#include <algorithm>

#include <stdlib.h>

#include "common.cuh"

namespace b200 {
namespace {
inline int env_int_or(const char* name, int dflt) {
  const char* v = getenv(name);
  return v != nullptr && *v ? atoi(v) : dflt;
}

inline bool aligned(const void* q, size_t a) { return (reinterpret_cast<uintptr_t>(q) % a) == 0; }

// ---- copy_blocks ---------------------------------------------------------------------------------------
// The per-layer cache pointers ride in the kernel parameter block (<= kLayersPerLaunch layers per launch),
// so nothing is staged through device memory and the call neither synchronises nor allocates.
constexpr int kLayersPerLaunch = 128;
struct CopyBlocksArgs {
  uint8_t* key[kLayersPerLaunch];
  uint8_t* value[kLayersPerLaunch];
  const int64_t* mapping;
  int64_t block_bytes;
};

// grid (pairs, layers * 2, slices): one CTA copies one slice of one block of the K or the V cache of a layer.
// NC = 2: key and value cache per layer (copy_blocks); NC = 1: one joint cache per layer (copy_blocks_mla)
template <bool VEC, int NC>
__global__ void __launch_bounds__(256) copy_blocks_kernel(const __grid_constant__ CopyBlocksArgs a) {
  const int pair = blockIdx.x;
  const int layer = blockIdx.y / NC;
  uint8_t* base = (NC == 2 && (blockIdx.y & 1)) ? a.value[layer] : a.key[layer];
  const int64_t src = a.mapping[2 * pair], dst = a.mapping[2 * pair + 1];
  const uint8_t* s = base + src * a.block_bytes;
  uint8_t* d = base + dst * a.block_bytes;
  if constexpr (VEC) {
    const int64_t n = a.block_bytes >> 4;
    for (int64_t i = (int64_t)blockIdx.z * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.z * blockDim.x)
      reinterpret_cast<Vec16*>(d)[i] = ld_global_16_coherent(s + (i << 4));  // src and dst blocks share the allocation
  } else {
    for (int64_t i = (int64_t)blockIdx.z * blockDim.x + threadIdx.x; i < a.block_bytes;
         i += (int64_t)gridDim.z * blockDim.x)
      d[i] = s[i];
  }
}

// ---- gather ----------------------------------------------------------------------------------------------
// One warp per destination row (token).  grid (batch, splits): a CTA of 8 warps walks the tokens of its
// slice of one sequence; the block-table entry is read once per token by every lane (broadcast).
//   MODE 0: same-width byte move, 16-byte vectors      MODE 1: same-width byte move, scalar elements
//   MODE 2: e4m3 -> T dequantisation, one 16-byte store per lane step   MODE 3: same, scalar
template <typename T, int MODE>
__global__ void __launch_bounds__(256) gather_cache_kernel(const b200_gather_cache_params_t p, const int elem_bytes) {
  const int b = blockIdx.x;
  const int seq_start = p.cu_seq_lens[b];
  const int seq_len = p.cu_seq_lens[b + 1] - seq_start;
  const int per = (seq_len + gridDim.y - 1) / gridDim.y;
  const int t0 = blockIdx.y * per;
  const int t1 = min(seq_len, t0 + per);
  if (t0 >= t1) return;
  int first = 0;
  if (p.seq_starts != nullptr) {
    first = p.seq_starts[b];
    if (!p.token_granular_starts) first = first / p.block_size * p.block_size;  // cache_kernels.cu:660-662
  }
  const int32_t* table = p.block_table + (int64_t)b * p.block_table_stride;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  const float scale = (MODE >= 2) ? *p.scale : 0.f;
  // A warp moves U rows at a time: the U loads of a step are independent, which keeps enough bytes in flight per SM
  // (one row per warp leaves the fp8 variant latency-bound, 256 B per warp-load: 3.6 -> 4.9 TB/s with four).
  constexpr int U = (MODE == 2) ? 4 : 1;  // measured: 2 rows per step slow the same-width copy down (5.5 -> 4.8 TB/s)
  for (int t = t0 + warp; t < t1; t += nwarps * U) {
    int64_t src_elt[U], dst_elt[U];
    bool ok[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int tu = t + u * nwarps;
      ok[u] = tu < t1;
      const int pos = first + (ok[u] ? tu : t);
      const int64_t blk = table[pos / p.block_size];
      const int off = pos % p.block_size;
      src_elt[u] = blk * p.cache_block_stride + (int64_t)off * p.cache_entry_stride;
      dst_elt[u] = (int64_t)(seq_start + tu) * p.dst_entry_stride;
    }
    if constexpr (MODE == 0) {
      const int n = (p.entry_size * elem_bytes) >> 4;
      for (int i = lane; i < n; i += 32) {
        Vec16 v[U];
#pragma unroll
        for (int u = 0; u < U; ++u)
          if (ok[u]) v[u] = ld_global_16(reinterpret_cast<const uint8_t*>(p.src_cache) + src_elt[u] * elem_bytes + (i << 4));
#pragma unroll
        for (int u = 0; u < U; ++u)
          if (ok[u]) reinterpret_cast<Vec16*>(reinterpret_cast<uint8_t*>(p.dst) + dst_elt[u] * elem_bytes)[i] = v[u];
      }
    } else if constexpr (MODE == 1) {
      const T* s = reinterpret_cast<const T*>(p.src_cache) + src_elt[0];
      T* d = reinterpret_cast<T*>(p.dst) + dst_elt[0];
      for (int i = lane; i < p.entry_size; i += 32) d[i] = s[i];
    } else if constexpr (MODE == 2) {
      // each lane turns kPer source bytes into one 16-byte store: a warp reads 32*kPer contiguous bytes and writes
      // 512 contiguous bytes per row and step.  Pairs of bytes go through cvt.rn.f16x2.e4m3x2 (exact), the scale is
      // applied in fp32 and the result rounded once to T — the same value as dequant_e4m3<T>.
      constexpr int kPer = 16 / sizeof(T);
      for (int i = lane * kPer; i < p.entry_size; i += 32 * kPer) {
        uint32_t raw[U][kPer / 4];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          if (!ok[u]) continue;
          const uint8_t* s = reinterpret_cast<const uint8_t*>(p.src_cache) + src_elt[u] + i;
          if constexpr (kPer == 8) {
            const uint2 v = __ldg(reinterpret_cast<const uint2*>(s));
            raw[u][0] = v.x; raw[u][1] = v.y;
          } else {
            raw[u][0] = __ldg(reinterpret_cast<const uint32_t*>(s));
          }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          if (!ok[u]) continue;
          Vec16 o;
          T* oe = reinterpret_cast<T*>(&o);
#pragma unroll
          for (int w = 0; w < kPer / 4; ++w) {
            uint32_t lo, hi;  // f16x2 of bytes (0,1) and (2,3)
            asm("{\n.reg .b16 l, h;\nmov.b32 {l, h}, %2;\ncvt.rn.f16x2.e4m3x2 %0, l;\ncvt.rn.f16x2.e4m3x2 %1, h;\n}"
                : "=r"(lo), "=r"(hi) : "r"(raw[u][w]));
            const float2 f0 = __half22float2(*reinterpret_cast<const __half2*>(&lo));
            const float2 f1 = __half22float2(*reinterpret_cast<const __half2*>(&hi));
            oe[4 * w + 0] = from_f32<T>(__fmul_rn(f0.x, scale));
            oe[4 * w + 1] = from_f32<T>(__fmul_rn(f0.y, scale));
            oe[4 * w + 2] = from_f32<T>(__fmul_rn(f1.x, scale));
            oe[4 * w + 3] = from_f32<T>(__fmul_rn(f1.y, scale));
          }
          *reinterpret_cast<Vec16*>(reinterpret_cast<T*>(p.dst) + dst_elt[u] + i) = o;
        }
      }
    } else {
      const uint8_t* s = reinterpret_cast<const uint8_t*>(p.src_cache) + src_elt[0];
      T* d = reinterpret_cast<T*>(p.dst) + dst_elt[0];
      for (int i = lane; i < p.entry_size; i += 32) d[i] = dequant_e4m3<T>(s[i], scale);
    }
  }
}

template <typename T, bool FP8>
void launch_gather(const b200_gather_cache_params_t& p, bool vec, int elem_bytes, dim3 grid, cudaStream_t s) {
  if constexpr (FP8) {
    if (vec) gather_cache_kernel<T, 2><<<grid, 256, 0, s>>>(p, elem_bytes);
    else gather_cache_kernel<T, 3><<<grid, 256, 0, s>>>(p, elem_bytes);
  } else {
    if (vec) gather_cache_kernel<T, 0><<<grid, 256, 0, s>>>(p, elem_bytes);
    else gather_cache_kernel<T, 1><<<grid, 256, 0, s>>>(p, elem_bytes);
  }
}

}  // namespace
}  // namespace b200

namespace b200 {
namespace {
template <int NC>
int copy_blocks_impl(void* const* key_caches, void* const* value_caches, int32_t num_layers, const int64_t* block_mapping,
                     int32_t num_pairs, int64_t block_bytes, void* stream, const char* what) {
  B200_REQUIRE(num_layers >= 0 && num_pairs >= 0 && block_bytes >= 0, B200_ERR_INVALID_ARGUMENT,
               "bad sizes: layers=%d pairs=%d block_bytes=%lld", num_layers, num_pairs, (long long)block_bytes);
  if (num_layers == 0 || num_pairs == 0 || block_bytes == 0) return B200_OK;  // cache_kernels.cu:118-120
  B200_REQUIRE(key_caches && (NC == 1 || value_caches) && block_mapping, B200_ERR_INVALID_ARGUMENT, "NULL pointer");
  B200_REQUIRE(num_pairs <= 0x7fffffff && num_layers <= 32767, B200_ERR_UNSUPPORTED, "too many layers / pairs");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  for (int l0 = 0; l0 < num_layers; l0 += kLayersPerLaunch) {
    const int nl = std::min(kLayersPerLaunch, num_layers - l0);
    CopyBlocksArgs a;
    memset(&a, 0, sizeof(a));
    bool vec = block_bytes % 16 == 0;
    for (int l = 0; l < nl; ++l) {
      B200_REQUIRE(key_caches[l0 + l] && (NC == 1 || value_caches[l0 + l]), B200_ERR_INVALID_ARGUMENT,
                   "NULL cache pointer");
      a.key[l] = static_cast<uint8_t*>(key_caches[l0 + l]);
      vec = vec && aligned(a.key[l], 16);
      if (NC == 2) {
        a.value[l] = static_cast<uint8_t*>(value_caches[l0 + l]);
        vec = vec && aligned(a.value[l], 16);
      }
    }
    a.mapping = block_mapping;
    a.block_bytes = block_bytes;
    // slices: enough CTAs to cover the machine a few times over, each moving >= 4 KiB
    const int64_t ctas = (int64_t)num_pairs * nl * NC;
    int slices = (int)((4 * (int64_t)device_sm_count() + ctas - 1) / ctas);
    const int max_slices = (int)std::min<int64_t>(65535, std::max<int64_t>(1, block_bytes / 4096));
    slices = std::max(1, std::min(slices, max_slices));
    dim3 grid(num_pairs, nl * NC, slices);
    if (vec) copy_blocks_kernel<true, NC><<<grid, 256, 0, s>>>(a);
    else copy_blocks_kernel<false, NC><<<grid, 256, 0, s>>>(a);
    count_launch();
    const int rc = check_launch(what);
    if (rc != B200_OK) return rc;
  }
  return B200_OK;
}

// ---- concat_and_cache_mla ------------------------------------------------------------------------------------
// Reference: csrc/cache_kernels.cu:346-389.  One CTA per token: entry = [kv_c (kv_lora_rank) | k_pe (pe_dim)] written at
// slot_mapping[token]; 16-byte source vectors when both parts and the entry are vector-aligned.
template <typename T, bool FP8, bool VEC>
__global__ void __launch_bounds__(256) concat_and_cache_mla_kernel(const b200_concat_mla_params_t p) {
  const int64_t token = blockIdx.x;
  const int64_t slot = p.slot_mapping[token];
  if (slot < 0) return;  // padded token (:364-367)
  const int64_t blk = slot / p.block_size, off = slot % p.block_size;
  const float scale = FP8 ? *p.scale : 0.f;
  constexpr int CB = FP8 ? 1 : (int)sizeof(T);
  constexpr int N = 16 / sizeof(T);
  uint8_t* dst = reinterpret_cast<uint8_t*>(p.kv_cache) + (blk * p.block_stride + off * p.entry_stride) * CB;
  const T* a = reinterpret_cast<const T*>(p.kv_c) + token * p.kv_c_stride;
  const T* b = reinterpret_cast<const T*>(p.k_pe) + token * p.k_pe_stride;
  const int total = p.kv_lora_rank + p.pe_dim;
  if constexpr (VEC) {
    for (int i = threadIdx.x * N; i < total; i += blockDim.x * N) {
      const Vec16 v = *reinterpret_cast<const Vec16*>(i < p.kv_lora_rank ? a + i : b + (i - p.kv_lora_rank));
      if constexpr (!FP8) {
        *reinterpret_cast<Vec16*>(dst + (int64_t)i * CB) = v;
      } else {
        const T* e = reinterpret_cast<const T*>(&v);
        alignas(8) uint8_t q[N];
#pragma unroll
        for (int j = 0; j < N; ++j) q[j] = quant_e4m3<T>(e[j], scale);
        if constexpr (N == 8) *reinterpret_cast<uint2*>(dst + i) = *reinterpret_cast<const uint2*>(q);
        else *reinterpret_cast<uint32_t*>(dst + i) = *reinterpret_cast<const uint32_t*>(q);
      }
    }
  } else {
    for (int i = threadIdx.x; i < total; i += blockDim.x) {
      const T x = i < p.kv_lora_rank ? a[i] : b[i - p.kv_lora_rank];
      if constexpr (!FP8) reinterpret_cast<T*>(dst)[i] = x;
      else dst[i] = quant_e4m3<T>(x, scale);
    }
  }
}

template <typename T>
int launch_concat_mla(const b200_concat_mla_params_t& p, cudaStream_t s) {
  constexpr int N = 16 / sizeof(T);
  const bool fp8 = p.kv_dtype == B200_KV_FP8_E4M3;
  const int cb = fp8 ? 1 : (int)sizeof(T);
  const bool vec = p.kv_lora_rank % N == 0 && p.pe_dim % N == 0 && p.kv_c_stride % N == 0 && p.k_pe_stride % N == 0 &&
                   aligned(p.kv_c, 16) && aligned(p.k_pe, 16) && aligned(p.kv_cache, (size_t)N * cb) &&
                   (p.block_stride * cb) % (N * cb) == 0 && (p.entry_stride * cb) % (N * cb) == 0;
  const int work = vec ? (p.kv_lora_rank + p.pe_dim) / N : p.kv_lora_rank + p.pe_dim;
  const int threads = std::max(32, std::min(256, (work + 31) / 32 * 32));
  if (fp8) {
    if (vec) concat_and_cache_mla_kernel<T, true, true><<<p.num_tokens, threads, 0, s>>>(p);
    else concat_and_cache_mla_kernel<T, true, false><<<p.num_tokens, threads, 0, s>>>(p);
  } else {
    if (vec) concat_and_cache_mla_kernel<T, false, true><<<p.num_tokens, threads, 0, s>>>(p);
    else concat_and_cache_mla_kernel<T, false, false><<<p.num_tokens, threads, 0, s>>>(p);
  }
  count_launch();
  return check_launch("concat_and_cache_mla");
}
}  // namespace
}  // namespace b200

extern "C" int b200_copy_blocks(void* const* key_caches, void* const* value_caches, int32_t num_layers,
                                const int64_t* block_mapping, int32_t num_pairs, int64_t block_bytes,
                                void* stream) {
  return b200::copy_blocks_impl<2>(key_caches, value_caches, num_layers, block_mapping, num_pairs, block_bytes, stream,
                                   "copy_blocks");
}

extern "C" int b200_copy_blocks_mla(void* const* kv_caches, int32_t num_layers, const int64_t* block_mapping,
                                    int32_t num_pairs, int64_t block_bytes, void* stream) {
  return b200::copy_blocks_impl<1>(kv_caches, nullptr, num_layers, block_mapping, num_pairs, block_bytes, stream,
                                   "copy_blocks_mla");
}

extern "C" int b200_concat_and_cache_mla(const b200_concat_mla_params_t* p, void* stream) {
  using namespace b200;
  B200_REQUIRE(p != nullptr, B200_ERR_INVALID_ARGUMENT, "params is NULL");
  B200_REQUIRE(p->num_tokens >= 0 && p->kv_lora_rank > 0 && p->pe_dim >= 0 && p->block_size > 0,
               B200_ERR_INVALID_ARGUMENT, "bad sizes: tokens=%d kv_lora_rank=%d pe_dim=%d block=%d", p->num_tokens,
               p->kv_lora_rank, p->pe_dim, p->block_size);
  if (p->num_tokens == 0) return B200_OK;
  B200_REQUIRE(p->kv_c && p->k_pe && p->kv_cache && p->slot_mapping, B200_ERR_INVALID_ARGUMENT, "NULL tensor pointer");
  B200_REQUIRE(p->kv_dtype == B200_KV_AUTO || p->kv_dtype == B200_KV_FP8_E4M3, B200_ERR_UNSUPPORTED,
               "Unsupported data type of kv cache: %d", p->kv_dtype);
  B200_REQUIRE(p->dtype == B200_DTYPE_F32 || p->dtype == B200_DTYPE_F16 || p->dtype == B200_DTYPE_BF16,
               B200_ERR_UNSUPPORTED, "Unsupported input type of kv cache: %d", p->dtype);
  if (p->kv_dtype != B200_KV_AUTO) B200_REQUIRE(p->scale != nullptr, B200_ERR_INVALID_ARGUMENT, "fp8 cache needs scale");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  switch (p->dtype) {
    case B200_DTYPE_F32: return launch_concat_mla<float>(*p, s);
    case B200_DTYPE_F16: return launch_concat_mla<__half>(*p, s);
    default: return launch_concat_mla<__nv_bfloat16>(*p, s);
  }
}

extern "C" int b200_swap_blocks(void* dst, const void* src, const int64_t* block_mapping, int64_t num_pairs,
                                int64_t block_bytes, int32_t kind, void* stream) {
  using namespace b200;
  B200_REQUIRE(num_pairs >= 0 && block_bytes >= 0, B200_ERR_INVALID_ARGUMENT, "bad sizes");
  if (num_pairs == 0 || block_bytes == 0) return B200_OK;
  B200_REQUIRE(dst && src && block_mapping, B200_ERR_INVALID_ARGUMENT, "NULL pointer");
  cudaMemcpyKind k;
  switch (kind) {
    case B200_COPY_D2D: k = cudaMemcpyDeviceToDevice; break;
    case B200_COPY_D2H: k = cudaMemcpyDeviceToHost; break;
    case B200_COPY_H2D: k = cudaMemcpyHostToDevice; break;
    default: B200_REQUIRE(false, B200_ERR_INVALID_ARGUMENT, "Invalid device combination"); return 0;
  }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  int64_t i = 0;
  while (i < num_pairs) {
    const int64_t sb = block_mapping[2 * i], db = block_mapping[2 * i + 1];
    B200_REQUIRE(sb >= 0 && db >= 0, B200_ERR_INVALID_ARGUMENT, "negative block number in block_mapping");
    int64_t run = 1;  // pairs consecutive on both sides travel as one copy
    while (i + run < num_pairs && block_mapping[2 * (i + run)] == sb + run &&
           block_mapping[2 * (i + run) + 1] == db + run)
      ++run;
    cudaError_t e = cudaMemcpyAsync(static_cast<char*>(dst) + db * block_bytes,
                                    static_cast<const char*>(src) + sb * block_bytes, (size_t)(run * block_bytes), k, s);
    if (e != cudaSuccess) {
      set_error("swap_blocks: cudaMemcpyAsync failed: %s", cudaGetErrorString(e));
      return B200_ERR_CUDA;
    }
    i += run;
  }
  return B200_OK;
}

extern "C" int b200_gather_cache(const b200_gather_cache_params_t* p, void* stream) {
  using namespace b200;
  B200_REQUIRE(p != nullptr, B200_ERR_INVALID_ARGUMENT, "params is NULL");
  B200_REQUIRE(p->batch_size >= 0 && p->block_size > 0 && p->entry_size > 0, B200_ERR_INVALID_ARGUMENT,
               "bad sizes: batch=%d block=%d entry=%d", p->batch_size, p->block_size, p->entry_size);
  if (p->batch_size == 0) return B200_OK;
  B200_REQUIRE(p->src_cache && p->dst && p->block_table && p->cu_seq_lens, B200_ERR_INVALID_ARGUMENT,
               "NULL tensor pointer");
  const bool fp8 = p->kv_dtype == B200_KV_FP8_E4M3;
  B200_REQUIRE(fp8 || p->kv_dtype == B200_KV_AUTO, B200_ERR_UNSUPPORTED, "Unsupported data type of kv cache: %d",
               p->kv_dtype);
  const int sb = dtype_bytes(p->src_dtype), db = dtype_bytes(p->dst_dtype);
  B200_REQUIRE(sb > 0 && db > 0, B200_ERR_UNSUPPORTED, "Unsupported data type: src=%d dst=%d", p->src_dtype,
               p->dst_dtype);
  if (fp8) {
    B200_REQUIRE(p->src_dtype == B200_DTYPE_U8 && p->dst_dtype != B200_DTYPE_U8, B200_ERR_UNSUPPORTED,
                 "fp8 gather needs a uint8 cache and a float/half/bf16 destination");
    B200_REQUIRE(p->scale != nullptr, B200_ERR_INVALID_ARGUMENT, "fp8 gather needs scale");
  } else {
    B200_REQUIRE(sb == db, B200_ERR_UNSUPPORTED, "src_cache and dst must have the same dtype");
  }
  // splits: a warp keeps one row (about 1 KiB) in flight, so the copy needs every warp slot of the machine (8 CTAs of
  // 8 warps per SM) and a few waves: ~32 CTAs per SM (the reference fixes 2/4/16 splits, cache_kernels.cu:747; with 4
  // CTAs per SM this kernel was latency-bound at 4.6-4.8 TB/s)
  static const int cover = env_int_or("B200_GATHER_COVER", 32);
  int splits = (cover * device_sm_count() + p->batch_size - 1) / p->batch_size;
  splits = std::max(1, std::min(splits, 1024));
  dim3 grid(p->batch_size, splits);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (fp8) {
    const bool vec = p->entry_size % 8 == 0 && p->cache_block_stride % 8 == 0 && p->cache_entry_stride % 8 == 0 &&
                     aligned(p->src_cache, 8) && aligned(p->dst, 16) && (p->dst_entry_stride * db) % 16 == 0;
    switch (p->dst_dtype) {
      case B200_DTYPE_F32: launch_gather<float, true>(*p, vec, 1, grid, s); break;
      case B200_DTYPE_F16: launch_gather<__half, true>(*p, vec, 1, grid, s); break;
      default: launch_gather<__nv_bfloat16, true>(*p, vec, 1, grid, s); break;
    }
  } else {
    const bool vec = ((int64_t)p->entry_size * sb) % 16 == 0 && (p->cache_block_stride * sb) % 16 == 0 &&
                     (p->cache_entry_stride * sb) % 16 == 0 && (p->dst_entry_stride * sb) % 16 == 0 &&
                     aligned(p->src_cache, 16) && aligned(p->dst, 16);
    switch (sb) {
      case 4: launch_gather<uint32_t, false>(*p, vec, 4, grid, s); break;
      case 2: launch_gather<uint16_t, false>(*p, vec, 2, grid, s); break;
      default: launch_gather<uint8_t, false>(*p, vec, 1, grid, s); break;
    }
  }
  count_launch();
  return check_launch("gather_cache");
}
