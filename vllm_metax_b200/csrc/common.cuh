// Shared device/host helpers for the B200 paged-attention path.
// Numerics contract: everything is computed in fp32; bf16/fp16
// <-> fp32 conversions are exact on load and round-to-nearest-even on store;
// FP8-e4m3 follows the semantics the reference's call sites intend
// (csrc/cache_kernels.cu:249-252, csrc/attention/attention_kernels.cuh:276-277):
//   quant   = e4m3_satfinite_rn(float(x) / scale)   (IEEE divide)
//   dequant = float(e4m3) * scale
#pragma once

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_fp8.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/b200_paged_attn.h"

namespace b200 {

constexpr int kWarpSize = 32;
constexpr int kPartitionSize = B200_PARTITION_SIZE;

// ---- error plumbing (host) -------------------------------------------------
void set_error(const char* fmt, ...);
void count_launch(int n = 1);
int check_launch(const char* what);  // cudaGetLastError -> status

#define B200_REQUIRE(cond, code, ...) \
  do {                                \
    if (!(cond)) {                    \
      ::b200::set_error(__VA_ARGS__); \
      return (code);                  \
    }                                 \
  } while (0)

// ---- scalar conversions ----------------------------------------------------
template <typename T>
__device__ __forceinline__ float to_f32(T v);
template <>
__device__ __forceinline__ float to_f32<float>(float v) { return v; }
template <>
__device__ __forceinline__ float to_f32<__half>(__half v) { return __half2float(v); }
template <>
__device__ __forceinline__ float to_f32<__nv_bfloat16>(__nv_bfloat16 v) {
  return __bfloat162float(v);
}

template <typename T>
__device__ __forceinline__ T from_f32(float v);
template <>
__device__ __forceinline__ float from_f32<float>(float v) { return v; }
template <>
__device__ __forceinline__ __half from_f32<__half>(float v) { return __float2half_rn(v); }
template <>
__device__ __forceinline__ __nv_bfloat16 from_f32<__nv_bfloat16>(float v) {
  return __float2bfloat16_rn(v);
}

// e4m3 byte -> fp32 (exact: every e4m3 value is an fp16 value).
__device__ __forceinline__ float e4m3_to_f32(uint8_t b) {
  __half_raw h = __nv_cvt_fp8_to_halfraw((__nv_fp8_storage_t)b, __NV_E4M3);
  return __half2float(*reinterpret_cast<__half*>(&h));
}

// fp32 -> e4m3 byte, round-to-nearest-even, saturate to +-448.
__device__ __forceinline__ uint8_t f32_to_e4m3(float v) {
  return (uint8_t)__nv_cvt_float_to_fp8(v, __NV_SATFINITE, __NV_E4M3);
}

// quantise one element: IEEE divide (never x * (1/scale)), then cast.
template <typename T>
__device__ __forceinline__ uint8_t quant_e4m3(T x, float scale) {
  return f32_to_e4m3(__fdiv_rn(to_f32<T>(x), scale));
}

// dequantise one element to T (round once, to T).
template <typename T>
__device__ __forceinline__ T dequant_e4m3(uint8_t b, float scale) {
  return from_f32<T>(__fmul_rn(e4m3_to_f32(b), scale));
}

// ---- 16-byte vector access ---------------------------------------------------
struct alignas(16) Vec16 {
  uint32_t w[4];
};

__device__ __forceinline__ Vec16 ld_global_16(const void* p) {
  Vec16 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.w[0]), "=r"(v.w[1]), "=r"(v.w[2]), "=r"(v.w[3])
               : "l"(p));
  return v;
}

// the same without the non-coherent path: for kernels that may read and write the same allocation (copy_blocks with
// chained mappings, merge_attn_states whose output aliases an input), where ld.global.nc would be undefined
__device__ __forceinline__ Vec16 ld_global_16_coherent(const void* p) {
  Vec16 v;
  asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.w[0]), "=r"(v.w[1]), "=r"(v.w[2]), "=r"(v.w[3])
               : "l"(p)
               : "memory");
  return v;
}

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int m = 16; m >= 1; m >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, m));
  return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int m = 16; m >= 1; m >>= 1) v += __shfl_xor_sync(0xffffffffu, v, m);
  return v;
}

inline int ceil_div(int a, int b) { return (a + b - 1) / b; }

// element size in bytes of a b200_dtype_t
inline int dtype_bytes(int dt) {
  switch (dt) {
    case B200_DTYPE_F32: return 4;
    case B200_DTYPE_F16: return 2;
    case B200_DTYPE_BF16: return 2;
    case B200_DTYPE_U8: return 1;
    default: return 0;
  }
}

// ---- split-KV partial results ---------------------------------------------
// Both attention kernel families cut a sequence into partitions of `chunk`
// tokens; partition j of (seq s, head h) leaves
//   m[s,h,j]  = max logit of the partition (natural-log units)
//   l[s,h,j]  = sum exp(logit - m)
//   o[s,h,j,:]= sum exp(logit - m) * v            (normalized == 0, fp32)
//             = that / (l + 1e-6), in the query dtype (normalized == 1)
// normalized == 1 with chunk == 512 is exactly the content the reference v2
// kernel leaves in max_logits / exp_sums / tmp_out
// (csrc/attention/attention_kernels.cuh:337-346, :472-484).
struct SplitOut {
  float* m;
  float* l;
  void* o;
  int num_partitions;  // P: partitions per (seq, head) in the layout
  int chunk;           // tokens per partition
  int normalized;      // 1: reference v2 format (o in query dtype)
};

// ---- flat (tile-balanced) schedule of the streaming kernel, used by paged_attention_v1 -----------------
// The tiles (32 tokens of one (seq, kv head) pair) of the whole batch form one flat sequence that is cut
// into equal contiguous ranges, one per CTA.  A pair that straddles a cut leaves one partial result per
// CTA it touches in that CTA's "head" slot (segment that does not start the pair) or "tail" slot (segment
// that starts the pair but does not finish it); the CTA that delivers the pair's last segment — found with
// an arrival counter per pair — merges them.  Slots and counters live in the caller's workspace:
//   [counters: kFlatCounterSlots ints, all zero between launches][slots: kFlatMaxRanges x 2 x kFlatMaxGroup x (D+2) f32]
// (a "range" is a CTA's static share or one dynamically claimed chunk of the tail region; the streaming kernel packs
// its slots by the launch's group size G: slot (range, head/tail, h) = (range * 2 + tail) * G + h).
constexpr int kFlatMaxCtas = 256;
constexpr int kFlatMaxGroup = 16;   // query heads per kv head served by the streaming kernel
constexpr int kFlatHeadSize = 128;
// Ranges of the flat tile sequence: one static range per CTA plus the dynamically claimed chunks of the tail
// region (stream_common.cuh: "guided" chunk sizes, at most kDynMaxLevels levels of one chunk per CTA each).
constexpr int kDynMaxLevels = 5;
constexpr int kFlatMaxRanges = kFlatMaxCtas * (1 + kDynMaxLevels);
constexpr uint64_t kFlatCounterSlots = 1u << 18;  // the last int is the claim counter of the dynamic region
constexpr uint64_t kFlatCounterBytes = kFlatCounterSlots * sizeof(int);
constexpr uint64_t kFlatSlotBytes =
    (uint64_t)kFlatMaxRanges * 2 * kFlatMaxGroup * (kFlatHeadSize + 2) * sizeof(float);
constexpr uint64_t kFlatWorkspaceBytes = kFlatCounterBytes + kFlatSlotBytes;

// ---- kernel-family entry points (defined in the .cu files) ------------------
int launch_attention_generic(const b200_paged_attn_params_t& p, const SplitOut& so, cudaStream_t s);
// v2 format / fixed-size partitions: work item = (partition, seq, kv head)
int launch_attention_stream(const b200_paged_attn_params_t& p, const SplitOut& so, cudaStream_t s);
// v1: flat schedule; `workspace` holds kFlatWorkspaceBytes laid out as above
int launch_attention_stream_flat(const b200_paged_attn_params_t& p, void* workspace, cudaStream_t s);
bool stream_variant_eligible(const b200_paged_attn_params_t& p, const char** why);
// flash layout [num_blocks, block_size, num_kv_heads, D] (NHD or HND strides), bf16: flat schedule, one launch
int launch_attention_stream_flash(const b200_paged_attn_params_t& p, void* workspace, cudaStream_t s);
bool stream_flash_eligible(const b200_paged_attn_params_t& p, const char** why);
// tcgen05 kernel for fp8 caches (attention_umma_fp8.cu): same two schedules
int launch_attention_umma_fp8(const b200_paged_attn_params_t& p, const SplitOut& so, cudaStream_t s);
int launch_attention_umma_fp8_flat(const b200_paged_attn_params_t& p, void* workspace, cudaStream_t s);
bool umma_fp8_variant_eligible(const b200_paged_attn_params_t& p, const char** why);
void set_stream_dynamic(int min_chunk_tiles);  // experiment switch of the flat schedule's dynamic tail region
int launch_split_reduce(const b200_paged_attn_params_t& p, const SplitOut& so, cudaStream_t s);
int device_sm_count();

}  // namespace b200
