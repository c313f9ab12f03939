// KV-cache writers of the decode path: reshape_and_cache (v1/v2 layout),
// reshape_and_cache_flash (NHD / HND flash layout) and convert_fp8, with the
// FP8-e4m3 quantisation fused into the scatter.
//
// Behaviour follows the reference ops (csrc/cache_kernels.cu:203-255, :271-344,
// :543-614): slot < 0 is padding and is skipped; num_tokens comes from
// slot_mapping; reshape_and_cache computes its destination from sizes
// (contiguous caches), reshape_and_cache_flash from strides, NHD vs HND by
// head_stride == head_size.  The kernels are written for B200: one CTA per
// token, 16-byte loads of the source row, 16-byte (auto) / 8-byte (fp8)
// stores wherever the destination run is contiguous, scalar path otherwise.
#include "common.cuh"

namespace b200 {
namespace {

template <typename T>
struct VecOf {
  static constexpr int kElems = 16 / sizeof(T);  // elements per 16-byte vector
};

// Store kElems cache elements (auto: same bits, fp8: quantised bytes).
template <typename T, bool FP8>
__device__ __forceinline__ void store_vec(void* dst, const Vec16& v, float scale) {
  constexpr int N = VecOf<T>::kElems;
  if constexpr (!FP8) {
    *reinterpret_cast<Vec16*>(dst) = v;
  } else {
    const T* e = reinterpret_cast<const T*>(&v);
    alignas(8) uint8_t q[N];
#pragma unroll
    for (int i = 0; i < N; ++i) q[i] = quant_e4m3<T>(e[i], scale);
    if constexpr (N == 8) {
      *reinterpret_cast<uint2*>(dst) = *reinterpret_cast<const uint2*>(q);
    } else {
      *reinterpret_cast<uint32_t*>(dst) = *reinterpret_cast<const uint32_t*>(q);
    }
  }
}

template <typename T, bool FP8>
__device__ __forceinline__ void store_one(void* dst_base, int64_t idx, T v, float scale) {
  if constexpr (!FP8) {
    reinterpret_cast<T*>(dst_base)[idx] = v;
  } else {
    reinterpret_cast<uint8_t*>(dst_base)[idx] = quant_e4m3<T>(v, scale);
  }
}

// ---- reshape_and_cache -------------------------------------------------------
// key_cache [NB, H, D/x, B, x], value_cache [NB, H, D, B], both contiguous.
// VECK: the x elements of one K unit are moved as 16-byte vector loads
// (requires x % kElems == 0 and 16-byte aligned source rows).
template <typename T, bool FP8, bool VECK>
__global__ void reshape_and_cache_kernel(b200_cache_write_params_t p) {
  // one CTA per (token, head): a decode step has few tokens, so the heads supply the parallelism
  const int64_t token = blockIdx.x;
  const int h = blockIdx.y;
  const int64_t slot = p.slot_mapping[token];
  if (slot < 0) return;  // padding token (csrc/cache_kernels.cu:217-220)
  const int64_t blk = slot / p.block_size;
  const int64_t off = slot % p.block_size;
  const int H = p.num_heads, D = p.head_size, B = p.block_size, x = p.x;
  const float ks = FP8 ? *p.k_scale : 0.f;
  const float vs = FP8 ? *p.v_scale : 0.f;
  const T* ksrc = reinterpret_cast<const T*>(p.key) + token * p.key_stride + (int64_t)h * D;
  const T* vsrc = reinterpret_cast<const T*>(p.value) + token * p.value_stride + (int64_t)h * D;
  constexpr int N = VecOf<T>::kElems;
  constexpr int CB = FP8 ? 1 : (int)sizeof(T);  // cache element bytes
  const int64_t kbase = (blk * H + h) * (int64_t)(D / x) * B * x + off * x;  // + (d/x)*B*x + d%x
  const int64_t vbase = (blk * H + h) * (int64_t)D * B + off;               // + d*B

  if constexpr (VECK) {
    // one 16-byte source vector = N consecutive dims inside one x-unit
    for (int v = threadIdx.x; v < D / N; v += blockDim.x) {
      const int d = v * N;
      const int64_t dst = kbase + (int64_t)(d / x) * B * x + d % x;
      Vec16 val = *reinterpret_cast<const Vec16*>(ksrc + d);
      store_vec<T, FP8>(reinterpret_cast<uint8_t*>(p.key_cache) + dst * CB, val, ks);
    }
  } else {
    for (int d = threadIdx.x; d < D; d += blockDim.x)
      store_one<T, FP8>(p.key_cache, kbase + (int64_t)(d / x) * B * x + d % x, ksrc[d], ks);
  }
  // V: destination elements of one token are B apart -> scalar scatter.
  for (int d = threadIdx.x; d < D; d += blockDim.x)
    store_one<T, FP8>(p.value_cache, vbase + (int64_t)d * B, vsrc[d], vs);
}

// ---- reshape_and_cache_flash -------------------------------------------------
// token row in the cache: base + blk*block_stride + off*page_stride
// (+ h*head_stride); every head's D elements are contiguous in both layouts,
// so one vector never straddles a head when D % kElems == 0.
template <typename T, bool FP8, bool VEC>
__global__ void reshape_and_cache_flash_kernel(b200_cache_write_params_t p) {
  const int64_t token = blockIdx.x;
  const int64_t slot = p.slot_mapping[token];
  if (slot < 0) return;  // padding token (csrc/cache_kernels.cu:285-289)
  const int64_t blk = slot / p.block_size;
  const int64_t off = slot % p.block_size;
  const int H = p.num_heads, D = p.head_size;
  const float ks = FP8 ? *p.k_scale : 0.f;
  const float vs = FP8 ? *p.v_scale : 0.f;
  const T* ksrc = reinterpret_cast<const T*>(p.key) + token * p.key_stride;
  const T* vsrc = reinterpret_cast<const T*>(p.value) + token * p.value_stride;
  const int64_t row = blk * p.block_stride + off * p.page_stride;
  const int n = H * D;
  constexpr int N = VecOf<T>::kElems;
  constexpr int CB = FP8 ? 1 : (int)sizeof(T);
  uint8_t* kdst = reinterpret_cast<uint8_t*>(p.key_cache);
  uint8_t* vdst = reinterpret_cast<uint8_t*>(p.value_cache);

  if constexpr (VEC) {
    for (int v = threadIdx.x; v < n / N; v += blockDim.x) {
      const int i = v * N;
      const int h = i / D, d = i % D;
      const int64_t dst = row + (int64_t)h * p.head_stride + d;
      Vec16 kv = *reinterpret_cast<const Vec16*>(ksrc + i);
      Vec16 vv = *reinterpret_cast<const Vec16*>(vsrc + i);
      store_vec<T, FP8>(kdst + dst * CB, kv, ks);
      store_vec<T, FP8>(vdst + dst * CB, vv, vs);
    }
  } else {
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      const int h = i / D, d = i % D;
      const int64_t dst = row + (int64_t)h * p.head_stride + d;
      store_one<T, FP8>(p.key_cache, dst, ksrc[i], ks);
      store_one<T, FP8>(p.value_cache, dst, vsrc[i], vs);
    }
  }
}

// ---- convert_fp8 ---------------------------------------------------------------
template <typename T>
__global__ void fp8_quant_kernel(uint8_t* __restrict__ dst, const T* __restrict__ src,
                                 int64_t n, float scale) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x)
    dst[i] = quant_e4m3<T>(src[i], scale);
}
template <typename T>
__global__ void fp8_dequant_kernel(T* __restrict__ dst, const uint8_t* __restrict__ src,
                                   int64_t n, float scale) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x)
    dst[i] = dequant_e4m3<T>(src[i], scale);
}

inline bool aligned(const void* p, size_t a) { return (reinterpret_cast<uintptr_t>(p) % a) == 0; }

template <typename T>
int launch_reshape_and_cache(const b200_cache_write_params_t& p, cudaStream_t s) {
  constexpr int N = VecOf<T>::kElems;
  const bool fp8 = p.kv_dtype == B200_KV_FP8_E4M3;
  const int n = p.num_heads * p.head_size;
  // vector path: x-units are whole vectors, source rows and the destination
  // units are vector aligned.
  const int cache_bytes = fp8 ? 1 : (int)sizeof(T);
  const bool veck = (p.x % N == 0) && (p.head_size % p.x == 0) && (p.key_stride % N == 0) && (p.head_size % N == 0) &&
                    aligned(p.key, 16) && aligned(p.key_cache, (size_t)N * cache_bytes);
  const int threads = max(32, min(256, (p.head_size + 31) / 32 * 32));
  dim3 grid(p.num_tokens, p.num_heads), block(threads);
  if (p.num_tokens == 0) return B200_OK;
#define LAUNCH(FP8, VECK) reshape_and_cache_kernel<T, FP8, VECK><<<grid, block, 0, s>>>(p)
  if (fp8) {
    if (veck) LAUNCH(true, true); else LAUNCH(true, false);
  } else {
    if (veck) LAUNCH(false, true); else LAUNCH(false, false);
  }
#undef LAUNCH
  count_launch();
  return check_launch("reshape_and_cache");
}

template <typename T>
int launch_reshape_and_cache_flash(const b200_cache_write_params_t& p, cudaStream_t s) {
  constexpr int N = VecOf<T>::kElems;
  const bool fp8 = p.kv_dtype == B200_KV_FP8_E4M3;
  const int n = p.num_heads * p.head_size;
  const int cache_bytes = fp8 ? 1 : (int)sizeof(T);
  const size_t dst_align = (size_t)N * cache_bytes;
  const bool vec = (p.head_size % N == 0) && (p.key_stride % N == 0) && (p.value_stride % N == 0) &&
                   (p.block_stride % N == 0) && (p.page_stride % N == 0) &&
                   (p.head_stride % N == 0) && aligned(p.key, 16) && aligned(p.value, 16) &&
                   aligned(p.key_cache, dst_align) && aligned(p.value_cache, dst_align);
  const int work = vec ? n / N : n;
  const int threads = max(32, min(512, (work + 31) / 32 * 32));
  dim3 grid(p.num_tokens), block(threads);
  if (p.num_tokens == 0) return B200_OK;
#define LAUNCH(FP8, VEC) reshape_and_cache_flash_kernel<T, FP8, VEC><<<grid, block, 0, s>>>(p)
  if (fp8) {
    if (vec) LAUNCH(true, true); else LAUNCH(true, false);
  } else {
    if (vec) LAUNCH(false, true); else LAUNCH(false, false);
  }
#undef LAUNCH
  count_launch();
  return check_launch("reshape_and_cache_flash");
}

int validate_cache_params(const b200_cache_write_params_t* p, bool flash) {
  B200_REQUIRE(p != nullptr, B200_ERR_INVALID_ARGUMENT, "params is NULL");
  B200_REQUIRE(p->num_tokens >= 0 && p->num_heads > 0 && p->head_size > 0 && p->block_size > 0,
               B200_ERR_INVALID_ARGUMENT, "bad sizes: tokens=%d heads=%d head_size=%d block=%d",
               p->num_tokens, p->num_heads, p->head_size, p->block_size);
  if (p->num_tokens == 0) return B200_OK;
  B200_REQUIRE(p->key && p->value && p->key_cache && p->value_cache && p->slot_mapping,
               B200_ERR_INVALID_ARGUMENT, "NULL tensor pointer");
  B200_REQUIRE(p->kv_dtype == B200_KV_AUTO || p->kv_dtype == B200_KV_FP8_E4M3,
               B200_ERR_UNSUPPORTED, "Unsupported data type of kv cache: %d", p->kv_dtype);
  B200_REQUIRE(p->dtype == B200_DTYPE_F32 || p->dtype == B200_DTYPE_F16 || p->dtype == B200_DTYPE_BF16,
               B200_ERR_UNSUPPORTED, "Unsupported input type of kv cache: %d", p->dtype);
  if (p->kv_dtype != B200_KV_AUTO)
    B200_REQUIRE(p->k_scale && p->v_scale, B200_ERR_INVALID_ARGUMENT, "fp8 cache needs k_scale/v_scale");
  if (!flash)
    B200_REQUIRE(p->x > 0 && p->head_size % p->x == 0, B200_ERR_INVALID_ARGUMENT,
                 "head_size %d not a multiple of x %d", p->head_size, p->x);
  return B200_OK;
}

}  // namespace
}  // namespace b200

extern "C" int b200_reshape_and_cache(const b200_cache_write_params_t* p, void* stream) {
  using namespace b200;
  int rc = validate_cache_params(p, false);
  if (rc != B200_OK) return rc;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  switch (p->dtype) {
    case B200_DTYPE_F32: return launch_reshape_and_cache<float>(*p, s);
    case B200_DTYPE_F16: return launch_reshape_and_cache<__half>(*p, s);
    default: return launch_reshape_and_cache<__nv_bfloat16>(*p, s);
  }
}

extern "C" int b200_reshape_and_cache_flash(const b200_cache_write_params_t* p, void* stream) {
  using namespace b200;
  int rc = validate_cache_params(p, true);
  if (rc != B200_OK) return rc;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  switch (p->dtype) {
    case B200_DTYPE_F32: return launch_reshape_and_cache_flash<float>(*p, s);
    case B200_DTYPE_F16: return launch_reshape_and_cache_flash<__half>(*p, s);
    default: return launch_reshape_and_cache_flash<__nv_bfloat16>(*p, s);
  }
}

extern "C" int b200_convert_fp8(void* dst, const void* src, int64_t n, float scale,
                                int32_t dst_dtype, int32_t src_dtype, int32_t kv_dtype,
                                void* stream) {
  using namespace b200;
  B200_REQUIRE(n >= 0, B200_ERR_INVALID_ARGUMENT, "negative element count");
  if (n == 0) return B200_OK;
  B200_REQUIRE(dst && src, B200_ERR_INVALID_ARGUMENT, "NULL tensor pointer");
  // "auto" is accepted by the reference's switch (csrc/cache_kernels.cu:581) but
  // routes to an assert(false) conversion; only e4m3 has defined semantics.
  B200_REQUIRE(kv_dtype == B200_KV_FP8_E4M3, B200_ERR_UNSUPPORTED,
               "Unsupported data type: convert_fp8 needs kv_cache_dtype fp8/fp8_e4m3");
  const bool quant = dst_dtype == B200_DTYPE_U8;
  const bool dequant = src_dtype == B200_DTYPE_U8;
  B200_REQUIRE(quant != dequant, B200_ERR_INVALID_ARGUMENT,
               "exactly one of dst/src must be uint8 (dst=%d src=%d)", dst_dtype, src_dtype);
  const int other = quant ? src_dtype : dst_dtype;
  B200_REQUIRE(other == B200_DTYPE_F32 || other == B200_DTYPE_F16 || other == B200_DTYPE_BF16,
               B200_ERR_UNSUPPORTED, "Unsupported data type: %d", other);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int threads = 256;
  const int64_t want = (n + threads - 1) / threads;
  const int blocks = (int)(want < 148 * 16 ? want : 148 * 16);
#define Q(T) fp8_quant_kernel<T><<<blocks, threads, 0, s>>>((uint8_t*)dst, (const T*)src, n, scale)
#define DQ(T) fp8_dequant_kernel<T><<<blocks, threads, 0, s>>>((T*)dst, (const uint8_t*)src, n, scale)
  if (quant) {
    if (other == B200_DTYPE_F32) Q(float);
    else if (other == B200_DTYPE_F16) Q(__half);
    else Q(__nv_bfloat16);
  } else {
    if (other == B200_DTYPE_F32) DQ(float);
    else if (other == B200_DTYPE_F16) DQ(__half);
    else DQ(__nv_bfloat16);
  }
#undef Q
#undef DQ
  count_launch();
  return check_launch("convert_fp8");
}
