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
#include <string.h>

#include <algorithm>

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
__device__ __forceinline__ void scatter_flash_token(const b200_cache_write_params_t& p, int64_t token) {
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

template <typename T, bool FP8, bool VEC>
__global__ void reshape_and_cache_flash_kernel(b200_cache_write_params_t p) {
  scatter_flash_token<T, FP8, VEC>(p, blockIdx.x);
}

// ---- rotary_embedding (+ fused flash cache write) ---------------------------------------------------------------
// Reference: csrc/pos_encoding_kernels.cu:9-96.  One CTA per token.  The rotation is written as the reference
// writes it in scalar_t arithmetic (:31-34): each product and the sum are rounded to T, which is also what the
// PyTorch-native implementation the reference tests against does — so results are bit-identical to that oracle.
template <typename T>
__device__ __forceinline__ T rnd(float v) { return from_f32<T>(v); }

template <typename T>
__device__ __forceinline__ void rotate_pair(T x, T y, T c, T sn, T& ox, T& oy) {
  const float xf = to_f32<T>(x), yf = to_f32<T>(y), cf = to_f32<T>(c), sf = to_f32<T>(sn);
  const float xc = to_f32<T>(rnd<T>(__fmul_rn(xf, cf))), ys = to_f32<T>(rnd<T>(__fmul_rn(yf, sf)));
  const float yc = to_f32<T>(rnd<T>(__fmul_rn(yf, cf))), xs = to_f32<T>(rnd<T>(__fmul_rn(xf, sf)));
  ox = rnd<T>(__fsub_rn(xc, ys));
  oy = rnd<T>(__fadd_rn(yc, xs));
}

template <typename T, bool NEOX>
__device__ __forceinline__ void rotate_heads(T* base, int64_t head_stride, int nheads, int embed, const T* cos_ptr,
                                             const T* sin_ptr) {
  const int n = nheads * embed;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const int head = i / embed, r = i % embed;
    T* arr = base + (int64_t)head * head_stride;
    const int xi = NEOX ? r : 2 * r;
    const int yi = NEOX ? embed + r : 2 * r + 1;
    rotate_pair<T>(arr[xi], arr[yi], cos_ptr[r], sin_ptr[r], arr[xi], arr[yi]);
  }
}

// 16-byte version: NeoX — one vector of x (dims r..r+N-1), the matching vector of y (embed + r..), one vector each of
// cos and sin; GPT-J — one vector holds N/2 adjacent (x, y) pairs.  Same arithmetic as rotate_heads.
template <typename T, bool NEOX>
__device__ __forceinline__ void rotate_heads_vec(T* base, int64_t head_stride, int nheads, int embed, const T* cos_ptr,
                                                 const T* sin_ptr) {
  constexpr int N = VecOf<T>::kElems;
  if constexpr (NEOX) {
    const int cph = embed / N;  // chunks per head
    for (int i = threadIdx.x; i < nheads * cph; i += blockDim.x) {
      const int head = i / cph, r = (i % cph) * N;
      T* arr = base + (int64_t)head * head_stride;
      Vec16 xv = *reinterpret_cast<const Vec16*>(arr + r);
      Vec16 yv = *reinterpret_cast<const Vec16*>(arr + embed + r);
      const Vec16 cv = *reinterpret_cast<const Vec16*>(cos_ptr + r);
      const Vec16 sv = *reinterpret_cast<const Vec16*>(sin_ptr + r);
      T* x = reinterpret_cast<T*>(&xv);
      T* y = reinterpret_cast<T*>(&yv);
      const T* c = reinterpret_cast<const T*>(&cv);
      const T* sn = reinterpret_cast<const T*>(&sv);
#pragma unroll
      for (int j = 0; j < N; ++j) rotate_pair<T>(x[j], y[j], c[j], sn[j], x[j], y[j]);
      *reinterpret_cast<Vec16*>(arr + r) = xv;
      *reinterpret_cast<Vec16*>(arr + embed + r) = yv;
    }
  } else {
    const int cph = 2 * embed / N;
    for (int i = threadIdx.x; i < nheads * cph; i += blockDim.x) {
      const int head = i / cph, e0 = (i % cph) * N;  // first element of the chunk; its first pair is e0 / 2
      T* arr = base + (int64_t)head * head_stride;
      Vec16 v = *reinterpret_cast<const Vec16*>(arr + e0);
      T* e = reinterpret_cast<T*>(&v);
#pragma unroll
      for (int j = 0; j < N / 2; ++j)
        rotate_pair<T>(e[2 * j], e[2 * j + 1], cos_ptr[e0 / 2 + j], sin_ptr[e0 / 2 + j], e[2 * j], e[2 * j + 1]);
      *reinterpret_cast<Vec16*>(arr + e0) = v;
    }
  }
}

// FUSE: 0 none, 1 auto scalar, 2 auto vec, 3 fp8 scalar, 4 fp8 vec (the flash cache write); RVEC: 16-byte rotation
template <typename T, bool NEOX, int FUSE, bool RVEC>
__global__ void rotary_embedding_kernel(const b200_rotary_params_t p, const b200_cache_write_params_t cw) {
  const int64_t token = blockIdx.x;
  const int64_t pos = p.positions[token];
  const int embed = p.rot_dim / 2;
  const T* cos_ptr = reinterpret_cast<const T*>(p.cos_sin_cache) + pos * p.rot_dim;
  const T* sin_ptr = cos_ptr + embed;
  T* q = reinterpret_cast<T*>(p.query) + token * p.query_stride;
  T* k = p.key != nullptr ? reinterpret_cast<T*>(p.key) + token * p.key_stride : nullptr;
  if constexpr (RVEC) {
    rotate_heads_vec<T, NEOX>(q, p.head_stride, p.num_heads, embed, cos_ptr, sin_ptr);
    if (k != nullptr) rotate_heads_vec<T, NEOX>(k, p.head_stride, p.num_kv_heads, embed, cos_ptr, sin_ptr);
  } else {
    rotate_heads<T, NEOX>(q, p.head_stride, p.num_heads, embed, cos_ptr, sin_ptr);
    if (k != nullptr) rotate_heads<T, NEOX>(k, p.head_stride, p.num_kv_heads, embed, cos_ptr, sin_ptr);
  }
  if constexpr (FUSE != 0) {
    if (token >= p.num_cache_tokens) return;  // CUDA-graph padding: slot_mapping is shorter than key
    __syncthreads();  // the rotated key row of this token is visible to the whole CTA
    scatter_flash_token<T, (FUSE >= 3), (FUSE == 2 || FUSE == 4)>(cw, token);
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

namespace b200 {
namespace {

int validate_rotary(const b200_rotary_params_t* p) {
  B200_REQUIRE(p != nullptr, B200_ERR_INVALID_ARGUMENT, "params is NULL");
  B200_REQUIRE(p->num_tokens >= 0 && p->num_heads > 0 && p->num_kv_heads > 0 && p->head_size > 0,
               B200_ERR_INVALID_ARGUMENT, "bad sizes: tokens=%d heads=%d kv_heads=%d head_size=%d", p->num_tokens,
               p->num_heads, p->num_kv_heads, p->head_size);
  B200_REQUIRE(p->rot_dim > 0 && p->rot_dim % 2 == 0 && p->rot_dim <= p->head_size, B200_ERR_INVALID_ARGUMENT,
               "rot_dim %d must be even and at most head_size %d", p->rot_dim, p->head_size);
  B200_REQUIRE(p->num_heads % p->num_kv_heads == 0, B200_ERR_INVALID_ARGUMENT,
               "num_heads %d is not a multiple of num_kv_heads %d", p->num_heads, p->num_kv_heads);  // :182
  B200_REQUIRE(p->dtype == B200_DTYPE_F32 || p->dtype == B200_DTYPE_F16 || p->dtype == B200_DTYPE_BF16,
               B200_ERR_UNSUPPORTED, "Unsupported data type: %d", p->dtype);
  if (p->num_tokens == 0) return B200_OK;
  B200_REQUIRE(p->positions && p->query && p->cos_sin_cache, B200_ERR_INVALID_ARGUMENT, "NULL tensor pointer");
  return B200_OK;
}

template <typename T>
int launch_rotary(const b200_rotary_params_t& p, bool fuse, cudaStream_t s) {
  b200_cache_write_params_t cw;
  memset(&cw, 0, sizeof(cw));
  int mode = 0;
  if (fuse) {
    cw.key = p.key; cw.value = p.value; cw.key_cache = p.key_cache; cw.value_cache = p.value_cache;
    cw.slot_mapping = p.slot_mapping; cw.k_scale = p.k_scale; cw.v_scale = p.v_scale;
    cw.key_stride = p.key_stride; cw.value_stride = p.value_stride;
    cw.block_stride = p.block_stride; cw.page_stride = p.page_stride; cw.head_stride = p.cache_head_stride;
    cw.num_tokens = p.num_cache_tokens; cw.num_heads = p.num_kv_heads; cw.head_size = p.head_size;
    cw.block_size = p.block_size; cw.dtype = p.dtype; cw.kv_dtype = p.kv_dtype;
    constexpr int N = VecOf<T>::kElems;
    const bool fp8 = p.kv_dtype == B200_KV_FP8_E4M3;
    const size_t dst_align = (size_t)N * (fp8 ? 1 : sizeof(T));
    // the vector path reads the key row as contiguous heads: needs head_stride == head_size
    const bool vec = (p.head_stride == p.head_size) && (p.head_size % N == 0) && (p.key_stride % N == 0) &&
                     (p.value_stride % N == 0) && (p.block_stride % N == 0) && (p.page_stride % N == 0) &&
                     (p.cache_head_stride % N == 0) && aligned(p.key, 16) && aligned(p.value, 16) &&
                     aligned(p.key_cache, dst_align) && aligned(p.value_cache, dst_align);
    mode = fp8 ? (vec ? 4 : 3) : (vec ? 2 : 1);
  }
  // 16-byte rotation: whole vectors inside each half (NeoX) / each head (GPT-J), vector-aligned rows and cos/sin
  constexpr int NV = VecOf<T>::kElems;
  const int embed = p.rot_dim / 2;
  const bool rvec = (p.is_neox ? embed % NV == 0 : p.rot_dim % NV == 0) && p.head_stride % NV == 0 &&
                    p.query_stride % NV == 0 && (p.key == nullptr || p.key_stride % NV == 0) && aligned(p.query, 16) &&
                    (p.key == nullptr || aligned(p.key, 16)) && aligned(p.cos_sin_cache, 16);
  const int work = rvec ? (p.num_heads + p.num_kv_heads) * p.rot_dim / NV : p.num_heads * embed;
  const int threads = std::max(32, std::min(rvec ? 256 : 512, (work + 31) / 32 * 32));  // csrc/pos_encoding_kernels.cu:191
  dim3 grid(p.num_tokens), block(threads);
#define ROT(NEOX, MODE) \
  (rvec ? rotary_embedding_kernel<T, NEOX, MODE, true><<<grid, block, 0, s>>>(p, cw) \
        : rotary_embedding_kernel<T, NEOX, MODE, false><<<grid, block, 0, s>>>(p, cw))
#define ROT_MODE(NEOX)                                                                              \
  switch (mode) {                                                                                   \
    case 0: ROT(NEOX, 0); break;                                                                    \
    case 1: ROT(NEOX, 1); break;                                                                    \
    case 2: ROT(NEOX, 2); break;                                                                    \
    case 3: ROT(NEOX, 3); break;                                                                    \
    default: ROT(NEOX, 4); break;                                                                   \
  }
  if (p.is_neox) { ROT_MODE(true) } else { ROT_MODE(false) }
#undef ROT_MODE
#undef ROT
  count_launch();
  return check_launch(fuse ? "rotary_embedding_and_cache_flash" : "rotary_embedding");
}

int dispatch_rotary(const b200_rotary_params_t* p, bool fuse, void* stream) {
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  switch (p->dtype) {
    case B200_DTYPE_F32: return launch_rotary<float>(*p, fuse, s);
    case B200_DTYPE_F16: return launch_rotary<__half>(*p, fuse, s);
    default: return launch_rotary<__nv_bfloat16>(*p, fuse, s);
  }
}

}  // namespace
}  // namespace b200

extern "C" int b200_rotary_embedding(const b200_rotary_params_t* p, void* stream) {
  using namespace b200;
  int rc = validate_rotary(p);
  if (rc != B200_OK || p->num_tokens == 0) return rc;
  return dispatch_rotary(p, false, stream);
}

extern "C" int b200_rotary_embedding_and_cache_flash(const b200_rotary_params_t* p, void* stream) {
  using namespace b200;
  int rc = validate_rotary(p);
  if (rc != B200_OK || p->num_tokens == 0) return rc;
  B200_REQUIRE(p->key && p->value && p->key_cache && p->value_cache && p->slot_mapping, B200_ERR_INVALID_ARGUMENT,
               "NULL tensor pointer (the fused cache write needs key, value, both caches and slot_mapping)");
  B200_REQUIRE(p->num_cache_tokens >= 0 && p->num_cache_tokens <= p->num_tokens && p->block_size > 0,
               B200_ERR_INVALID_ARGUMENT, "bad sizes: cache tokens=%d tokens=%d block=%d", p->num_cache_tokens,
               p->num_tokens, p->block_size);
  B200_REQUIRE(p->kv_dtype == B200_KV_AUTO || p->kv_dtype == B200_KV_FP8_E4M3, B200_ERR_UNSUPPORTED,
               "Unsupported data type of kv cache: %d", p->kv_dtype);
  if (p->kv_dtype != B200_KV_AUTO)
    B200_REQUIRE(p->k_scale && p->v_scale, B200_ERR_INVALID_ARGUMENT, "fp8 cache needs k_scale/v_scale");
  // the cache writer reads a token's heads as one contiguous run (as reshape_and_cache_flash does)
  B200_REQUIRE(p->head_stride == p->head_size, B200_ERR_UNSUPPORTED,
               "the fused cache write needs contiguous heads (head_stride %lld != head_size %d)",
               (long long)p->head_stride, p->head_size);
  return dispatch_rotary(p, true, stream);
}
