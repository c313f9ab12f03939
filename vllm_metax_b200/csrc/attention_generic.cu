// Generic paged-attention decode kernel (FMA) and the split-KV reduce kernel.
//
// This is the "any supported shape" family: every head size / block size /
// dtype the reference op accepts (csrc/attention/paged_attention_v1.cu:90-158,
// csrc/quantization/fp8/metax/quant_utils.cuh:29-39), ALiBi, fp8-e4m3 caches.
// The D=128 bf16/fp8 decode shapes are served by the streaming kernel in
// attention_stream.cu instead; both leave their split-KV partials in the same
// SplitOut format and share the reduce kernel below.
//
// Work decomposition: one CTA per (kv head, seq, partition of `chunk` tokens);
// the CTA serves all G query heads of its kv head (8 at a time), so K/V are
// read once per GQA group, not once per query head as in the reference
// (grid.x = num_heads, csrc/attention/paged_attention_v1.cu:86).
//
// Numerics: fp32 logits / softmax / accumulation, 1/(sum + 1e-6) normaliser
// (csrc/attention/attention_kernels.cuh:331), tokens past seq_len never read.
#include <float.h>

#include <type_traits>

#include "common.cuh"

namespace b200 {
namespace {

constexpr int kThreads = 128;
constexpr int kWarps = kThreads / kWarpSize;
constexpr int kHeadsPerPass = 8;
constexpr int kMaxChunk = 512;
constexpr int kMaxD = 256;

struct GenericParams {
  const void* q;
  const void* kc;
  const void* vc;
  const int32_t* bt;
  const int32_t* seq_lens;
  const float* alibi;
  const float* k_scale;
  const float* v_scale;
  void* out;
  float* lse;  // optional [H, S] log-sum-exp of the scaled logits (natural log); written with the final output
  SplitOut so;
  int64_t q_stride, kv_block_stride, kv_head_stride;
  int S, H, Hkv, G, D, B, max_blocks;
  float scale;
};

template <typename CT>
struct CacheTraits;
template <>
struct CacheTraits<float> {
  static constexpr int X = 4;
  static __device__ __forceinline__ float get(const Vec16& v, int i) {
    return reinterpret_cast<const float*>(&v)[i];
  }
};
template <>
struct CacheTraits<__half> {
  static constexpr int X = 8;
  static __device__ __forceinline__ float get(const Vec16& v, int i) {
    return __half2float(reinterpret_cast<const __half*>(&v)[i]);
  }
};
template <>
struct CacheTraits<__nv_bfloat16> {
  static constexpr int X = 8;
  static __device__ __forceinline__ float get(const Vec16& v, int i) {
    return __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(&v)[i]);
  }
};
template <>
struct CacheTraits<uint8_t> {  // fp8-e4m3 bytes
  static constexpr int X = 16;
  static __device__ __forceinline__ float get(const Vec16& v, int i) {
    return e4m3_to_f32(reinterpret_cast<const uint8_t*>(&v)[i]);
  }
};

// T: query/out dtype, CT: cache element (T, or uint8_t for fp8)
template <typename T, typename CT>
__global__ void __launch_bounds__(kThreads)
paged_attention_generic_kernel(GenericParams p) {
  constexpr int X = CacheTraits<CT>::X;
  constexpr bool FP8 = sizeof(CT) == 1;
  const int kvh = blockIdx.x, s = blockIdx.y, part = blockIdx.z;
  const int L = p.seq_lens[s];
  const int chunk = p.so.chunk;
  const int t0 = part * chunk;
  if (t0 >= L) {
    // an empty sequence (CUDA-graph padding): the reference kernels leave zeros (acc = 0 times 1/(0 + 1e-6),
    // csrc/attention/attention_kernels.cuh:331); lse = -inf so that a merge ignores the row
    if (L <= 0 && part == 0) {
      T* o = reinterpret_cast<T*>(p.out) + ((int64_t)s * p.H + (int64_t)kvh * p.G) * p.D;
      for (int i = threadIdx.x; i < p.G * p.D; i += kThreads) o[i] = from_f32<T>(0.f);
      if (p.lse != nullptr)
        for (int h = threadIdx.x; h < p.G; h += kThreads) p.lse[(int64_t)(kvh * p.G + h) * p.S + s] = -INFINITY;
    }
    return;
  }
  const int t1 = min(t0 + chunk, L);
  const int n = t1 - t0;
  const int D = p.D, B = p.B, G = p.G;
  const int tid = threadIdx.x, lane = tid % kWarpSize, warp = tid / kWarpSize;
  const bool single = L <= chunk;  // the whole sequence is this partition

  __shared__ float q_s[kHeadsPerPass][kMaxD];
  __shared__ float logit_s[kHeadsPerPass][kMaxChunk];
  __shared__ float m_s[kHeadsPerPass], l_s[kHeadsPerPass];

  const float ks = FP8 ? *p.k_scale : 1.f;
  const float vs = FP8 ? *p.v_scale : 1.f;
  const float qk_scale = p.scale * ks;
  const int32_t* bt = p.bt + (int64_t)s * p.max_blocks;
  const CT* kc = reinterpret_cast<const CT*>(p.kc) + (int64_t)kvh * p.kv_head_stride;
  const CT* vc = reinterpret_cast<const CT*>(p.vc) + (int64_t)kvh * p.kv_head_stride;

  for (int h0 = 0; h0 < G; h0 += kHeadsPerPass) {
    const int nh = min(kHeadsPerPass, G - h0);
    const int head0 = kvh * G + h0;
    __syncthreads();  // previous pass done with q_s / logit_s
    for (int i = tid; i < nh * D; i += kThreads) {
      const int hh = i / D, d = i % D;
      q_s[hh][d] = to_f32<T>(
          reinterpret_cast<const T*>(p.q)[(int64_t)s * p.q_stride + (int64_t)(head0 + hh) * D + d]);
    }
    __syncthreads();

    // ---- logits: one token per thread ----------------------------------------
    for (int ti = tid; ti < n; ti += kThreads) {
      const int t = t0 + ti;
      const int64_t blk = bt[t / B];
      const int off = t % B;
      const CT* kp = kc + blk * p.kv_block_stride + (int64_t)off * X;
      float acc[kHeadsPerPass];
#pragma unroll
      for (int hh = 0; hh < kHeadsPerPass; ++hh) acc[hh] = 0.f;
      for (int j = 0; j < D / X; ++j) {
        const Vec16 kv = *reinterpret_cast<const Vec16*>(kp + (int64_t)j * B * X);
#pragma unroll
        for (int e = 0; e < X; ++e) {
          const float kf = CacheTraits<CT>::get(kv, e);
#pragma unroll
          for (int hh = 0; hh < kHeadsPerPass; ++hh) acc[hh] = fmaf(q_s[hh][j * X + e], kf, acc[hh]);
        }
      }
#pragma unroll
      for (int hh = 0; hh < kHeadsPerPass; ++hh) {
        if (hh < nh) {
          float lg = acc[hh] * qk_scale;
          if (p.alibi) {
            const float slope = p.alibi[head0 + hh];
            lg += (slope != 0.f) ? slope * (float)(t - L + 1) : 0.f;
          }
          logit_s[hh][ti] = lg;
        }
      }
    }
    __syncthreads();

    // ---- softmax statistics: one warp per head --------------------------------
    for (int hh = warp; hh < nh; hh += kWarps) {
      float m = -FLT_MAX;
      for (int ti = lane; ti < n; ti += kWarpSize) m = fmaxf(m, logit_s[hh][ti]);
      m = warp_max(m);
      float l = 0.f;
      for (int ti = lane; ti < n; ti += kWarpSize) {
        const float e = __expf(logit_s[hh][ti] - m);
        logit_s[hh][ti] = e;
        l += e;
      }
      l = warp_sum(l);
      if (lane == 0) {
        m_s[hh] = m;
        l_s[hh] = l;
      }
    }
    __syncthreads();

    // ---- P.V: one (or two) head dims per thread --------------------------------
    float o[kHeadsPerPass][kMaxD / kThreads];
#pragma unroll
    for (int hh = 0; hh < kHeadsPerPass; ++hh)
#pragma unroll
      for (int r = 0; r < kMaxD / kThreads; ++r) o[hh][r] = 0.f;
    constexpr int VPB = 16 / sizeof(CT);  // cache elements per 16-byte load
#pragma unroll
    for (int r = 0; r < kMaxD / kThreads; ++r) {
      const int d = tid + r * kThreads;
      if (d >= D) break;
      for (int tb = t0; tb < t1; tb += B) {  // t0 is a multiple of B
        const int64_t blk = bt[tb / B];
        const CT* vp = vc + blk * p.kv_block_stride + (int64_t)d * B;
        const int nt = min(B, t1 - tb);
        if (B < VPB) {  // fp8 with block_size 8: rows are shorter than a vector
          for (int e = 0; e < nt; ++e) {
            Vec16 one;
            reinterpret_cast<CT*>(&one)[0] = vp[e];
            const float vf = CacheTraits<CT>::get(one, 0);
#pragma unroll
            for (int hh = 0; hh < kHeadsPerPass; ++hh)
              o[hh][r] = fmaf(logit_s[hh][tb - t0 + e], vf, o[hh][r]);
          }
          continue;
        }
        for (int v0 = 0; v0 < nt; v0 += VPB) {
          const Vec16 vv = *reinterpret_cast<const Vec16*>(vp + v0);
#pragma unroll
          for (int e = 0; e < VPB; ++e) {
            if (v0 + e < nt) {  // never touch tokens past seq_len (stale cache)
              const float vf = CacheTraits<CT>::get(vv, e);
              const int ti = tb - t0 + v0 + e;
#pragma unroll
              for (int hh = 0; hh < kHeadsPerPass; ++hh) o[hh][r] = fmaf(logit_s[hh][ti], vf, o[hh][r]);
            }
          }
        }
      }
    }

    // ---- write ------------------------------------------------------------------
#pragma unroll
    for (int r = 0; r < kMaxD / kThreads; ++r) {
      const int d = tid + r * kThreads;
      if (d >= D) break;
#pragma unroll
      for (int hh = 0; hh < kHeadsPerPass; ++hh) {
        if (hh >= nh) break;
        const int head = head0 + hh;
        const float l = l_s[hh];
        const float ov = o[hh][r] * vs;
        const float on = ov * __fdividef(1.f, l + 1e-6f);
        if (single)
          reinterpret_cast<T*>(p.out)[((int64_t)s * p.H + head) * D + d] = from_f32<T>(on);
        if (p.so.o != nullptr && (!single || p.so.normalized)) {
          const int64_t pi = ((int64_t)s * p.H + head) * p.so.num_partitions + part;
          if (p.so.normalized)
            reinterpret_cast<T*>(p.so.o)[pi * D + d] = from_f32<T>(on);
          else
            reinterpret_cast<float*>(p.so.o)[pi * D + d] = ov;
        }
      }
    }
    if (tid < nh && single && p.lse != nullptr)
      p.lse[(int64_t)(head0 + tid) * p.S + s] = l_s[tid] > 0.f ? m_s[tid] + logf(l_s[tid]) : -INFINITY;
    if (tid < nh && p.so.m != nullptr && (!single || p.so.normalized)) {
      const int64_t pi = ((int64_t)s * p.H + head0 + tid) * p.so.num_partitions + part;
      p.so.m[pi] = m_s[tid];
      p.so.l[pi] = l_s[tid];
    }
  }
}

// ---- split-KV reduce -----------------------------------------------------------
// Reference: paged_attention_v2_reduce_kernel
// (csrc/attention/attention_kernels.cuh:556-658).  Sequences that fit in one
// partition were written to `out` by the main kernel already.
struct ReduceParams {
  const int32_t* seq_lens;
  void* out;
  float* lse;  // optional [H, S]
  SplitOut so;
  int S, H, D;
};

template <typename T, bool NORMALIZED>
__global__ void __launch_bounds__(kThreads) split_reduce_kernel(ReduceParams p) {
  const int head = blockIdx.x, s = blockIdx.y;
  const int L = p.seq_lens[s];
  const int n = (L + p.so.chunk - 1) / p.so.chunk;
  if (n <= 1) return;
  const int P = p.so.num_partitions;
  const int64_t base = ((int64_t)s * p.H + head) * P;
  extern __shared__ float w_s[];  // [n] rescaled weights
  __shared__ float red[kWarps];
  const int tid = threadIdx.x, lane = tid % kWarpSize, warp = tid / kWarpSize;

  float m = -FLT_MAX;
  for (int j = tid; j < n; j += kThreads) m = fmaxf(m, p.so.m[base + j]);
  m = warp_max(m);
  if (lane == 0) red[warp] = m;
  __syncthreads();
  m = red[0];
#pragma unroll
  for (int w = 1; w < kWarps; ++w) m = fmaxf(m, red[w]);
  __syncthreads();

  float sum = 0.f;
  for (int j = tid; j < n; j += kThreads) {
    const float e = expf(p.so.m[base + j] - m);
    // normalized partials are weighted by their rescaled exp-sum; raw ones
    // only by the max correction
    w_s[j] = NORMALIZED ? p.so.l[base + j] * e : e;
    sum += p.so.l[base + j] * e;
  }
  sum = warp_sum(sum);
  if (lane == 0) red[warp] = sum;
  __syncthreads();
  sum = 0.f;
#pragma unroll
  for (int w = 0; w < kWarps; ++w) sum += red[w];
  const float inv = __fdividef(1.f, sum + 1e-6f);
  if (tid == 0 && p.lse != nullptr) p.lse[(int64_t)head * p.S + s] = m + logf(sum);

  // Weighted sum of the partitions' rows.  The threads split into groups of D / kPack lanes: a group reads whole rows
  // with 16-byte loads and takes the partitions j == group (mod groups), so that at P = 256 (ctx 131072) a thread
  // walks 32 rows with independent loads in flight instead of 256 one after the other (the first version: 35 us of a
  // 670 us step at C5); the groups' sums meet in shared memory.
  using TO = typename std::conditional<NORMALIZED, T, float>::type;
  constexpr int kPack = 16 / sizeof(TO);
  __shared__ float acc_s[kThreads * kPack];  // [groups][D]
  const TO* o = reinterpret_cast<const TO*>(p.so.o) + base * p.D;
  const int lanes = p.D / kPack;           // 16-byte packs per row (head sizes are multiples of 8)
  const int groups = kThreads / lanes;
  const int grp = tid / lanes, pk = tid - grp * lanes;
  __syncthreads();  // w_s is complete
  if (grp < groups) {
    float acc[kPack];
#pragma unroll
    for (int i = 0; i < kPack; ++i) acc[i] = 0.f;
#pragma unroll 4
    for (int j = grp; j < n; j += groups) {
      const Vec16 v = ld_global_16(o + (int64_t)j * p.D + pk * kPack);
      const float w = w_s[j];
      const TO* e = reinterpret_cast<const TO*>(&v);
#pragma unroll
      for (int i = 0; i < kPack; ++i) acc[i] = fmaf(to_f32<TO>(e[i]), w, acc[i]);
    }
#pragma unroll
    for (int i = 0; i < kPack; ++i) acc_s[grp * p.D + pk * kPack + i] = acc[i];
  }
  __syncthreads();
  for (int d = tid; d < p.D; d += kThreads) {
    float a = 0.f;
    for (int g = 0; g < groups; ++g) a += acc_s[g * p.D + d];
    reinterpret_cast<T*>(p.out)[((int64_t)s * p.H + head) * p.D + d] = from_f32<T>(a * inv);
  }
}

template <typename T, typename CT>
int launch_generic_t(const GenericParams& gp, int S, int P, cudaStream_t s) {
  dim3 grid(gp.Hkv, S, P), block(kThreads);
  paged_attention_generic_kernel<T, CT><<<grid, block, 0, s>>>(gp);
  count_launch();
  return check_launch("paged_attention_generic");
}

}  // namespace

int launch_attention_generic(const b200_paged_attn_params_t& p, const SplitOut& so, cudaStream_t s) {
  B200_REQUIRE(so.chunk > 0 && so.chunk <= kMaxChunk && so.chunk % p.block_size == 0,
               B200_ERR_INVALID_ARGUMENT, "generic kernel: bad chunk %d", so.chunk);
  B200_REQUIRE(p.head_size <= kMaxD, B200_ERR_UNSUPPORTED, "Unsupported head size: %d", p.head_size);
  GenericParams gp;
  gp.q = p.query; gp.kc = p.key_cache; gp.vc = p.value_cache;
  gp.bt = p.block_tables; gp.seq_lens = p.seq_lens; gp.alibi = p.alibi_slopes;
  gp.k_scale = p.k_scale; gp.v_scale = p.v_scale; gp.out = p.out; gp.lse = p.out_lse; gp.so = so;
  gp.q_stride = p.q_stride; gp.kv_block_stride = p.kv_block_stride; gp.kv_head_stride = p.kv_head_stride;
  gp.S = p.num_seqs; gp.H = p.num_heads; gp.Hkv = p.num_kv_heads; gp.G = p.num_heads / p.num_kv_heads;
  gp.D = p.head_size; gp.B = p.block_size; gp.max_blocks = p.max_num_blocks_per_seq;
  gp.scale = p.scale;
  const int P = so.num_partitions;
  const bool fp8 = p.kv_dtype == B200_KV_FP8_E4M3;
  switch (p.dtype) {
    case B200_DTYPE_F32:
      return fp8 ? launch_generic_t<float, uint8_t>(gp, p.num_seqs, P, s)
                 : launch_generic_t<float, float>(gp, p.num_seqs, P, s);
    case B200_DTYPE_F16:
      return fp8 ? launch_generic_t<__half, uint8_t>(gp, p.num_seqs, P, s)
                 : launch_generic_t<__half, __half>(gp, p.num_seqs, P, s);
    case B200_DTYPE_BF16:
      return fp8 ? launch_generic_t<__nv_bfloat16, uint8_t>(gp, p.num_seqs, P, s)
                 : launch_generic_t<__nv_bfloat16, __nv_bfloat16>(gp, p.num_seqs, P, s);
    default:
      set_error("Unsupported data type: %d", p.dtype);
      return B200_ERR_UNSUPPORTED;
  }
}

int launch_split_reduce(const b200_paged_attn_params_t& p, const SplitOut& so, cudaStream_t s) {
  if (so.num_partitions <= 1) return B200_OK;  // every sequence was written directly
  ReduceParams rp;
  rp.seq_lens = p.seq_lens; rp.out = p.out; rp.lse = p.out_lse; rp.so = so; rp.S = p.num_seqs; rp.H = p.num_heads;
  rp.D = p.head_size;
  dim3 grid(p.num_heads, p.num_seqs), block(kThreads);
  const size_t smem = (size_t)so.num_partitions * sizeof(float);
#define RUN(T)                                                                   \
  do {                                                                           \
    if (so.normalized) split_reduce_kernel<T, true><<<grid, block, smem, s>>>(rp); \
    else split_reduce_kernel<T, false><<<grid, block, smem, s>>>(rp);            \
  } while (0)
  switch (p.dtype) {
    case B200_DTYPE_F32: RUN(float); break;
    case B200_DTYPE_F16: RUN(__half); break;
    default: RUN(__nv_bfloat16); break;
  }
#undef RUN
  count_launch();
  return check_launch("paged_attention_split_reduce");
}

}  // namespace b200
