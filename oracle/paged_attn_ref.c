/*
 * CPU restatement of the reference's paged-attention decode path in plain C — TEST INFRASTRUCTURE.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg may load
 * this; the product (vllm_metax_b200/) never does and has no CPU fallback.
 *
 * It follows the reference's CUDA kernels phase by phase, on host cores (OpenMP over (seq, kv head)):
 *   paged attention   csrc/attention/attention_kernels.cuh:79-485
 *     logits  = scale * <q, k_t> (+ alibi_slope * (t - L + 1))            :283-286
 *     softmax with exp(logit - max) and 1 / (sum + 1e-6)                  :302-335
 *     out     = sum_t p_t * v_t, tokens past seq_len never contribute     :348-423
 *     cache addressing through block_tables, kv_block_stride, kv_head_stride :246-271,391-397
 *   reshape_and_cache        csrc/cache_kernels.cu:203-255
 *   reshape_and_cache_flash  csrc/cache_kernels.cu:271-344
 *   fp8 e4m3: quant = satfinite_rn(x / scale), dequant = e4m3 * scale     call sites :249-252, attention_kernels.cuh:276-277
 * Arithmetic is fp32 throughout (the reference rounds probabilities to bf16 before P.V; the parity bar
 * of 2e-3 covers the difference, see tests/test_oracle_c.py).  Parity status: pinned through
 * torch_oracle (tests/test_oracle_c.py compares the two; torch_oracle is pinned to the reference's own
 * Python oracles by tests/golden).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

enum { DT_F32 = 0, DT_F16 = 1, DT_BF16 = 2, DT_U8 = 3 };

static inline float bf16_to_f32(uint16_t h) {
  uint32_t u = (uint32_t)h << 16;
  float f;
  memcpy(&f, &u, 4);
  return f;
}
static inline uint16_t f32_to_bf16(float f) { /* round to nearest even */
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40);
  u += 0x7fffu + ((u >> 16) & 1u);
  return (uint16_t)(u >> 16);
}
static inline float f16_to_f32(uint16_t h) {
  uint32_t s = (uint32_t)(h & 0x8000u) << 16, e = (h >> 10) & 0x1f, m = h & 0x3ffu, u;
  if (e == 0) {
    if (m == 0) u = s;
    else {
      int sh = 0;
      while (!(m & 0x400u)) { m <<= 1; ++sh; }
      u = s | ((uint32_t)(113 - sh) << 23) | ((m & 0x3ffu) << 13);
    }
  } else if (e == 31) u = s | 0x7f800000u | (m << 13);
  else u = s | ((e + 112) << 23) | (m << 13);
  float f;
  memcpy(&f, &u, 4);
  return f;
}
static inline uint16_t f32_to_f16(float f) { /* round to nearest even */
  uint32_t u;
  memcpy(&u, &f, 4);
  uint32_t s = (u >> 16) & 0x8000u, a = u & 0x7fffffffu;
  if (a > 0x7f800000u) return (uint16_t)(s | 0x7e00u);
  if (a >= 0x47800000u) return (uint16_t)(s | 0x7c00u); /* overflow -> inf */
  if (a < 0x38800000u) {                                 /* subnormal or zero */
    if (a < 0x33000000u) return (uint16_t)s;
    int e = (int)(a >> 23);
    uint32_t m = (a & 0x7fffffu) | 0x800000u;
    int shift = 126 - e; /* 14..24 */
    uint32_t r = m >> shift, rem = m & ((1u << shift) - 1), half = 1u << (shift - 1);
    if (rem > half || (rem == half && (r & 1))) ++r;
    return (uint16_t)(s | r);
  }
  uint32_t r = a - 0x38000000u; /* rebias */
  uint32_t m = r >> 13, rem = r & 0x1fffu;
  if (rem > 0x1000u || (rem == 0x1000u && (m & 1))) ++m;
  return (uint16_t)(s | m);
}

static float g_e4m3_lut[256];
static int g_lut_ready = 0;
static void init_lut(void) {
  if (g_lut_ready) return;
  for (int c = 0; c < 256; ++c) {
    int e = (c >> 3) & 15, m = c & 7;
    float v = e == 0 ? ldexpf((float)m, -9) : ldexpf(1.0f + (float)m / 8.0f, e - 7);
    if (e == 15 && m == 7) v = NAN;
    g_e4m3_lut[c] = (c & 0x80) ? -v : v;
  }
  g_lut_ready = 1;
}

/* fp32 -> e4m3 byte, round to nearest even, saturate to +-448 (NaN -> 0x7f | sign) */
uint8_t ref_f32_to_e4m3(float x) {
  uint32_t u;
  memcpy(&u, &x, 4);
  uint8_t sign = (uint8_t)((u >> 24) & 0x80);
  uint32_t a = u & 0x7fffffffu;
  if (a > 0x7f800000u) return sign | 0x7f;
  float ax = fabsf(x);
  if (ax >= 448.0f) return sign | 0x7e;
  if (ax < ldexpf(1.0f, -10)) return sign; /* below half the smallest subnormal (ties to even -> 0) */
  int ex;
  (void)frexpf(ax, &ex); /* ax = f * 2^ex, f in [0.5,1) -> unbiased exponent ex-1 */
  int qexp = ex - 1 < -6 ? -6 : ex - 1;
  float q = ax * ldexpf(1.0f, 3 - qexp); /* in units of the grid step; exact scaling */
  float r = nearbyintf(q);              /* default rounding mode: nearest even */
  int k = (int)r;                       /* 0..16 */
  int code;
  if (qexp == -6 && k < 8) code = k;
  else if (k >= 16) code = ((qexp + 1 + 7) << 3);
  else code = ((qexp + 7) << 3) | (k - 8);
  if (code > 0x7e) code = 0x7e;
  return sign | (uint8_t)code;
}

static inline float load_elem(const void* base, int64_t idx, int dt) {
  switch (dt) {
    case DT_F32: return ((const float*)base)[idx];
    case DT_F16: return f16_to_f32(((const uint16_t*)base)[idx]);
    case DT_BF16: return bf16_to_f32(((const uint16_t*)base)[idx]);
    default: return g_e4m3_lut[((const uint8_t*)base)[idx]];
  }
}
static inline void store_elem(void* base, int64_t idx, int dt, float v) {
  switch (dt) {
    case DT_F32: ((float*)base)[idx] = v; break;
    case DT_F16: ((uint16_t*)base)[idx] = f32_to_f16(v); break;
    case DT_BF16: ((uint16_t*)base)[idx] = f32_to_bf16(v); break;
    default: ((uint8_t*)base)[idx] = ref_f32_to_e4m3(v); break;
  }
}

/*
 * out[S,H,D] (dtype) <- paged attention.  cache_dt: element type of the caches (dtype, or DT_U8 = e4m3
 * with k_scale / v_scale).  x = 16 / sizeof(cache element).  Strides in elements.
 */
int ref_paged_attention(void* out, const void* query, const void* key_cache, const void* value_cache,
                        const int32_t* block_tables, const int32_t* seq_lens, const float* alibi_slopes,
                        float scale, float k_scale, float v_scale, int S, int H, int Hkv, int D, int B,
                        int max_blocks, int64_t q_stride, int64_t kv_block_stride, int64_t kv_head_stride, int dtype,
                        int cache_dt, int num_threads) {
  init_lut();
  const int G = H / Hkv;
  const int x = cache_dt == DT_U8 ? 16 : (cache_dt == DT_F32 ? 4 : 8);
  const float kmul = cache_dt == DT_U8 ? k_scale : 1.0f, vmul = cache_dt == DT_U8 ? v_scale : 1.0f;
  int status = 0;
#pragma omp parallel for collapse(2) schedule(dynamic, 1) num_threads(num_threads)
  for (int s = 0; s < S; ++s) {
    for (int kvh = 0; kvh < Hkv; ++kvh) {
      const int L = seq_lens[s];
      if (L <= 0) continue;
      float* q = (float*)malloc(sizeof(float) * (size_t)G * D);
      float* logits = (float*)malloc(sizeof(float) * (size_t)G * L);
      float* acc = (float*)calloc((size_t)G * D, sizeof(float));
      float* kblk = (float*)malloc(sizeof(float) * (size_t)D * B);
      if (!q || !logits || !acc || !kblk) { status = 1; free(q); free(logits); free(acc); free(kblk); continue; }
      for (int g = 0; g < G; ++g)
        for (int d = 0; d < D; ++d) q[g * D + d] = load_elem(query, s * q_stride + (int64_t)(kvh * G + g) * D + d, dtype);
      const int nblk = (L + B - 1) / B;
      /* phase 1: logits (attention_kernels.cuh:216-297) */
      for (int b = 0; b < nblk; ++b) {
        const int64_t base = (int64_t)block_tables[(int64_t)s * max_blocks + b] * kv_block_stride + (int64_t)kvh * kv_head_stride;
        const int nt = (b + 1) * B <= L ? B : L - b * B;
        for (int i = 0; i < D * B; ++i) kblk[i] = load_elem(key_cache, base + i, cache_dt); /* [D/x][B][x] */
        for (int g = 0; g < G; ++g)
          for (int t = 0; t < nt; ++t) {
            float dot = 0.f;
            for (int j = 0; j < D / x; ++j)
              for (int e = 0; e < x; ++e) dot += q[g * D + j * x + e] * kblk[(j * B + t) * x + e];
            const int tok = b * B + t;
            float lg = dot * scale * kmul;
            if (alibi_slopes) lg += alibi_slopes[kvh * G + g] * (float)(tok - L + 1);
            logits[(size_t)g * L + tok] = lg;
          }
      }
      /* phase 2: softmax statistics (:302-335) */
      for (int g = 0; g < G; ++g) {
        float m = -INFINITY, sum = 0.f;
        for (int t = 0; t < L; ++t) m = fmaxf(m, logits[(size_t)g * L + t]);
        for (int t = 0; t < L; ++t) {
          const float e = expf(logits[(size_t)g * L + t] - m);
          logits[(size_t)g * L + t] = e;
          sum += e;
        }
        const float inv = 1.0f / (sum + 1e-6f);
        for (int t = 0; t < L; ++t) logits[(size_t)g * L + t] *= inv;
      }
      /* phase 3: P.V (:348-423); value_cache block is [D][B] */
      for (int b = 0; b < nblk; ++b) {
        const int64_t base = (int64_t)block_tables[(int64_t)s * max_blocks + b] * kv_block_stride + (int64_t)kvh * kv_head_stride;
        const int nt = (b + 1) * B <= L ? B : L - b * B;
        for (int i = 0; i < D * B; ++i) kblk[i] = load_elem(value_cache, base + i, cache_dt);
        for (int g = 0; g < G; ++g)
          for (int d = 0; d < D; ++d) {
            float a = 0.f;
            for (int t = 0; t < nt; ++t) a += logits[(size_t)g * L + b * B + t] * kblk[d * B + t];
            acc[g * D + d] += a;
          }
      }
      for (int g = 0; g < G; ++g)
        for (int d = 0; d < D; ++d) store_elem(out, ((int64_t)s * H + kvh * G + g) * D + d, dtype, acc[g * D + d] * vmul);
      free(q); free(logits); free(acc); free(kblk);
    }
  }
  return status;
}

/* key/value [T, H, D] (token strides key_stride/value_stride) -> v1/v2 layout caches (contiguous). */
void ref_reshape_and_cache(const void* key, const void* value, void* key_cache, void* value_cache,
                           const int64_t* slot_mapping, float k_scale, float v_scale, int T, int H, int D, int B, int x,
                           int64_t key_stride, int64_t value_stride, int dtype, int cache_dt) {
  init_lut();
  for (int t = 0; t < T; ++t) {
    const int64_t slot = slot_mapping[t];
    if (slot < 0) continue;
    const int64_t blk = slot / B, off = slot % B;
    for (int h = 0; h < H; ++h)
      for (int d = 0; d < D; ++d) {
        float kv = load_elem(key, t * key_stride + (int64_t)h * D + d, dtype);
        float vv = load_elem(value, t * value_stride + (int64_t)h * D + d, dtype);
        if (cache_dt == DT_U8) { kv = kv / k_scale; vv = vv / v_scale; }
        const int64_t kdst = (((blk * H + h) * (D / x) + d / x) * B + off) * x + d % x;
        const int64_t vdst = ((blk * H + h) * D + d) * B + off;
        if (cache_dt == DT_U8 || cache_dt != dtype) {
          store_elem(key_cache, kdst, cache_dt, kv);
          store_elem(value_cache, vdst, cache_dt, vv);
        } else { /* plain copy keeps the bits */
          const int eb = dtype == DT_F32 ? 4 : 2;
          memcpy((char*)key_cache + kdst * eb, (const char*)key + (t * key_stride + (int64_t)h * D + d) * eb, eb);
          memcpy((char*)value_cache + vdst * eb, (const char*)value + (t * value_stride + (int64_t)h * D + d) * eb, eb);
        }
      }
  }
}

/* flash layout: token row at blk*block_stride + off*page_stride + h*head_stride (elements). */
void ref_reshape_and_cache_flash(const void* key, const void* value, void* key_cache, void* value_cache,
                                 const int64_t* slot_mapping, float k_scale, float v_scale, int T, int H, int D, int B,
                                 int64_t key_stride, int64_t value_stride, int64_t block_stride, int64_t page_stride,
                                 int64_t head_stride, int dtype, int cache_dt) {
  init_lut();
  for (int t = 0; t < T; ++t) {
    const int64_t slot = slot_mapping[t];
    if (slot < 0) continue;
    const int64_t blk = slot / B, off = slot % B;
    for (int h = 0; h < H; ++h)
      for (int d = 0; d < D; ++d) {
        const int64_t dst = blk * block_stride + off * page_stride + (int64_t)h * head_stride + d;
        const int64_t ksrc = t * key_stride + (int64_t)h * D + d, vsrc = t * value_stride + (int64_t)h * D + d;
        if (cache_dt == DT_U8) {
          store_elem(key_cache, dst, DT_U8, load_elem(key, ksrc, dtype) / k_scale);
          store_elem(value_cache, dst, DT_U8, load_elem(value, vsrc, dtype) / v_scale);
        } else {
          const int eb = dtype == DT_F32 ? 4 : 2;
          memcpy((char*)key_cache + dst * eb, (const char*)key + ksrc * eb, eb);
          memcpy((char*)value_cache + dst * eb, (const char*)value + vsrc * eb, eb);
        }
      }
  }
}

/* bulk fp8 conversion (csrc/cache_kernels.cu:543-614): one side is DT_U8 */
void ref_convert_fp8(void* dst, const void* src, int64_t n, float scale, int dst_dt, int src_dt) {
  init_lut();
  for (int64_t i = 0; i < n; ++i) {
    if (dst_dt == DT_U8) store_elem(dst, i, DT_U8, load_elem(src, i, src_dt) / scale);
    else store_elem(dst, i, dst_dt, load_elem(src, i, DT_U8) * scale);
  }
}
