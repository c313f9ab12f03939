// merge_attn_states: combine two partial attention results (prefix / suffix,
// disjoint key sets) from their outputs and log-sum-exps — section 2.2 of
// arXiv 2501.01005, as the reference op does (csrc/attention/merge_attn_states.cu:14-88).
//
//   m = max(p_lse, s_lse)            (+inf lse is read as -inf, :45-46)
//   w_p = exp(p_lse - m), w_s = exp(s_lse - m)
//   out = prefix * w_p/(w_p+w_s) + suffix * w_s/(w_p+w_s)     fp32, rounded once
//   out_lse = log(w_p + w_s) + m
//
// HBM-bound elementwise op (3 x N x H x D elements moved, 2 lse reads per head):
// one CTA per token, 16-byte packs, two packs in flight per thread, 32-bit index
// arithmetic.  The per-(token, head) weights are recomputed by each of the D/pack
// threads of a head (the lse values are broadcast hits in L1); that keeps the
// kernel barrier-free.
#include "common.cuh"

namespace b200 {
namespace {

// one CTA per token; thread t walks the token's packs t, t + 256, ... (32-bit index arithmetic only), two packs per
// step with all four 16-byte loads issued before the arithmetic
template <typename T>
__global__ void __launch_bounds__(256) merge_attn_states_kernel(b200_merge_attn_states_params_t p) {
  constexpr int kPack = 16 / sizeof(T);
  const uint32_t packs_per_head = p.head_size / kPack;
  const uint32_t packs = p.num_heads * packs_per_head;     // per token
  const uint64_t token = blockIdx.x;
  const T* pbase = reinterpret_cast<const T*>(p.prefix_output) + token * p.prefix_stride;
  const T* sbase = reinterpret_cast<const T*>(p.suffix_output) + token * p.suffix_stride;
  T* obase = reinterpret_cast<T*>(p.output) + token * p.output_stride;

  auto weights = [&](uint32_t head, float& pw, float& sw, float& lse) {
    float pl = p.prefix_lse[(uint64_t)head * p.num_tokens + token];
    float sl = p.suffix_lse[(uint64_t)head * p.num_tokens + token];
    pl = isinf(pl) ? -INFINITY : pl;
    sl = isinf(sl) ? -INFINITY : sl;
    const float m = fmaxf(pl, sl);
    const float pse = expf(pl - m);
    const float sse = expf(sl - m);
    const float se = pse + sse;
    pw = pse / se;
    sw = sse / se;
    lse = logf(se) + m;  // :84-85
  };
  auto combine = [&](const Vec16& pv, const Vec16& sv, float pw, float sw) {
    Vec16 ov;
    const T* pe = reinterpret_cast<const T*>(&pv);
    const T* se_ = reinterpret_cast<const T*>(&sv);
    T* oe = reinterpret_cast<T*>(&ov);
#pragma unroll
    for (int i = 0; i < kPack; ++i)
      oe[i] = from_f32<T>(__fmaf_rn(to_f32<T>(pe[i]), pw, __fmul_rn(to_f32<T>(se_[i]), sw)));  // :73
    return ov;
  };
  uint32_t i = threadIdx.x;
  for (; i + 256 < packs; i += 512) {
    const uint32_t j = i + 256;
    const Vec16 pv0 = ld_global_16_coherent(pbase + (uint64_t)i * kPack), sv0 = ld_global_16_coherent(sbase + (uint64_t)i * kPack);
    const Vec16 pv1 = ld_global_16_coherent(pbase + (uint64_t)j * kPack), sv1 = ld_global_16_coherent(sbase + (uint64_t)j * kPack);
    const uint32_t h0 = i / packs_per_head, h1 = j / packs_per_head;
    float pw0, sw0, l0, pw1, sw1, l1;
    weights(h0, pw0, sw0, l0);
    weights(h1, pw1, sw1, l1);
    *reinterpret_cast<Vec16*>(obase + (uint64_t)i * kPack) = combine(pv0, sv0, pw0, sw0);
    *reinterpret_cast<Vec16*>(obase + (uint64_t)j * kPack) = combine(pv1, sv1, pw1, sw1);
    if (p.output_lse != nullptr) {
      if (i % packs_per_head == 0) p.output_lse[(uint64_t)h0 * p.num_tokens + token] = l0;
      if (j % packs_per_head == 0) p.output_lse[(uint64_t)h1 * p.num_tokens + token] = l1;
    }
  }
  for (; i < packs; i += 256) {
    const Vec16 pv = ld_global_16_coherent(pbase + (uint64_t)i * kPack), sv = ld_global_16_coherent(sbase + (uint64_t)i * kPack);
    const uint32_t h = i / packs_per_head;
    float pw, sw, l;
    weights(h, pw, sw, l);
    *reinterpret_cast<Vec16*>(obase + (uint64_t)i * kPack) = combine(pv, sv, pw, sw);
    if (p.output_lse != nullptr && i % packs_per_head == 0) p.output_lse[(uint64_t)h * p.num_tokens + token] = l;
  }
}

inline bool aligned16(const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; }

}  // namespace
}  // namespace b200

extern "C" int b200_merge_attn_states(const b200_merge_attn_states_params_t* p, void* stream) {
  using namespace b200;
  B200_REQUIRE(p != nullptr, B200_ERR_INVALID_ARGUMENT, "params is NULL");
  B200_REQUIRE(p->num_tokens >= 0 && p->num_heads > 0 && p->head_size > 0, B200_ERR_INVALID_ARGUMENT,
               "bad sizes: tokens=%d heads=%d head_size=%d", p->num_tokens, p->num_heads, p->head_size);
  B200_REQUIRE(p->dtype == B200_DTYPE_F32 || p->dtype == B200_DTYPE_F16 || p->dtype == B200_DTYPE_BF16,
               B200_ERR_UNSUPPORTED, "Unsupported data type of O: %d", p->dtype);
  if (p->num_tokens == 0) return B200_OK;
  B200_REQUIRE(p->output && p->prefix_output && p->prefix_lse && p->suffix_output && p->suffix_lse,
               B200_ERR_INVALID_ARGUMENT, "NULL tensor pointer");
  const int pack = 16 / dtype_bytes(p->dtype);
  // csrc/attention/merge_attn_states.cu:138-139
  B200_REQUIRE(p->head_size % pack == 0, B200_ERR_UNSUPPORTED, "headsize must be multiple of pack_size:%d", pack);
  const int64_t row = (int64_t)p->num_heads * p->head_size;
  B200_REQUIRE(p->output_stride >= row && p->prefix_stride >= row && p->suffix_stride >= row,
               B200_ERR_INVALID_ARGUMENT, "token strides smaller than num_heads*head_size");
  B200_REQUIRE(aligned16(p->output) && aligned16(p->prefix_output) && aligned16(p->suffix_output) &&
                   p->output_stride % pack == 0 && p->prefix_stride % pack == 0 && p->suffix_stride % pack == 0,
               B200_ERR_INVALID_ARGUMENT, "outputs must be 16-byte aligned with 16-byte aligned token rows");
  const int threads = 256;
  const uint64_t blocks = (uint64_t)p->num_tokens;   // one CTA per token
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  switch (p->dtype) {
    case B200_DTYPE_F32: merge_attn_states_kernel<float><<<(unsigned)blocks, threads, 0, s>>>(*p); break;
    case B200_DTYPE_F16: merge_attn_states_kernel<__half><<<(unsigned)blocks, threads, 0, s>>>(*p); break;
    default: merge_attn_states_kernel<__nv_bfloat16><<<(unsigned)blocks, threads, 0, s>>>(*p); break;
  }
  count_launch();
  return check_launch("merge_attn_states");
}
