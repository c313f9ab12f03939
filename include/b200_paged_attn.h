/*
 * b200_paged_attn.h — C ABI of the B200-native paged-attention decode path.
 *
 * This is the drop-in boundary.  Every entry point takes plain
 * device pointers, sizes, element strides and a cudaStream_t passed as void*;
 * no torch type crosses it.  The reference has no C ABI for this path: it
 * registers torch-dispatcher ops (csrc/torch_bindings.cpp:45-69, :394-411,
 * :423-426) whose C++ prototypes are csrc/ops.h:32-53 and csrc/cache.h:21-40.
 * Each function below names the reference op it stands in for; the torch shim
 * (vllm_metax_b200/csrc/torch_shim.cpp) turns the op's tensors into these
 * arguments, exactly as the reference launchers turn them into kernel
 * arguments (csrc/attention/paged_attention_v1.cu:43-125,
 * csrc/cache_kernels.cu:407-488).
 *
 * Conventions
 *   - all strides are in ELEMENTS of the tensor they belong to;
 *   - every pointer is a device pointer unless stated otherwise;
 *   - calls only enqueue work on `stream`; they never synchronise, never
 *     allocate device memory and never read device memory on the host, so
 *     they are CUDA-graph capturable;
 *   - return value: B200_OK (0) or an error code; b200_last_error() gives
 *     the message of the last failure on the calling thread.  The library
 *     never calls exit()/abort().
 */
#ifndef B200_PAGED_ATTN_H_
#define B200_PAGED_ATTN_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_PAGED_ATTN_ABI_VERSION 3

/* v2 partition size: template default of the reference launcher
 * (csrc/attention/paged_attention_v2.cu:45) and of its callers
 * (tests/kernels/attention/test_attention.py:30).  It fixes the layout of
 * exp_sums / max_logits / tmp_out, so it is ABI. */
#define B200_PARTITION_SIZE 512

typedef enum b200_status {
  B200_OK = 0,
  B200_ERR_INVALID_ARGUMENT = 1, /* null pointer, negative size, bad enum     */
  B200_ERR_UNSUPPORTED = 2,      /* head size / block size / dtype / sparse   */
  B200_ERR_CUDA = 3,             /* a CUDA runtime call or launch failed      */
  B200_ERR_WORKSPACE = 4         /* workspace missing or too small            */
} b200_status_t;

/* dtype of query / out / key / value (reference: float, half, bf16 —
 * csrc/quantization/fp8/metax/quant_utils.cuh:29-39). */
typedef enum b200_dtype {
  B200_DTYPE_F32 = 0,
  B200_DTYPE_F16 = 1,
  B200_DTYPE_BF16 = 2,
  B200_DTYPE_U8 = 3 /* only as the fp8 side of b200_convert_fp8 */
} b200_dtype_t;

/* kv_cache_dtype string of the ops ("auto" | "fp8" | "fp8_e4m3";
 * csrc/cache_kernels.cu:581,595).  "fp8_e5m2" is listed by the reference enum
 * (csrc/attention/dtype_fp8.cuh:9-13) but implemented nowhere: unsupported. */
typedef enum b200_kv_dtype {
  B200_KV_AUTO = 0,
  B200_KV_FP8_E4M3 = 1
} b200_kv_dtype_t;

/* Which kernel family served the last attention call (for tests / bench). */
typedef enum b200_attn_variant {
  B200_VARIANT_AUTO = 0,      /* choose: stream-MMA when eligible, else generic (tcgen05 fp8: opt-in)  */
  B200_VARIANT_GENERIC = 1,   /* FMA kernel, any supported shape                */
  B200_VARIANT_STREAM_MMA = 2,/* TMA ring + mma.sync persistent kernel          */
  B200_VARIANT_UMMA_FP8 = 3   /* TMA ring + tcgen05.mma (fp8 cache, block 16, GQA group <= 8) */
} b200_attn_variant_t;

/*
 * Arguments of paged_attention_v1 / paged_attention_v2.
 *   reference ops : csrc/torch_bindings.cpp:45-55 (v1), :58-69 (v2)
 *   reference C++ : csrc/ops.h:32-53
 * Layouts (csrc/attention/attention_kernels.cuh:79-100):
 *   out         [num_seqs, num_heads, head_size]              dtype, contiguous
 *   query       [num_seqs, num_heads, head_size]              dtype, row stride q_stride
 *   key_cache   [num_blocks, num_kv_heads, head_size/x, block_size, x]
 *   value_cache [num_blocks, num_kv_heads, head_size, block_size]
 *               x = 16 / sizeof(cache element); caches are addressed with
 *               kv_block_stride / kv_head_stride taken from key_cache
 *               (paged_attention_v1.cu:59-60), cache element = dtype (auto)
 *               or one e4m3 byte (fp8)
 *   block_tables[num_seqs, max_num_blocks_per_seq] int32
 *   seq_lens    [num_seqs] int32
 *   alibi_slopes[num_heads] float32 or NULL
 *   k_scale, v_scale : one float32 each, ON THE DEVICE (read by the kernel)
 * v2 only (caller-allocated scratch, P = ceil(max_seq_len / 512)):
 *   exp_sums, max_logits [num_seqs, num_heads, P] float32
 *   tmp_out              [num_seqs, num_heads, P, head_size] dtype
 * workspace: v1 needs scratch for its split-KV partials (the reference op
 *   has none because its v1 never splits).  Size it with
 *   b200_paged_attention_workspace_bytes().  It must be ZERO-FILLED ONCE before
 *   its first use (its first 1 MiB holds arrival counters that every completed
 *   launch leaves at zero again, whatever the shapes) and may then be reused by
 *   calls of any shape on the same stream; calls that may overlap need separate
 *   workspaces.  v2 needs none (NULL, 0).
 */
typedef struct b200_paged_attn_params {
  void* out;
  const void* query;
  const void* key_cache;
  const void* value_cache;
  const int32_t* block_tables;
  const int32_t* seq_lens;
  const float* alibi_slopes; /* nullable */
  const float* k_scale;
  const float* v_scale;
  float* exp_sums;   /* v2 */
  float* max_logits; /* v2 */
  void* tmp_out;     /* v2 */
  void* workspace;   /* v1 */
  uint64_t workspace_bytes;
  int64_t q_stride;
  int64_t kv_block_stride;
  int64_t kv_head_stride;
  int32_t num_seqs;
  int32_t num_heads;
  int32_t num_kv_heads;
  int32_t head_size;
  int32_t block_size;
  int32_t max_seq_len;
  int32_t max_num_blocks_per_seq;
  int32_t dtype;    /* b200_dtype_t of query/out */
  int32_t kv_dtype; /* b200_kv_dtype_t           */
  float scale;
  /* block-sparse arguments of the op signature; dense only
   * (blocksparse_vert_stride <= 1), as in every reference test
   * (tests/kernels/attention/test_attention.py:217-222). */
  int32_t tp_rank;
  int32_t blocksparse_local_blocks;
  int32_t blocksparse_vert_stride;
  int32_t blocksparse_block_size;
  int32_t blocksparse_head_sliding_step;
  /* ABI v2.  Optional log-sum-exp of the scaled logits per (head, seq), float32
   * [num_heads, num_seqs] (the layout merge_attn_states reads: prefix_lse /
   * suffix_lse [NUM_HEADS, NUM_TOKENS], csrc/attention/merge_attn_states.cu:43-44),
   * natural-log units; NULL = not wanted (the reference ops have no such
   * output).  A sequence of length 0 yields -inf.  Feeds b200_merge_attn_states
   * (cascade / decode-context-parallel, vllm_metax/v1/attention/backends/flash_attn.py:794-872). */
  float* out_lse;
  /* ABI v2, b200_paged_attention_flash only: stride between consecutive tokens
   * of a block (key_cache.stride(1)), elements. */
  int64_t kv_page_stride;
  /* ABI v2, b200_paged_attention_flash only: key_cache.size(0).  Bounds the TMA
   * tensor map of the cache (no block id at or above it is ever dereferenced). */
  int64_t kv_num_blocks;
} b200_paged_attn_params_t;

/* Replaces torch.ops._C.paged_attention_v1 (csrc/torch_bindings.cpp:45-55;
 * csrc/attention/paged_attention_v1.cu:160-182). */
int b200_paged_attention_v1(const b200_paged_attn_params_t* params, void* stream);

/* Replaces torch.ops._C.paged_attention_v2 incl. its reduce kernel
 * (csrc/torch_bindings.cpp:58-69; csrc/attention/paged_attention_v2.cu:167-192;
 * csrc/attention/attention_kernels.cuh:556-658). */
int b200_paged_attention_v2(const b200_paged_attn_params_t* params, void* stream);

/*
 * Decode attention over the FLASH cache layout — the layout reshape_and_cache_flash
 * writes and the reference's serving path reads through the external flash-attention
 * library: flash_attn_with_kvcache(q, k_cache, v_cache, block_table, cache_seqlens,
 * softmax_scale, causal=True, alibi_slopes, ...) with one query token per sequence
 * (vllm_metax/v1/attention/backends/flash_attn.py:742-757; import in
 * vllm_metax/attention/utils/fa_utils.py:12).  Same math and arguments as
 * b200_paged_attention_v1, except
 *   key_cache / value_cache [num_blocks, block_size, num_kv_heads, head_size]
 *       addressed by kv_block_stride (stride 0), kv_page_stride (stride 1) and
 *       kv_head_stride (stride 2): NHD and HND storage both work (csrc/cache_kernels.cu:303);
 *   supported: bf16 query and cache (kv_dtype auto — the reference's flash path has
 *       no fp8 cache: fa_utils.py:26-30), head_size 128, block_size a multiple of 16,
 *       GQA group <= 16; anything else -> B200_ERR_UNSUPPORTED;
 *   out_lse (optional) as for v1; workspace as for v1.
 */
int b200_paged_attention_flash(const b200_paged_attn_params_t* params, void* stream);

/* Bytes of workspace b200_paged_attention_v1 needs for these shapes (pointer
 * fields of params are ignored).  Host-only, no CUDA call. */
uint64_t b200_paged_attention_workspace_bytes(const b200_paged_attn_params_t* params);

/* Re-zero the arrival counters at the head of a v1 / flash workspace (enqueued on `stream`; graph-capturable).
 * Never needed after launches that completed (each leaves the counters at zero).  It exists for hosts that recycle a
 * workspace from an unknown state (a pool, or a buffer some other code wrote into).  Note that a launch that
 * FAULTED cannot poison later calls through the workspace: a device fault is a sticky context error, after which no
 * call on that context runs at all.  No reference counterpart (the reference's v1 never splits). */
int b200_paged_attention_workspace_reset(void* workspace, uint64_t workspace_bytes, void* stream);

/*
 * b200_paged_attention_v1 without a caller workspace and without any device allocation: for code that cannot call
 * cudaMalloc (CUDA-graph capture, tight serving loops, harnesses that forbid it).  NO COUNTERPART IN THE
 * REFERENCE.
 *   - Same arguments and results as b200_paged_attention_v1; params->workspace / workspace_bytes are ignored.
 *   - The split-KV scratch is a static device array of B200_STATIC_SCRATCH_BYTES (32 MiB) inside the library; every
 *     GPU that loads the library holds it until the process exits.
 *   - Shapes for which b200_paged_attention_workspace_bytes() exceeds that size -> B200_ERR_WORKSPACE (every shape the
 *     streaming kernels serve fits: their workspace is a constant ~26.6 MB; the generic kernel's split-KV partials get
 *     the remaining ~6.9 MB).
 *   - ALL calls on a device share the array: calls that may overlap on the GPU (different streams) race on it —
 *     serialise them or use b200_paged_attention_v1 with one workspace per stream.  The array keeps the last call's
 *     contents; each call re-zeroes the 1 MiB of arrival counters on `stream` before its kernel.
 *   - Calls no cudaMalloc / cudaFree / allocating library routine.  The device of the call is the CURRENT device.
 */
#define B200_STATIC_SCRATCH_BYTES (32u << 20)
int b200_paged_attention_v1_static(const b200_paged_attn_params_t* params, void* stream);
uint64_t b200_paged_attention_static_scratch_bytes(void);

/*
 * Arguments of reshape_and_cache / reshape_and_cache_flash.
 *   reference ops : csrc/torch_bindings.cpp:394-400, :403-411
 *   reference C++ : csrc/cache.h:21-40; csrc/cache_kernels.cu:407-433, :450-488
 *   key, value  [num_tokens(+padding), num_heads, head_size] dtype,
 *               token strides key_stride / value_stride, heads*dims contiguous
 *   slot_mapping[num_tokens] int64; slot < 0 = padding, skipped
 *               (csrc/cache_kernels.cu:217-220, :285-289)
 * reshape_and_cache: contiguous v1/v2 layout, x = key_cache.size(4)
 *               (index math from sizes, csrc/cache_kernels.cu:235-242).
 * reshape_and_cache_flash: caches [num_blocks, block_size, num_heads, head_size]
 *               addressed by block_stride / page_stride / head_stride
 *               (= key_cache.stride(0/1/2)); NHD iff head_stride == head_size
 *               (csrc/cache_kernels.cu:303), else per-head rows (HND).
 */
typedef struct b200_cache_write_params {
  const void* key;
  const void* value;
  void* key_cache;
  void* value_cache;
  const int64_t* slot_mapping;
  const float* k_scale; /* device; read only when kv_dtype != auto */
  const float* v_scale;
  int64_t key_stride;
  int64_t value_stride;
  int64_t block_stride; /* flash only */
  int64_t page_stride;  /* flash only */
  int64_t head_stride;  /* flash only */
  int32_t num_tokens;   /* = slot_mapping.size(0) (csrc/cache_kernels.cu:417,469) */
  int32_t num_heads;
  int32_t head_size;
  int32_t block_size;
  int32_t x;        /* reshape_and_cache only */
  int32_t dtype;    /* b200_dtype_t of key/value */
  int32_t kv_dtype; /* b200_kv_dtype_t           */
} b200_cache_write_params_t;

/* Replaces torch.ops._C_cache_ops.reshape_and_cache. */
int b200_reshape_and_cache(const b200_cache_write_params_t* params, void* stream);

/* Replaces torch.ops._C_cache_ops.reshape_and_cache_flash. */
int b200_reshape_and_cache_flash(const b200_cache_write_params_t* params, void* stream);

/* Replaces torch.ops._C_cache_ops.convert_fp8 (csrc/torch_bindings.cpp:423-426;
 * csrc/cache_kernels.cu:543-614).  Element-wise over num_elements
 * (= size(0) * stride(0) of the op's tensors).  Exactly one of
 * dst_dtype / src_dtype is B200_DTYPE_U8: src U8 = dequantise
 * (float(e4m3) * scale -> dst dtype), dst U8 = quantise
 * (e4m3_satfinite_rn(float(x) / scale)).  scale is a HOST float. */
int b200_convert_fp8(void* dst, const void* src, int64_t num_elements, float scale,
                     int32_t dst_dtype, int32_t src_dtype, int32_t kv_dtype, void* stream);

/*
 * Replaces torch.ops._C.merge_attn_states (csrc/torch_bindings.cpp:74-82;
 * csrc/attention/merge_attn_states.cu:14-88, :129-168): section 2.2 of
 * arXiv 2501.01005 — combine two partial attention results over disjoint key
 * sets:  m = max(p_lse, s_lse), w_p = exp(p_lse - m), w_s = exp(s_lse - m),
 * out = (prefix*w_p + suffix*w_s) / (w_p + w_s), out_lse = log(w_p + w_s) + m,
 * with +inf lse treated as -inf (:45-46).  fp32 arithmetic.
 *   output / prefix_output / suffix_output [num_tokens, num_heads, head_size]
 *       dtype (f32/f16/bf16), heads*dims contiguous, token strides given in
 *       elements (the reference assumes num_heads*head_size, :36-37)
 *   prefix_lse / suffix_lse / output_lse [num_heads, num_tokens] float32,
 *       output_lse nullable
 *   head_size must be a multiple of 16 / sizeof(dtype) (:138-139) and the
 *   tensors 16-byte aligned.
 */
typedef struct b200_merge_attn_states_params {
  void* output;
  float* output_lse; /* nullable */
  const void* prefix_output;
  const float* prefix_lse;
  const void* suffix_output;
  const float* suffix_lse;
  int64_t output_stride; /* token strides, elements */
  int64_t prefix_stride;
  int64_t suffix_stride;
  int32_t num_tokens;
  int32_t num_heads;
  int32_t head_size;
  int32_t dtype; /* b200_dtype_t */
} b200_merge_attn_states_params_t;

int b200_merge_attn_states(const b200_merge_attn_states_params_t* params, void* stream);

/*
 * Replaces torch.ops._C_cache_ops.copy_blocks (csrc/torch_bindings.cpp:384-387;
 * csrc/cache_kernels.cu:64-92, :113-163): for every layer and every
 * (src, dst) pair copy block src onto block dst inside the same key cache and
 * the same value cache.
 *   key_caches / value_caches : HOST arrays of num_layers DEVICE pointers (the
 *       reference stages them through a synchronous H2D copy, :139-145; here
 *       they travel as kernel arguments — no copy, no sync, graph-capturable)
 *   block_mapping : DEVICE int64 [num_pairs, 2]
 *   block_bytes   : bytes of one block (= numel of cache[0] * element size, :148)
 */
int b200_copy_blocks(void* const* key_caches, void* const* value_caches, int32_t num_layers,
                     const int64_t* block_mapping, int32_t num_pairs, int64_t block_bytes, void* stream);

/* Replaces torch.ops._C_cache_ops.copy_blocks_mla (csrc/torch_bindings.cpp:389-391; csrc/cache_kernels.cu:94-110,
 * :165-201): as b200_copy_blocks for one joint cache per layer; block_bytes = stride(0) * element size (:186-189). */
int b200_copy_blocks_mla(void* const* kv_caches, int32_t num_layers, const int64_t* block_mapping, int32_t num_pairs,
                         int64_t block_bytes, void* stream);

/*
 * Replaces torch.ops._C_cache_ops.concat_and_cache_mla (csrc/torch_bindings.cpp:415-420;
 * csrc/cache_kernels.cu:346-389, :490-539): cache entry of token t at slot_mapping[t] =
 * [kv_c[t] (kv_lora_rank) | k_pe[t] (pe_dim)], optionally e4m3-quantised with one scale; slot < 0 skipped;
 * num_tokens = slot_mapping.size(0).
 *   kv_c [.., kv_lora_rank] row stride kv_c_stride; k_pe [.., pe_dim] row stride k_pe_stride (elements)
 *   kv_cache [num_blocks, block_size, kv_lora_rank + pe_dim] addressed by block_stride / entry_stride
 */
typedef struct b200_concat_mla_params {
  const void* kv_c;
  const void* k_pe;
  void* kv_cache;
  const int64_t* slot_mapping;
  const float* scale; /* device; read only when kv_dtype != auto */
  int64_t kv_c_stride;
  int64_t k_pe_stride;
  int64_t block_stride;
  int64_t entry_stride;
  int32_t num_tokens;
  int32_t kv_lora_rank;
  int32_t pe_dim;
  int32_t block_size;
  int32_t dtype;    /* b200_dtype_t of kv_c / k_pe */
  int32_t kv_dtype; /* b200_kv_dtype_t */
} b200_concat_mla_params_t;

int b200_concat_and_cache_mla(const b200_concat_mla_params_t* params, void* stream);

/* memcpy direction of b200_swap_blocks (csrc/cache_kernels.cu:22-34). */
typedef enum b200_copy_kind {
  B200_COPY_D2D = 0,
  B200_COPY_D2H = 1,
  B200_COPY_H2D = 2
} b200_copy_kind_t;

/*
 * Replaces torch.ops._C_cache_ops.swap_blocks (csrc/torch_bindings.cpp:380-381;
 * csrc/cache_kernels.cu:18-60): dst block mapping[i][1] <- src block mapping[i][0],
 * block_bytes each (= element_size * stride(0), :46), enqueued on `stream`.
 *   block_mapping : HOST int64 [num_pairs, 2] (:40)
 * Runs of pairs that are consecutive in both src and dst go out as ONE copy
 * (the reference issues one cudaMemcpyAsync per block, :52-59).
 */
int b200_swap_blocks(void* dst, const void* src, const int64_t* block_mapping, int64_t num_pairs,
                     int64_t block_bytes, int32_t kind, void* stream);

/*
 * Arguments of gather_and_maybe_dequant_cache / cp_gather_cache: copy the
 * cache entries of each sequence, in token order, into a packed destination.
 *   reference ops : csrc/torch_bindings.cpp:431-442
 *   reference C++ : csrc/cache_kernels.cu:618-773 (gather_and_maybe_dequant_cache),
 *                   :775-839, :1075-1134 (cp_gather_cache)
 *   src_cache   [num_blocks, block_size, entry_size] addressed by cache_block_stride /
 *               cache_entry_stride (elements), entry contiguous
 *   dst         [total_tokens, entry_size], row stride dst_entry_stride
 *   block_table [batch, ...] int32, row stride block_table_stride
 *   cu_seq_lens [batch + 1] int32; sequence b fills dst rows [cu[b], cu[b+1])
 *   seq_starts  [batch] int32 or NULL: first token of each sequence inside its
 *               block table.  token_granular_starts == 0: rounded DOWN to a block
 *               boundary (gather_and_maybe_dequant_cache, :660-662);
 *               != 0: used as is (cp_gather_cache, :806-811)
 *   scale       device float, read only when kv_dtype != auto
 *   src_dtype / dst_dtype: b200_dtype_t; auto: equal (any 1/2/4-byte element is
 *               moved as bytes: pass U8/F16/F32 by width); fp8: src U8, dst f32/f16/bf16
 *               = float(e4m3) * scale.
 */
typedef struct b200_gather_cache_params {
  const void* src_cache;
  void* dst;
  const int32_t* block_table;
  const int32_t* cu_seq_lens;
  const int32_t* seq_starts; /* nullable */
  const float* scale;        /* nullable when kv_dtype == auto */
  int64_t block_table_stride;
  int64_t cache_block_stride;
  int64_t cache_entry_stride;
  int64_t dst_entry_stride;
  int32_t batch_size;
  int32_t block_size;
  int32_t entry_size;
  int32_t src_dtype;
  int32_t dst_dtype;
  int32_t kv_dtype;
  int32_t token_granular_starts;
} b200_gather_cache_params_t;

/* Replaces torch.ops._C_cache_ops.gather_and_maybe_dequant_cache and
 * torch.ops._C_cache_ops.cp_gather_cache (see the struct). */
int b200_gather_cache(const b200_gather_cache_params_t* params, void* stream);

/*
 * Arguments of rotary_embedding, optionally fused with the flash-layout cache write that follows it in the
 * serving step.
 *   reference op  : csrc/torch_bindings.cpp:216-219
 *   reference C++ : csrc/pos_encoding_kernels.cu:9-96 (kernels), :126-209 (host)
 *   positions     [num_tokens] int64
 *   query         [num_tokens, num_heads, head_size] dtype, rotated IN PLACE; token stride query_stride,
 *                 head stride head_stride (the reference uses the query's head stride for the key too, :187-188)
 *   key           nullable, [num_tokens, num_kv_heads, head_size], rotated in place; token stride key_stride
 *   cos_sin_cache [max_position, rot_dim] dtype: cos in [0, rot_dim/2), sin in [rot_dim/2, rot_dim)
 *   is_neox       1: pairs (i, i + rot_dim/2) (GPT-NeoX); 0: pairs (2i, 2i+1) (GPT-J)   (:15-29)
 * Arithmetic as the reference kernel writes it (:31-34): x*cos - y*sin and y*cos + x*sin with every product and
 * the sum rounded to dtype — bit-identical to the PyTorch-native implementation the reference tests against.
 * Fused cache write (b200_rotary_embedding_and_cache_flash only): after the rotation the kernel scatters the
 * rotated key and the value of the first num_cache_tokens tokens into the flash-layout caches exactly as
 * b200_reshape_and_cache_flash would (same fields as b200_cache_write_params_t) — one launch instead of two.
 */
typedef struct b200_rotary_params {
  const int64_t* positions;
  void* query;
  void* key; /* nullable */
  const void* cos_sin_cache;
  int64_t query_stride;
  int64_t key_stride;
  int64_t head_stride;
  int32_t num_tokens;
  int32_t num_heads;
  int32_t num_kv_heads;
  int32_t head_size;
  int32_t rot_dim;
  int32_t is_neox;
  int32_t dtype; /* b200_dtype_t */
  /* fused cache write */
  int32_t num_cache_tokens; /* slot_mapping.size(0) <= num_tokens */
  const void* value;
  void* key_cache;
  void* value_cache;
  const int64_t* slot_mapping;
  const float* k_scale;
  const float* v_scale;
  int64_t value_stride;
  int64_t block_stride;
  int64_t page_stride;
  int64_t cache_head_stride;
  int32_t block_size;
  int32_t kv_dtype; /* b200_kv_dtype_t */
} b200_rotary_params_t;

/* Replaces torch.ops._C.rotary_embedding (the cache-write fields are ignored). */
int b200_rotary_embedding(const b200_rotary_params_t* params, void* stream);

/* rotary_embedding followed by reshape_and_cache_flash in one launch (no reference op: the reference runs the
 * two ops back to back, vllm_metax/v1/attention/backends/flash_attn.py:678-687 after the model's rotary layer). */
int b200_rotary_embedding_and_cache_flash(const b200_rotary_params_t* params, void* stream);

/*
 * Decode context parallel (DCP): exchange + combine of the per-rank partial
 * attention results in ONE launch over peer-mapped memory.  Replaces the
 * collective pair + correction the reference runs after its DCP attention call
 * (vllm_metax/v1/attention/backends/flash_attn.py:836-841, upstream
 * `cp_lse_ag_out_rs`: all-gather of the lse, reduce-scatter of the corrected
 * outputs), with the combination rule of merge_attn_states
 * (csrc/attention/merge_attn_states.cu:14-88) applied to world partials at once.
 *
 * Every rank r has written, with an attention call that precedes this one on
 * `stream`, its partial over its own token shard for ALL heads of the group:
 *   peer_out[r]   [num_seqs, world*heads_local, head_size] dtype, contiguous
 *   peer_lse[r]   [world*heads_local, num_seqs] float32
 * (pointers valid in THIS process: the local buffer for r == rank, an IPC
 * mapping otherwise — b200_peer_alloc / b200_peer_open).  peer_flags[r] is rank
 * r's array of `world` uint32 arrival flags, zero-initialised, writable through
 * the mapping.  The launch publishes `epoch` to flag[rank] of every peer, waits
 * for flag[r] >= epoch of every r, pulls the slices of heads
 * [rank*heads_local, (rank+1)*heads_local) and writes
 *   out [num_seqs, heads_local, head_size] dtype,
 *   out_lse [heads_local, num_seqs] float32 (nullable).
 * `epoch` starts at 1 and grows by one per step on every rank; the partial
 * buffers must alternate between two allocations by epoch parity (a peer may
 * be one step ahead).  A peer that does not arrive within
 * B200_DCP_WAIT_SECONDS (default 120 s) traps the kernel.
 * world <= B200_DCP_MAX_WORLD.
 *
 * Step number in device memory (ABI v3, CUDA-graph capture): with
 * step_counter != NULL the launch takes its step from *step_counter — the word
 * the b200_dcp_gather_query call of the same step has just advanced — and
 * `epoch` is ignored, so nothing in the launch changes from step to step and
 * the gather / attention / fold triple can be captured once and replayed.
 * With a gather in every step ONE set of buffers is enough (no parity): a
 * peer's gather of step k+1 returns only after every rank has published its
 * gather flag k+1, which every rank does after its fold of step k has
 * finished reading.  With step_counter == NULL (host-side epoch) the call
 * refuses to be captured: a replay would wait for a step long passed.
 */
#define B200_DCP_MAX_WORLD 16
typedef struct b200_dcp_fold_params {
  const void* const* peer_out;  /* [world] host array of device pointers */
  const float* const* peer_lse; /* [world] */
  uint32_t* const* peer_flags;  /* [world] */
  void* out;
  float* out_lse; /* nullable */
  int32_t rank;
  int32_t world;
  int32_t num_seqs;
  int32_t heads_local;
  int32_t head_size;
  int32_t dtype; /* b200_dtype_t of the partial and final outputs */
  uint32_t epoch;
  /* ABI v3, nullable: the step counter of the paired b200_dcp_gather_query (local device memory), read only */
  const uint32_t* step_counter;
} b200_dcp_fold_params_t;
int b200_dcp_exchange_fold(const b200_dcp_fold_params_t* params, void* stream);

/*
 * Query all-gather of the DCP step over peer memory.  Replaces
 * `get_dcp_group().all_gather(query, dim=1)`
 * (vllm_metax/v1/attention/backends/flash_attn.py:812).  Stores this rank's
 * query [num_seqs, heads_local, head_size] (q_stride elements between
 * sequences, heads contiguous) into heads [rank*heads_local, ...) of every
 * rank's gathered query peer_q[r] [num_seqs, world*heads_local, head_size],
 * publishes `epoch` to flag[rank] of every peer and returns (on the stream)
 * once every rank has published: the attention call that follows reads a whole
 * gathered query.  peer_flags are arrays of `world` uint32 distinct from the
 * fold's; cta_counter is one zero-initialised uint32 of local device memory.
 * The gathered-query buffers alternate by epoch parity like the partials.
 * ABI v3: with step_counter != NULL (one zero-initialised uint32 of local
 * device memory) the step is *step_counter + 1, which the launch stores back
 * before it completes; `epoch` is ignored and the call can be captured into a
 * CUDA graph (see b200_dcp_exchange_fold).
 */
typedef struct b200_dcp_gather_params {
  void* const* peer_q;         /* [world] host array of device pointers */
  uint32_t* const* peer_flags; /* [world] */
  const void* query;
  uint32_t* cta_counter;
  int64_t q_stride;
  int32_t rank;
  int32_t world;
  int32_t num_seqs;
  int32_t heads_local;
  int32_t head_size;
  int32_t dtype;
  uint32_t epoch;
  uint32_t* step_counter; /* ABI v3, nullable */
} b200_dcp_gather_params_t;
int b200_dcp_gather_query(const b200_dcp_gather_params_t* params, void* stream);

/* Peer-mappable device memory for the DCP buffers: one cudaMalloc (zero-filled)
 * and its 64-byte CUDA IPC handle; b200_peer_open maps a handle received from
 * another process of the same node (peer access enabled lazily) and returns a
 * pointer valid in the calling process.  Synchronous host calls. */
int b200_peer_alloc(uint64_t bytes, void** ptr, void* handle64);
int b200_peer_open(const void* handle64, void** ptr);
int b200_peer_close(void* ptr);
int b200_peer_free(void* ptr);

/*
 * MLA decode attention.  Replaces torch.ops._C.cutlass_mla_decode (csrc/torch_bindings.cpp:287-290; entry
 * csrc/attention/mla/cutlass_mla_entry.cu:28-38; checks and launch csrc/attention/mla/cutlass_mla_kernels.cu:154-225)
 * and the Triton split-K MLA decode (vllm_metax/attention/ops/triton_decode_attention.py:673-721) for the MLA shape:
 *   out[s, h, :512] = softmax_t(scale * [q_nope[s,h] | q_pe[s,h]] . kv[t, :576]) . kv[t, :512],  t < seq_lens[s]
 * where row t of sequence s is kv_cache[page_table[s, t / page_size], t % page_size, :] (one shared latent head).
 *   q_nope [num_seqs, num_heads, 512], q_pe [num_seqs, num_heads, 64], out [num_seqs, num_heads, 512]: element type
 *     `dtype` (bf16 / f16), innermost contiguous, seq / head strides in elements (16-byte aligned rows)
 *   kv_cache [num_blocks, page_size, 576] same element type; block_stride / entry_stride in elements
 *   kv_dtype FP8_E4M3: q_nope, q_pe and kv_cache hold e4m3 bytes instead (out stays `dtype`), as the reference op accepts
 *   seq_lens int32[num_seqs]; page_table int32[num_seqs, max_pages] (row stride page_table_stride)
 *   lse (optional) float32 [num_seqs, num_heads] (row stride lse_stride): natural-log sum of exp of the scaled logits
 *   num_heads <= 128 (the reference requires exactly 128 and pads; any count works here), page_size a power of two >= 16
 *   workspace: b200_mla_decode_workspace_bytes(num_seqs, num_heads, max_pages * page_size, num_kv_splits) bytes, used when the
 *     launch splits sequences along tokens (fp32 partial outputs + LSEs, merged by a second kernel); may be NULL when
 *     num_kv_splits == 1.  num_kv_splits: 0 = choose, 1 = never split, n = at most n splits per sequence.
 * A sequence of length 0 writes zeros (lse = -inf).  All pointers are device pointers; nothing is allocated.
 */
typedef struct b200_mla_decode_params {
  void* out;
  float* lse;
  const void* q_nope;
  const void* q_pe;
  const void* kv_cache;
  const int32_t* seq_lens;
  const int32_t* page_table;
  void* workspace;
  uint64_t workspace_bytes;
  int64_t out_seq_stride, out_head_stride;
  int64_t q_nope_seq_stride, q_nope_head_stride;
  int64_t q_pe_seq_stride, q_pe_head_stride;
  int64_t block_stride, entry_stride;
  int64_t page_table_stride;
  int64_t lse_stride;
  int64_t num_blocks;
  float scale;
  int32_t num_seqs;
  int32_t num_heads;
  int32_t page_size;
  int32_t max_pages;
  int32_t dtype; /* b200_dtype_t of out: BF16 or F16 (and of q / cache when kv_dtype is auto) */
  int32_t num_kv_splits;
  float* debug_logits; /* NULL, or [128 x 64] fp32: raw S^T of the first tile of CTA 0 (tools/ only) */
  int32_t kv_dtype;    /* b200_kv_dtype_t: AUTO = q / cache have the out dtype; FP8_E4M3 = q AND cache are e4m3 bytes
                        * (the reference op's float8_e4m3fn form, descale 1; strides then count bytes) */
  int32_t reserved;
} b200_mla_decode_params_t;

int b200_mla_decode(const b200_mla_decode_params_t* params, void* stream);
uint64_t b200_mla_decode_workspace_bytes(int32_t num_seqs, int32_t num_heads, int64_t max_seq_len, int32_t num_kv_splits);

/* "auto" -> B200_KV_AUTO, "fp8" | "fp8_e4m3" -> B200_KV_FP8_E4M3, anything
 * else (incl. "fp8_e5m2") -> -1. */
int b200_kv_dtype_from_string(const char* kv_cache_dtype);

/* Message of the last error on this thread ("" if none). */
const char* b200_last_error(void);

/* ABI version of the loaded library (B200_PAGED_ATTN_ABI_VERSION). */
int b200_abi_version(void);

/* Test/bench hooks: force a kernel family (b200_attn_variant_t; AUTO
 * restores dispatch) and read back which one the last attention call on this
 * thread used.  Forcing STREAM_MMA on an ineligible shape fails with
 * B200_ERR_UNSUPPORTED instead of falling back. */
int b200_set_attention_variant(int variant);
int b200_last_attention_variant(void);
/* Experiment switch (no reference counterpart): min_chunk_tiles > 0 lets the flat schedule of the streaming kernel hand
 * out the last ~15 % of the tiles dynamically in chunks of at least that many 32-token tiles; 0 = static split
 * (default); < 0 = follow the B200_STREAM_DYN environment variable.  Results are identical either way. */
int b200_set_stream_dynamic(int min_chunk_tiles);

/* Number of kernel launches the library has enqueued from this process since
 * load (bench.py reports it as gpu_launches). */
uint64_t b200_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* B200_PAGED_ATTN_H_ */
