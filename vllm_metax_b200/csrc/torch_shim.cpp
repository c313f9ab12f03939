// Thin torch-dispatcher shim: registers the reference's operator schemas and
// forwards to the C ABI in include/b200_paged_attn.h.
//
// The schema strings are byte-identical to the reference's registrations
// (csrc/torch_bindings.cpp:45-55, :58-69, :394-400, :403-411, :423-426), under
// the same library names `_C` and `_C_cache_ops`, with CUDA-key
// implementations only (the reference registers torch::kCUDA only: :55, :69,
// :400, :410).  There is deliberately no CPU implementation: on a tensor that
// is not on a CUDA device the dispatcher raises, exactly as with the reference.
//
// If another library already defined the schemas in this process (upstream
// vllm._C, whose strings are identical), only the CUDA kernels are overridden.
#include <ATen/cuda/CUDAContext.h>
#include <Python.h>
#include <c10/cuda/CUDAGuard.h>
#include <torch/library.h>
#include <torch/types.h>

#include <string.h>

#include <algorithm>
#include <memory>
#include <mutex>
#include <optional>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/b200_paged_attn.h"

namespace {

int to_dtype(at::ScalarType t) {
  switch (t) {
    case at::kFloat: return B200_DTYPE_F32;
    case at::kHalf: return B200_DTYPE_F16;
    case at::kBFloat16: return B200_DTYPE_BF16;
    case at::kByte: return B200_DTYPE_U8;
    case at::kFloat8_e4m3fn: return B200_DTYPE_U8;
    default: return -1;
  }
}

int kv_dtype_or_throw(const std::string& s) {
  const int kv = b200_kv_dtype_from_string(s.c_str());
  // csrc/quantization/fp8/metax/quant_utils.cuh:37-41
  TORCH_CHECK(kv >= 0, "Unsupported data type of kv cache: ", s);
  return kv;
}

void check_status(int rc, const char* op) {
  TORCH_CHECK(rc == B200_OK, op, ": ", b200_last_error());
}

bool is_cache_byte_type(at::ScalarType t) { return t == at::kByte || t == at::kFloat8_e4m3fn; }

// v1 split-KV scratch: one buffer per (device, stream).  The op has no workspace argument
// (csrc/torch_bindings.cpp:45-55) and the reference's ops allocate nothing (SURVEY.md §8b "Ownership"), so the
// allocation lives OUTSIDE the attention ops: `_C_b200::reserve_workspace` (called by import_kernels()-time warm-up /
// the plugin's profile run) sizes it; the attention ops only look it up.  For the streaming kernels the size is a
// constant (kFlatWorkspaceBytes, ~27 MB); only the generic kernel's split-KV partials grow with the shape.  A
// call that finds no (or too small a) buffer allocates it lazily — unless the stream is being captured into a CUDA
// graph, where an allocation would be illegal: that raises and names reserve_workspace.  Work on one stream is
// ordered, so reuse across calls is safe.
struct WorkspaceCache {
  std::mutex mu;
  std::unordered_map<uint64_t, at::Tensor> map;
};
WorkspaceCache& workspace_cache() {
  static WorkspaceCache c;
  return c;
}
uint64_t workspace_key(int device, cudaStream_t stream) {
  return (static_cast<uint64_t>(device) << 56) ^ reinterpret_cast<uint64_t>(stream);
}
bool stream_is_capturing(cudaStream_t stream) {
  cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(stream, &st) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return st != cudaStreamCaptureStatusNone;
}
at::Tensor& grow_workspace(uint64_t bytes, const at::TensorOptions& opts, int device, cudaStream_t stream, const char* op) {
  WorkspaceCache& c = workspace_cache();
  at::Tensor& t = c.map[workspace_key(device, stream)];
  if (!t.defined() || static_cast<uint64_t>(t.numel()) < bytes) {
    TORCH_CHECK(!stream_is_capturing(stream), op, ": the split-KV workspace of this stream (", t.defined() ? t.numel() : 0,
                " bytes) is smaller than the ", bytes, " bytes this shape needs and the stream is being captured into a CUDA "
                "graph; call torch.ops._C_b200.reserve_workspace(query, max_seq_len) once before capture");
    const uint64_t want = bytes + bytes / 4;
    // zero-filled once: the arrival counters inside must start at zero; every launch leaves them at zero
    t = at::zeros({static_cast<int64_t>(want)}, opts.dtype(at::kByte));
  }
  return t;
}
void* v1_workspace(uint64_t bytes, const at::Tensor& like, cudaStream_t stream, const char* op = "paged_attention") {
  if (bytes == 0) return nullptr;
  std::lock_guard<std::mutex> lock(workspace_cache().mu);
  return grow_workspace(bytes, like.options(), like.get_device(), stream, op).data_ptr();
}

// _C_b200::reserve_workspace(Tensor query_like, int max_seq_len, bool reset) -> ()
// Sizes (and with reset = true re-zeroes) the current stream's workspace for decode batches shaped like `query_like`
// [num_seqs, num_heads, head_size] with contexts up to max_seq_len.  Call it at warm-up, outside graph capture.
void reserve_workspace(const at::Tensor& query_like, int64_t max_seq_len, bool reset) {
  TORCH_CHECK(query_like.is_cuda() && query_like.dim() == 3, "reserve_workspace: query_like must be a CUDA tensor [num_seqs, num_heads, head_size]");
  b200_paged_attn_params_t p;
  memset(&p, 0, sizeof(p));
  p.num_seqs = static_cast<int32_t>(query_like.size(0));
  p.num_heads = static_cast<int32_t>(query_like.size(1));
  p.num_kv_heads = p.num_heads;
  p.head_size = static_cast<int32_t>(query_like.size(2));
  p.max_seq_len = static_cast<int32_t>(max_seq_len);
  const at::cuda::OptionalCUDAGuard guard(device_of(query_like));
  const cudaStream_t stream = at::cuda::getCurrentCUDAStream();
  const uint64_t bytes = b200_paged_attention_workspace_bytes(&p);
  if (bytes == 0) return;
  std::lock_guard<std::mutex> lock(workspace_cache().mu);
  at::Tensor& t = grow_workspace(bytes, query_like.options(), query_like.get_device(), stream, "reserve_workspace");
  if (reset) check_status(b200_paged_attention_workspace_reset(t.data_ptr(), static_cast<uint64_t>(t.numel()), stream), "reserve_workspace");
}

void fill_attn_params(b200_paged_attn_params_t& p, at::Tensor& out, const at::Tensor& query,
                      const at::Tensor& key_cache, const at::Tensor& value_cache, int64_t num_kv_heads,
                      double scale, const at::Tensor& block_tables, const at::Tensor& seq_lens,
                      int64_t block_size, int64_t max_seq_len, const std::optional<at::Tensor>& alibi_slopes,
                      const std::string& kv_cache_dtype, const at::Tensor& k_scale, const at::Tensor& v_scale,
                      int64_t tp_rank, int64_t bs_local, int64_t bs_vert, int64_t bs_block, int64_t bs_slide) {
  TORCH_CHECK(query.dim() == 3 && out.dim() == 3, "query/out must be [num_seqs, num_heads, head_size]");
  TORCH_CHECK(out.scalar_type() == query.scalar_type(), "out and query dtypes differ");
  TORCH_CHECK(out.is_contiguous(), "out must be contiguous");
  TORCH_CHECK(query.stride(2) == 1 && query.stride(1) == query.size(2), "query heads must be contiguous");
  TORCH_CHECK(key_cache.dim() == 5 && value_cache.dim() == 4,
              "key_cache must be [nb, nkv, hs/x, bs, x] and value_cache [nb, nkv, hs, bs]");
  TORCH_CHECK(block_tables.scalar_type() == at::kInt && block_tables.dim() == 2 && block_tables.stride(1) == 1,
              "block_tables must be a 2-D int32 tensor with contiguous rows");
  TORCH_CHECK(seq_lens.scalar_type() == at::kInt && seq_lens.is_contiguous(), "seq_lens must be contiguous int32");
  TORCH_CHECK(k_scale.scalar_type() == at::kFloat && v_scale.scalar_type() == at::kFloat,
              "k_scale / v_scale must be float32 tensors");
  TORCH_CHECK(query.is_cuda() && key_cache.is_cuda() && value_cache.is_cuda() && block_tables.is_cuda() &&
                  seq_lens.is_cuda() && k_scale.is_cuda() && v_scale.is_cuda(),
              "all tensors must be on the CUDA device");
  const int kv = kv_dtype_or_throw(kv_cache_dtype);
  if (kv == B200_KV_AUTO) {
    TORCH_CHECK(key_cache.scalar_type() == query.scalar_type() && value_cache.scalar_type() == query.scalar_type(),
                "kv_cache_dtype 'auto' needs caches of the query dtype");
  } else {
    TORCH_CHECK(is_cache_byte_type(key_cache.scalar_type()) && is_cache_byte_type(value_cache.scalar_type()),
                "fp8 kv cache must be stored as uint8");
  }
  TORCH_CHECK(value_cache.stride(0) == key_cache.stride(0) && value_cache.stride(1) == key_cache.stride(1),
              "key_cache and value_cache must share block/head strides");
  const int dt = to_dtype(query.scalar_type());
  TORCH_CHECK(dt == B200_DTYPE_F32 || dt == B200_DTYPE_F16 || dt == B200_DTYPE_BF16, "Unsupported data type: ",
              query.scalar_type());
  memset(&p, 0, sizeof(p));
  p.out = out.data_ptr();
  p.query = query.data_ptr();
  p.key_cache = key_cache.data_ptr();
  p.value_cache = value_cache.data_ptr();
  p.block_tables = block_tables.data_ptr<int32_t>();
  p.seq_lens = seq_lens.data_ptr<int32_t>();
  if (alibi_slopes.has_value() && alibi_slopes->defined()) {
    TORCH_CHECK(alibi_slopes->scalar_type() == at::kFloat && alibi_slopes->is_contiguous() &&
                    alibi_slopes->numel() == query.size(1) && alibi_slopes->is_cuda(),
                "alibi_slopes must be a contiguous float32 CUDA tensor of num_heads entries");
    p.alibi_slopes = alibi_slopes->data_ptr<float>();
  }
  p.k_scale = k_scale.data_ptr<float>();
  p.v_scale = v_scale.data_ptr<float>();
  p.q_stride = query.stride(0);
  p.kv_block_stride = key_cache.stride(0);
  p.kv_head_stride = key_cache.stride(1);
  p.num_seqs = static_cast<int32_t>(query.size(0));
  p.num_heads = static_cast<int32_t>(query.size(1));
  p.num_kv_heads = static_cast<int32_t>(num_kv_heads);
  p.head_size = static_cast<int32_t>(query.size(2));
  p.block_size = static_cast<int32_t>(block_size);
  p.max_seq_len = static_cast<int32_t>(max_seq_len);
  p.max_num_blocks_per_seq = static_cast<int32_t>(block_tables.size(1));
  p.dtype = dt;
  p.kv_dtype = kv;
  p.scale = static_cast<float>(scale);
  p.tp_rank = static_cast<int32_t>(tp_rank);
  p.blocksparse_local_blocks = static_cast<int32_t>(bs_local);
  p.blocksparse_vert_stride = static_cast<int32_t>(bs_vert);
  p.blocksparse_block_size = static_cast<int32_t>(bs_block);
  p.blocksparse_head_sliding_step = static_cast<int32_t>(bs_slide);
  TORCH_CHECK(value_cache.size(3) == block_size && value_cache.size(2) == query.size(2) &&
                  value_cache.size(1) == num_kv_heads,
              "value_cache shape does not match num_kv_heads / head_size / block_size");
  TORCH_CHECK(block_tables.size(0) >= query.size(0) && seq_lens.numel() >= query.size(0),
              "block_tables / seq_lens have fewer rows than query");
}

// csrc/ops.h:32-41
void paged_attention_v1(at::Tensor& out, at::Tensor& query, at::Tensor& key_cache, at::Tensor& value_cache,
                        int64_t num_kv_heads, double scale, at::Tensor& block_tables, at::Tensor& seq_lens,
                        int64_t block_size, int64_t max_seq_len, const std::optional<at::Tensor>& alibi_slopes,
                        const std::string& kv_cache_dtype, at::Tensor& k_scale, at::Tensor& v_scale,
                        const int64_t tp_rank, const int64_t blocksparse_local_blocks,
                        const int64_t blocksparse_vert_stride, const int64_t blocksparse_block_size,
                        const int64_t blocksparse_head_sliding_step) {
  b200_paged_attn_params_t p;
  fill_attn_params(p, out, query, key_cache, value_cache, num_kv_heads, scale, block_tables, seq_lens, block_size,
                   max_seq_len, alibi_slopes, kv_cache_dtype, k_scale, v_scale, tp_rank, blocksparse_local_blocks,
                   blocksparse_vert_stride, blocksparse_block_size, blocksparse_head_sliding_step);
  const at::cuda::OptionalCUDAGuard guard(device_of(query));
  const cudaStream_t stream = at::cuda::getCurrentCUDAStream();
  p.workspace_bytes = b200_paged_attention_workspace_bytes(&p);
  p.workspace = v1_workspace(p.workspace_bytes, query, stream);
  check_status(b200_paged_attention_v1(&p, stream), "paged_attention_v1");
}

// Not a reference op: paged_attention_v1 that also returns the log-sum-exp of the scaled logits,
// lse [num_heads, num_seqs] float32 — the second operand merge_attn_states needs (csrc/attention/merge_attn_states.cu:43-44)
// when a decode result is merged with another partial result (cascade attention, decode context parallel:
// vllm_metax/v1/attention/backends/flash_attn.py:794-872).  Registered under `_C_b200`, outside the reference namespaces.
void paged_attention_v1_lse(at::Tensor& out, at::Tensor& lse, at::Tensor& query, at::Tensor& key_cache,
                            at::Tensor& value_cache, int64_t num_kv_heads, double scale, at::Tensor& block_tables,
                            at::Tensor& seq_lens, int64_t block_size, int64_t max_seq_len,
                            const std::optional<at::Tensor>& alibi_slopes, const std::string& kv_cache_dtype,
                            at::Tensor& k_scale, at::Tensor& v_scale) {
  b200_paged_attn_params_t p;
  fill_attn_params(p, out, query, key_cache, value_cache, num_kv_heads, scale, block_tables, seq_lens, block_size,
                   max_seq_len, alibi_slopes, kv_cache_dtype, k_scale, v_scale, 0, 0, 0, 64, 0);
  TORCH_CHECK(lse.is_cuda() && lse.scalar_type() == at::kFloat && lse.is_contiguous() && lse.dim() == 2 &&
                  lse.size(0) == query.size(1) && lse.size(1) == query.size(0),
              "lse must be a contiguous float32 CUDA tensor [num_heads, num_seqs]");
  p.out_lse = lse.data_ptr<float>();
  const at::cuda::OptionalCUDAGuard guard(device_of(query));
  const cudaStream_t stream = at::cuda::getCurrentCUDAStream();
  p.workspace_bytes = b200_paged_attention_workspace_bytes(&p);
  p.workspace = v1_workspace(p.workspace_bytes, query, stream);
  check_status(b200_paged_attention_v1(&p, stream), "paged_attention_v1_lse");
}

// Not a reference op: decode attention over the flash cache layout [num_blocks, block_size, num_kv_heads, head_size]
// (NHD or HND strides) — what the reference's serving path gets from flash_attn_with_kvcache of the external
// flash-attention library (vllm_metax/v1/attention/backends/flash_attn.py:742-757).  query/out [num_seqs, num_heads, D].
void paged_attention_flash(at::Tensor& out, std::optional<at::Tensor> lse, const at::Tensor& query,
                           const at::Tensor& key_cache, const at::Tensor& value_cache, double scale,
                           const at::Tensor& block_tables, const at::Tensor& seq_lens, int64_t max_seq_len,
                           const std::optional<at::Tensor>& alibi_slopes) {
  TORCH_CHECK(query.dim() == 3 && out.dim() == 3 && out.sizes() == query.sizes(),
              "query/out must be [num_seqs, num_heads, head_size]");
  TORCH_CHECK(out.scalar_type() == query.scalar_type() && out.is_contiguous(), "out must be contiguous, of the query dtype");
  TORCH_CHECK(query.stride(2) == 1 && query.stride(1) == query.size(2), "query heads must be contiguous");
  TORCH_CHECK(key_cache.dim() == 4 && value_cache.dim() == 4 && key_cache.sizes() == value_cache.sizes(),
              "key_cache / value_cache must be [num_blocks, block_size, num_kv_heads, head_size]");
  TORCH_CHECK(key_cache.strides() == value_cache.strides() && key_cache.stride(3) == 1,
              "key_cache and value_cache must share strides with contiguous head dims");
  TORCH_CHECK(key_cache.scalar_type() == query.scalar_type() && value_cache.scalar_type() == query.scalar_type(),
              "flash-layout decode needs caches of the query dtype");
  TORCH_CHECK(key_cache.size(3) == query.size(2), "head_size of the cache and the query differ");
  TORCH_CHECK(block_tables.scalar_type() == at::kInt && block_tables.dim() == 2 && block_tables.stride(1) == 1,
              "block_tables must be a 2-D int32 tensor with contiguous rows");
  TORCH_CHECK(seq_lens.scalar_type() == at::kInt && seq_lens.is_contiguous(), "seq_lens must be contiguous int32");
  TORCH_CHECK(query.is_cuda() && key_cache.is_cuda() && value_cache.is_cuda() && block_tables.is_cuda() && seq_lens.is_cuda(),
              "all tensors must be on the CUDA device");
  TORCH_CHECK(block_tables.size(0) >= query.size(0) && seq_lens.numel() >= query.size(0),
              "block_tables / seq_lens have fewer rows than query");
  const int dt = to_dtype(query.scalar_type());
  TORCH_CHECK(dt == B200_DTYPE_BF16, "paged_attention_flash: query dtype is not bf16");
  b200_paged_attn_params_t p;
  memset(&p, 0, sizeof(p));
  p.out = out.data_ptr();
  p.query = query.data_ptr();
  p.key_cache = key_cache.data_ptr();
  p.value_cache = value_cache.data_ptr();
  p.block_tables = block_tables.data_ptr<int32_t>();
  p.seq_lens = seq_lens.data_ptr<int32_t>();
  if (alibi_slopes.has_value() && alibi_slopes->defined()) {
    TORCH_CHECK(alibi_slopes->scalar_type() == at::kFloat && alibi_slopes->is_contiguous() &&
                    alibi_slopes->numel() == query.size(1) && alibi_slopes->is_cuda(),
                "alibi_slopes must be a contiguous float32 CUDA tensor of num_heads entries");
    p.alibi_slopes = alibi_slopes->data_ptr<float>();
  }
  if (lse.has_value() && lse->defined()) {
    TORCH_CHECK(lse->is_cuda() && lse->scalar_type() == at::kFloat && lse->is_contiguous() && lse->dim() == 2 &&
                    lse->size(0) == query.size(1) && lse->size(1) == query.size(0),
                "lse must be a contiguous float32 CUDA tensor [num_heads, num_seqs]");
    p.out_lse = lse->data_ptr<float>();
  }
  p.q_stride = query.stride(0);
  p.kv_block_stride = key_cache.stride(0);
  p.kv_page_stride = key_cache.stride(1);
  p.kv_num_blocks = key_cache.size(0);
  p.kv_head_stride = key_cache.stride(2);
  p.num_seqs = static_cast<int32_t>(query.size(0));
  p.num_heads = static_cast<int32_t>(query.size(1));
  p.num_kv_heads = static_cast<int32_t>(key_cache.size(2));
  p.head_size = static_cast<int32_t>(query.size(2));
  p.block_size = static_cast<int32_t>(key_cache.size(1));
  p.max_seq_len = static_cast<int32_t>(max_seq_len);
  p.max_num_blocks_per_seq = static_cast<int32_t>(block_tables.size(1));
  p.dtype = dt;
  p.kv_dtype = B200_KV_AUTO;
  p.scale = static_cast<float>(scale);
  p.blocksparse_block_size = 64;
  const at::cuda::OptionalCUDAGuard guard(device_of(query));
  const cudaStream_t stream = at::cuda::getCurrentCUDAStream();
  p.workspace_bytes = b200_paged_attention_workspace_bytes(&p);
  p.workspace = v1_workspace(p.workspace_bytes, query, stream);
  check_status(b200_paged_attention_flash(&p, stream), "paged_attention_flash");
}

void fill_cache_params(b200_cache_write_params_t& p, const at::Tensor& key, const at::Tensor& value,
                       const at::Tensor& key_cache, const at::Tensor& value_cache, const at::Tensor& slot_mapping,
                       const std::string& kv_cache_dtype, const at::Tensor& k_scale, const at::Tensor& v_scale);

// csrc/ops.h (rotary_embedding) ; csrc/pos_encoding_kernels.cu:126-209
void fill_rotary_params(b200_rotary_params_t& p, const at::Tensor& positions, at::Tensor& query,
                        std::optional<at::Tensor>& key, int64_t head_size, const at::Tensor& cos_sin_cache,
                        bool is_neox) {
  const int64_t num_tokens = positions.numel();
  const int64_t pd = positions.dim();
  TORCH_CHECK(pd == 1 || pd == 2, "positions must have shape [num_tokens] or [batch_size, seq_len]");
  const bool has_key = key.has_value() && key->defined();
  if (pd == 1) {
    TORCH_CHECK(query.size(0) == positions.size(0) && (!has_key || key->size(0) == positions.size(0)),
                "query, key and positions must have the same number of tokens");
  } else {
    TORCH_CHECK(query.size(0) == positions.size(0) && (!has_key || key->size(0) == positions.size(0)) &&
                    query.size(1) == positions.size(1) && (!has_key || key->size(1) == positions.size(1)),
                "query, key and positions must have the same batch_size and seq_len");
  }
  TORCH_CHECK(positions.scalar_type() == at::kLong && positions.is_contiguous(), "positions must be contiguous int64");
  TORCH_CHECK(query.is_cuda() && positions.is_cuda() && cos_sin_cache.is_cuda() && (!has_key || key->is_cuda()),
              "all tensors must be on the CUDA device");
  TORCH_CHECK(cos_sin_cache.scalar_type() == query.scalar_type() && (!has_key || key->scalar_type() == query.scalar_type()),
              "query, key and cos_sin_cache dtypes differ");
  TORCH_CHECK(cos_sin_cache.dim() == 2 && cos_sin_cache.is_contiguous(), "cos_sin_cache must be contiguous [max_position, rot_dim]");
  memset(&p, 0, sizeof(p));
  if (num_tokens == 0) return;
  const int64_t q_hidden = query.numel() / num_tokens;
  const int64_t k_hidden = has_key ? key->numel() / num_tokens : 0;
  TORCH_CHECK(q_hidden % head_size == 0);
  TORCH_CHECK(k_hidden % head_size == 0);
  const int64_t num_heads = q_hidden / head_size;
  const int64_t num_kv_heads = has_key ? k_hidden / head_size : num_heads;
  TORCH_CHECK(num_heads % num_kv_heads == 0);
  const int64_t seq_dim = pd - 1;
  if (pd == 2) {  // [batch, seq, ...]: token t = b * seq_len + s must be one stride apart
    TORCH_CHECK(query.stride(0) == query.stride(1) * query.size(1) &&
                    (!has_key || key->stride(0) == key->stride(1) * key->size(1)),
                "batched query/key must be dense over (batch, seq)");
  }
  const int dt = to_dtype(query.scalar_type());
  TORCH_CHECK(dt == B200_DTYPE_F32 || dt == B200_DTYPE_F16 || dt == B200_DTYPE_BF16, "Unsupported data type: ",
              query.scalar_type());
  p.positions = positions.data_ptr<int64_t>();
  p.query = query.data_ptr();
  p.key = has_key ? key->data_ptr() : nullptr;
  p.cos_sin_cache = cos_sin_cache.data_ptr();
  p.query_stride = query.stride(seq_dim);
  p.key_stride = has_key ? key->stride(seq_dim) : 0;
  p.head_stride = (query.dim() == pd + 2) ? query.stride(-2) : head_size;  // pos_encoding_kernels.cu:187-188
  p.num_tokens = static_cast<int32_t>(num_tokens);
  p.num_heads = static_cast<int32_t>(num_heads);
  p.num_kv_heads = static_cast<int32_t>(num_kv_heads);
  p.head_size = static_cast<int32_t>(head_size);
  p.rot_dim = static_cast<int32_t>(cos_sin_cache.size(1));
  p.is_neox = is_neox ? 1 : 0;
  p.dtype = dt;
}

void rotary_embedding(at::Tensor& positions, at::Tensor& query, std::optional<at::Tensor> key, int64_t head_size,
                      at::Tensor& cos_sin_cache, bool is_neox) {
  b200_rotary_params_t p;
  fill_rotary_params(p, positions, query, key, head_size, cos_sin_cache, is_neox);
  if (positions.numel() == 0) return;
  const at::cuda::OptionalCUDAGuard guard(device_of(query));
  check_status(b200_rotary_embedding(&p, at::cuda::getCurrentCUDAStream()), "rotary_embedding");
}

// Not a reference op: rotary_embedding and the reshape_and_cache_flash that follows it in the serving step, one launch.
void rotary_embedding_and_cache_flash(at::Tensor& positions, at::Tensor& query, at::Tensor& key, at::Tensor& value,
                                      int64_t head_size, at::Tensor& cos_sin_cache, bool is_neox, at::Tensor& key_cache,
                                      at::Tensor& value_cache, at::Tensor& slot_mapping, const std::string& kv_cache_dtype,
                                      at::Tensor& k_scale, at::Tensor& v_scale) {
  TORCH_CHECK(positions.dim() == 1 && key.dim() == 3 && value.dim() == 3,
              "fused op takes positions [num_tokens] and key/value [num_tokens, num_kv_heads, head_size]");
  b200_rotary_params_t p;
  std::optional<at::Tensor> okey = key;
  fill_rotary_params(p, positions, query, okey, head_size, cos_sin_cache, is_neox);
  b200_cache_write_params_t c;
  fill_cache_params(c, key, value, key_cache, value_cache, slot_mapping, kv_cache_dtype, k_scale, v_scale);
  TORCH_CHECK(key_cache.dim() == 4 && value_cache.dim() == 4, "caches must be [nb, bs, nh, hs]");
  TORCH_CHECK(key_cache.stride(0) == value_cache.stride(0) && key_cache.stride(1) == value_cache.stride(1) &&
                  key_cache.stride(2) == value_cache.stride(2) && key_cache.stride(3) == 1 && value_cache.stride(3) == 1,
              "key_cache and value_cache must share strides with contiguous head dims");
  TORCH_CHECK(key.stride(0) == value.stride(0), "key and value token strides differ");
  TORCH_CHECK(key_cache.size(2) == key.size(1) && key_cache.size(3) == key.size(2), "cache shapes do not match key");
  if (positions.numel() == 0) return;
  p.num_cache_tokens = c.num_tokens;
  p.value = c.value;
  p.key_cache = c.key_cache;
  p.value_cache = c.value_cache;
  p.slot_mapping = c.slot_mapping;
  p.k_scale = c.k_scale;
  p.v_scale = c.v_scale;
  p.value_stride = c.value_stride;
  p.block_stride = key_cache.stride(0);
  p.page_stride = key_cache.stride(1);
  p.cache_head_stride = key_cache.stride(2);
  p.block_size = static_cast<int32_t>(key_cache.size(1));
  p.kv_dtype = c.kv_dtype;
  const at::cuda::OptionalCUDAGuard guard(device_of(query));
  check_status(b200_rotary_embedding_and_cache_flash(&p, at::cuda::getCurrentCUDAStream()),
               "rotary_embedding_and_cache_flash");
}

// csrc/ops.h (merge_attn_states) ; csrc/attention/merge_attn_states.cu:129-178
void merge_attn_states(at::Tensor& output, std::optional<at::Tensor> output_lse, const at::Tensor& prefix_output,
                       const at::Tensor& prefix_lse, const at::Tensor& suffix_output, const at::Tensor& suffix_lse) {
  const int dt = to_dtype(output.scalar_type());
  TORCH_CHECK(dt == B200_DTYPE_F32 || dt == B200_DTYPE_F16 || dt == B200_DTYPE_BF16,
              "Unsupported data type of O: ", output.scalar_type());  // merge_attn_states.cu:104
  TORCH_CHECK(output.dim() == 3 && prefix_output.sizes() == output.sizes() && suffix_output.sizes() == output.sizes(),
              "output / prefix_output / suffix_output must share the shape [num_tokens, num_heads, head_size]");
  TORCH_CHECK(prefix_output.scalar_type() == output.scalar_type() && suffix_output.scalar_type() == output.scalar_type(),
              "output / prefix_output / suffix_output dtypes differ");
  const int64_t N = output.size(0), H = output.size(1), D = output.size(2);
  const int64_t pack = 16 / output.element_size();
  TORCH_CHECK(D % pack == 0, "headsize must be multiple of pack_size:", pack);  // :138-139
  TORCH_CHECK(output.stride(-2) == D && output.stride(-1) == 1, "output heads must be contiguous in memory");
  TORCH_CHECK(prefix_output.stride(-2) == D && prefix_output.stride(-1) == 1,
              "prefix_output heads must be contiguous in memory");
  TORCH_CHECK(suffix_output.stride(-2) == D && suffix_output.stride(-1) == 1,
              "suffix_output heads must be contiguous in memory");
  TORCH_CHECK(prefix_lse.scalar_type() == at::kFloat && suffix_lse.scalar_type() == at::kFloat &&
                  prefix_lse.is_contiguous() && suffix_lse.is_contiguous() && prefix_lse.numel() == H * N &&
                  suffix_lse.numel() == H * N,
              "prefix_lse / suffix_lse must be contiguous float32 [num_heads, num_tokens]");
  TORCH_CHECK(output.is_cuda() && prefix_output.is_cuda() && suffix_output.is_cuda() && prefix_lse.is_cuda() &&
                  suffix_lse.is_cuda(),
              "all tensors must be on the CUDA device");
  b200_merge_attn_states_params_t p;
  memset(&p, 0, sizeof(p));
  p.output = output.data_ptr();
  if (output_lse.has_value() && output_lse->defined()) {
    TORCH_CHECK(output_lse->scalar_type() == at::kFloat && output_lse->is_contiguous() && output_lse->numel() == H * N &&
                    output_lse->is_cuda(),
                "output_lse must be contiguous float32 [num_heads, num_tokens]");
    p.output_lse = output_lse->data_ptr<float>();
  }
  p.prefix_output = prefix_output.data_ptr();
  p.prefix_lse = prefix_lse.data_ptr<float>();
  p.suffix_output = suffix_output.data_ptr();
  p.suffix_lse = suffix_lse.data_ptr<float>();
  p.output_stride = output.stride(0);
  p.prefix_stride = prefix_output.stride(0);
  p.suffix_stride = suffix_output.stride(0);
  p.num_tokens = static_cast<int32_t>(N);
  p.num_heads = static_cast<int32_t>(H);
  p.head_size = static_cast<int32_t>(D);
  p.dtype = dt;
  const at::cuda::OptionalCUDAGuard guard(device_of(prefix_output));
  check_status(b200_merge_attn_states(&p, at::cuda::getCurrentCUDAStream()), "merge_attn_states");
}

// csrc/ops.h:43-53
void paged_attention_v2(at::Tensor& out, at::Tensor& exp_sums, at::Tensor& max_logits, at::Tensor& tmp_out,
                        at::Tensor& query, at::Tensor& key_cache, at::Tensor& value_cache, int64_t num_kv_heads,
                        double scale, at::Tensor& block_tables, at::Tensor& seq_lens, int64_t block_size,
                        int64_t max_seq_len, const std::optional<at::Tensor>& alibi_slopes,
                        const std::string& kv_cache_dtype, at::Tensor& k_scale, at::Tensor& v_scale,
                        const int64_t tp_rank, const int64_t blocksparse_local_blocks,
                        const int64_t blocksparse_vert_stride, const int64_t blocksparse_block_size,
                        const int64_t blocksparse_head_sliding_step) {
  b200_paged_attn_params_t p;
  fill_attn_params(p, out, query, key_cache, value_cache, num_kv_heads, scale, block_tables, seq_lens, block_size,
                   max_seq_len, alibi_slopes, kv_cache_dtype, k_scale, v_scale, tp_rank, blocksparse_local_blocks,
                   blocksparse_vert_stride, blocksparse_block_size, blocksparse_head_sliding_step);
  const int64_t P = (max_seq_len + B200_PARTITION_SIZE - 1) / B200_PARTITION_SIZE;
  const int64_t S = query.size(0), H = query.size(1), D = query.size(2);
  TORCH_CHECK(exp_sums.scalar_type() == at::kFloat && max_logits.scalar_type() == at::kFloat &&
                  exp_sums.is_contiguous() && max_logits.is_contiguous() && exp_sums.numel() >= S * H * P &&
                  max_logits.numel() >= S * H * P,
              "exp_sums / max_logits must be contiguous float32 [num_seqs, num_heads, max_num_partitions]");
  TORCH_CHECK(tmp_out.scalar_type() == query.scalar_type() && tmp_out.is_contiguous() &&
                  tmp_out.numel() >= S * H * P * D,
              "tmp_out must be contiguous [num_seqs, num_heads, max_num_partitions, head_size] of the query dtype");
  TORCH_CHECK(exp_sums.dim() == 3 && exp_sums.size(2) == P,
              "exp_sums.size(2) must be ceil(max_seq_len / 512)");
  p.exp_sums = exp_sums.data_ptr<float>();
  p.max_logits = max_logits.data_ptr<float>();
  p.tmp_out = tmp_out.data_ptr();
  const at::cuda::OptionalCUDAGuard guard(device_of(query));
  const cudaStream_t stream = at::cuda::getCurrentCUDAStream();
  check_status(b200_paged_attention_v2(&p, stream), "paged_attention_v2");
}

void fill_cache_params(b200_cache_write_params_t& p, const at::Tensor& key, const at::Tensor& value,
                       const at::Tensor& key_cache, const at::Tensor& value_cache, const at::Tensor& slot_mapping,
                       const std::string& kv_cache_dtype, const at::Tensor& k_scale, const at::Tensor& v_scale) {
  TORCH_CHECK(key.dim() == 3 && value.dim() == 3, "key/value must be [num_tokens, num_heads, head_size]");
  TORCH_CHECK(key.scalar_type() == value.scalar_type(), "key and value dtypes differ");
  TORCH_CHECK(key.stride(2) == 1 && key.stride(1) == key.size(2) && value.stride(2) == 1 &&
                  value.stride(1) == value.size(2),
              "key/value heads must be contiguous");
  TORCH_CHECK(slot_mapping.scalar_type() == at::kLong && slot_mapping.is_contiguous(),
              "slot_mapping must be contiguous int64");
  TORCH_CHECK(slot_mapping.size(0) <= key.size(0), "slot_mapping longer than key");
  TORCH_CHECK(k_scale.scalar_type() == at::kFloat && v_scale.scalar_type() == at::kFloat,
              "k_scale / v_scale must be float32 tensors");
  TORCH_CHECK(key.is_cuda() && value.is_cuda() && key_cache.is_cuda() && value_cache.is_cuda() &&
                  slot_mapping.is_cuda() && k_scale.is_cuda() && v_scale.is_cuda(),
              "all tensors must be on the CUDA device");
  const int kv = kv_dtype_or_throw(kv_cache_dtype);
  if (kv == B200_KV_AUTO) {
    TORCH_CHECK(key_cache.scalar_type() == key.scalar_type() && value_cache.scalar_type() == key.scalar_type(),
                "kv_cache_dtype 'auto' needs caches of the key dtype");
  } else {
    TORCH_CHECK(is_cache_byte_type(key_cache.scalar_type()) && is_cache_byte_type(value_cache.scalar_type()),
                "fp8 kv cache must be stored as uint8");
  }
  const int dt = to_dtype(key.scalar_type());
  TORCH_CHECK(dt == B200_DTYPE_F32 || dt == B200_DTYPE_F16 || dt == B200_DTYPE_BF16,
              "Unsupported input type of kv cache: ", key.scalar_type());
  memset(&p, 0, sizeof(p));
  p.key = key.data_ptr();
  p.value = value.data_ptr();
  p.key_cache = key_cache.data_ptr();
  p.value_cache = value_cache.data_ptr();
  p.slot_mapping = slot_mapping.data_ptr<int64_t>();
  p.k_scale = k_scale.data_ptr<float>();
  p.v_scale = v_scale.data_ptr<float>();
  p.key_stride = key.stride(0);
  p.value_stride = value.stride(0);
  p.num_tokens = static_cast<int32_t>(slot_mapping.size(0));
  p.num_heads = static_cast<int32_t>(key.size(1));
  p.head_size = static_cast<int32_t>(key.size(2));
  p.dtype = dt;
  p.kv_dtype = kv;
}

// csrc/cache.h:21-26 ; csrc/cache_kernels.cu:407-433
void reshape_and_cache(at::Tensor& key, at::Tensor& value, at::Tensor& key_cache, at::Tensor& value_cache,
                       at::Tensor& slot_mapping, const std::string& kv_cache_dtype, at::Tensor& k_scale,
                       at::Tensor& v_scale) {
  b200_cache_write_params_t p;
  fill_cache_params(p, key, value, key_cache, value_cache, slot_mapping, kv_cache_dtype, k_scale, v_scale);
  TORCH_CHECK(key_cache.dim() == 5 && value_cache.dim() == 4,
              "key_cache must be [nb, nh, hs/x, bs, x] and value_cache [nb, nh, hs, bs]");
  // the reference computes destinations from sizes (csrc/cache_kernels.cu:235-242)
  TORCH_CHECK(key_cache.is_contiguous() && value_cache.is_contiguous(), "reshape_and_cache needs contiguous caches");
  p.block_size = static_cast<int32_t>(key_cache.size(3));
  p.x = static_cast<int32_t>(key_cache.size(4));
  TORCH_CHECK(key_cache.size(1) == key.size(1) && key_cache.size(2) * key_cache.size(4) == key.size(2) &&
                  value_cache.size(1) == key.size(1) && value_cache.size(2) == key.size(2) &&
                  value_cache.size(3) == key_cache.size(3),
              "cache shapes do not match key/value");
  const at::cuda::OptionalCUDAGuard guard(device_of(key));
  check_status(b200_reshape_and_cache(&p, at::cuda::getCurrentCUDAStream()), "reshape_and_cache");
}

// csrc/cache.h:28-35 ; csrc/cache_kernels.cu:450-488
void reshape_and_cache_flash(at::Tensor& key, at::Tensor& value, at::Tensor& key_cache, at::Tensor& value_cache,
                             at::Tensor& slot_mapping, const std::string& kv_cache_dtype, at::Tensor& k_scale,
                             at::Tensor& v_scale) {
  b200_cache_write_params_t p;
  fill_cache_params(p, key, value, key_cache, value_cache, slot_mapping, kv_cache_dtype, k_scale, v_scale);
  TORCH_CHECK(key_cache.dim() == 4 && value_cache.dim() == 4, "caches must be [nb, bs, nh, hs]");
  TORCH_CHECK(key_cache.stride(0) == value_cache.stride(0) && key_cache.stride(1) == value_cache.stride(1) &&
                  key_cache.stride(2) == value_cache.stride(2) && key_cache.stride(3) == 1 &&
                  value_cache.stride(3) == 1,
              "key_cache and value_cache must share strides with contiguous head dims");
  TORCH_CHECK(key.stride(0) == value.stride(0), "key and value token strides differ");  // cache_kernels.cu:479
  p.block_size = static_cast<int32_t>(key_cache.size(1));
  p.block_stride = key_cache.stride(0);
  p.page_stride = key_cache.stride(1);
  p.head_stride = key_cache.stride(2);
  TORCH_CHECK(key_cache.size(2) == key.size(1) && key_cache.size(3) == key.size(2), "cache shapes do not match key");
  const at::cuda::OptionalCUDAGuard guard(device_of(key));
  check_status(b200_reshape_and_cache_flash(&p, at::cuda::getCurrentCUDAStream()), "reshape_and_cache_flash");
}

// csrc/cache.h:37-40 ; csrc/cache_kernels.cu:543-614
void convert_fp8(at::Tensor& dst_cache, at::Tensor& src_cache, const double scale, const std::string& kv_cache_dtype) {
  TORCH_CHECK(src_cache.device() == dst_cache.device(), "src and dst must be on the same device");
  TORCH_CHECK(src_cache.is_cuda(), "src must be on a GPU");  // cache_kernels.cu:568-570
  TORCH_CHECK(dst_cache.numel() == src_cache.numel(), "src and dst sizes differ");
  TORCH_CHECK(dst_cache.is_contiguous() && src_cache.is_contiguous(), "convert_fp8 needs contiguous tensors");
  // "auto" is accepted by the reference's string switch (cache_kernels.cu:581) but
  // lands in an assert(false) conversion; only fp8 / fp8_e4m3 have defined results.
  TORCH_CHECK(b200_kv_dtype_from_string(kv_cache_dtype.c_str()) == B200_KV_FP8_E4M3,
              "Unsupported data type: ", kv_cache_dtype);
  const at::cuda::OptionalCUDAGuard guard(device_of(src_cache));
  check_status(b200_convert_fp8(dst_cache.data_ptr(), src_cache.data_ptr(), src_cache.numel(),
                                static_cast<float>(scale), to_dtype(dst_cache.scalar_type()),
                                to_dtype(src_cache.scalar_type()), B200_KV_FP8_E4M3,
                                at::cuda::getCurrentCUDAStream()),
               "convert_fp8");
}

// csrc/cache.h ; csrc/cache_kernels.cu:18-60
void swap_blocks(at::Tensor& src, at::Tensor& dst, const at::Tensor& block_mapping) {
  int kind;
  if (src.is_cuda() && dst.is_cuda()) {
    TORCH_CHECK(src.get_device() == dst.get_device(), "src and dst must be on the same GPU");
    kind = B200_COPY_D2D;
  } else if (src.is_cuda() && dst.is_cpu()) {
    kind = B200_COPY_D2H;
  } else if (src.is_cpu() && dst.is_cuda()) {
    kind = B200_COPY_H2D;
  } else {
    TORCH_CHECK(false, "Invalid device combination");
  }
  TORCH_CHECK(block_mapping.device().is_cpu(), "block_mapping must be on CPU");
  TORCH_CHECK(block_mapping.scalar_type() == at::kLong && block_mapping.dim() == 2 && block_mapping.size(1) == 2,
              "block_mapping must be an int64 tensor [num_pairs, 2]");
  const at::Tensor mapping = block_mapping.contiguous();
  // stride, not numel: blocks may be padded (cache_kernels.cu:43-46)
  const int64_t block_bytes = src.element_size() * src.stride(0);
  TORCH_CHECK(dst.element_size() * dst.stride(0) == block_bytes, "src and dst blocks differ in size");
  const at::cuda::OptionalCUDAGuard guard(src.is_cuda() ? src.device() : dst.device());
  check_status(b200_swap_blocks(dst.data_ptr(), src.data_ptr(), mapping.data_ptr<int64_t>(), mapping.size(0),
                                block_bytes, kind, at::cuda::getCurrentCUDAStream()),
               "swap_blocks");
}

// csrc/cache.h ; csrc/cache_kernels.cu:113-163
void copy_blocks(std::vector<at::Tensor> const& key_caches, std::vector<at::Tensor> const& value_caches,
                 const at::Tensor& block_mapping) {
  const size_t num_layers = key_caches.size();
  TORCH_CHECK(num_layers == value_caches.size());
  if (num_layers == 0) return;
  TORCH_CHECK(key_caches[0].is_cuda());
  TORCH_CHECK(block_mapping.is_cuda() && block_mapping.scalar_type() == at::kLong && block_mapping.is_contiguous() &&
                  block_mapping.dim() == 2 && block_mapping.size(1) == 2,
              "block_mapping must be a contiguous int64 CUDA tensor [num_pairs, 2]");
  std::vector<void*> kp(num_layers), vp(num_layers);
  const int64_t block_bytes = key_caches[0][0].numel() * key_caches[0].element_size();  // cache_kernels.cu:148
  for (size_t l = 0; l < num_layers; ++l) {
    TORCH_CHECK(key_caches[l].device() == key_caches[0].device() && value_caches[l].device() == key_caches[0].device(),
                "all caches must be on the same device");
    kp[l] = key_caches[l].data_ptr();
    vp[l] = value_caches[l].data_ptr();
  }
  const at::cuda::OptionalCUDAGuard guard(key_caches[0].device());
  check_status(b200_copy_blocks(kp.data(), vp.data(), static_cast<int32_t>(num_layers), block_mapping.data_ptr<int64_t>(),
                                static_cast<int32_t>(block_mapping.size(0)), block_bytes,
                                at::cuda::getCurrentCUDAStream()),
               "copy_blocks");
}

// csrc/cache.h ; csrc/cache_kernels.cu:165-201
void copy_blocks_mla(std::vector<at::Tensor> const& kv_caches, const at::Tensor& block_mapping) {
  const size_t num_layers = kv_caches.size();
  if (num_layers == 0) return;
  TORCH_CHECK(kv_caches[0].is_cuda(), "kv_cache must be on CUDA");
  TORCH_CHECK(block_mapping.is_cuda() && block_mapping.scalar_type() == at::kLong && block_mapping.is_contiguous() &&
                  block_mapping.dim() == 2 && block_mapping.size(1) == 2,
              "block_mapping must be a contiguous int64 CUDA tensor [num_pairs, 2]");
  std::vector<void*> ptrs(num_layers);
  for (size_t l = 0; l < num_layers; ++l) {
    TORCH_CHECK(kv_caches[l].device() == kv_caches[0].device(), "all caches must be on the same device");
    ptrs[l] = kv_caches[l].data_ptr();
  }
  // stride, not numel: blocks may be padded (cache_kernels.cu:186-189)
  const int64_t block_bytes = kv_caches[0].stride(0) * kv_caches[0].element_size();
  const at::cuda::OptionalCUDAGuard guard(kv_caches[0].device());
  check_status(b200_copy_blocks_mla(ptrs.data(), static_cast<int32_t>(num_layers), block_mapping.data_ptr<int64_t>(),
                                    static_cast<int32_t>(block_mapping.size(0)), block_bytes,
                                    at::cuda::getCurrentCUDAStream()),
               "copy_blocks_mla");
}

// csrc/cache.h ; csrc/cache_kernels.cu:490-539
void concat_and_cache_mla(at::Tensor& kv_c, at::Tensor& k_pe, at::Tensor& kv_cache, at::Tensor& slot_mapping,
                          const std::string& kv_cache_dtype, at::Tensor& scale) {
  TORCH_CHECK(kv_c.dim() == 2 && k_pe.dim() == 2 && kv_cache.dim() == 3, "kv_c [T, rank], k_pe [T, pe], kv_cache [nb, bs, rank + pe]");
  TORCH_CHECK(kv_cache.size(2) == kv_c.size(1) + k_pe.size(1));  // :511
  TORCH_CHECK(kv_c.scalar_type() == k_pe.scalar_type(), "kv_c and k_pe dtypes differ");
  TORCH_CHECK(kv_c.stride(1) == 1 && k_pe.stride(1) == 1 && kv_cache.stride(2) == 1, "innermost dims must be contiguous");
  TORCH_CHECK(slot_mapping.scalar_type() == at::kLong && slot_mapping.is_contiguous(), "slot_mapping must be contiguous int64");
  TORCH_CHECK(slot_mapping.size(0) <= kv_c.size(0) && slot_mapping.size(0) <= k_pe.size(0), "slot_mapping longer than kv_c / k_pe");
  TORCH_CHECK(kv_c.is_cuda() && k_pe.is_cuda() && kv_cache.is_cuda() && slot_mapping.is_cuda() && scale.is_cuda(),
              "all tensors must be on the CUDA device");
  const int kv = kv_dtype_or_throw(kv_cache_dtype);
  if (kv == B200_KV_AUTO) {
    TORCH_CHECK(kv_cache.scalar_type() == kv_c.scalar_type(), "kv_cache_dtype 'auto' needs a cache of the kv_c dtype");
  } else {
    TORCH_CHECK(is_cache_byte_type(kv_cache.scalar_type()), "fp8 kv cache must be stored as uint8");
    TORCH_CHECK(scale.scalar_type() == at::kFloat, "scale must be a float32 tensor");
  }
  const int dt = to_dtype(kv_c.scalar_type());
  TORCH_CHECK(dt == B200_DTYPE_F32 || dt == B200_DTYPE_F16 || dt == B200_DTYPE_BF16,
              "Unsupported input type of kv cache: ", kv_c.scalar_type());
  b200_concat_mla_params_t p;
  memset(&p, 0, sizeof(p));
  p.kv_c = kv_c.data_ptr();
  p.k_pe = k_pe.data_ptr();
  p.kv_cache = kv_cache.data_ptr();
  p.slot_mapping = slot_mapping.data_ptr<int64_t>();
  p.scale = kv == B200_KV_AUTO ? nullptr : scale.data_ptr<float>();
  p.kv_c_stride = kv_c.stride(0);
  p.k_pe_stride = k_pe.stride(0);
  p.block_stride = kv_cache.stride(0);
  p.entry_stride = kv_cache.stride(1);
  p.num_tokens = static_cast<int32_t>(slot_mapping.size(0));
  p.kv_lora_rank = static_cast<int32_t>(kv_c.size(1));
  p.pe_dim = static_cast<int32_t>(k_pe.size(1));
  p.block_size = static_cast<int32_t>(kv_cache.size(1));
  p.dtype = dt;
  p.kv_dtype = kv;
  const at::cuda::OptionalCUDAGuard guard(device_of(kv_c));
  check_status(b200_concat_and_cache_mla(&p, at::cuda::getCurrentCUDAStream()), "concat_and_cache_mla");
}

// ---- MLA decode attention -------------------------------------------------------------------------------------
// csrc/attention/mla/cutlass_mla_kernels.cu:154-202 (the reference's checks; it requires exactly 128 heads and the
// caller pads, any count <= 128 is accepted here) and :141-147 (the reference allocates its workspace inside the op).
void mla_decode_common(const at::Tensor& out, const at::Tensor* lse, const at::Tensor& q_nope, const at::Tensor& q_pe,
                       const at::Tensor& cache, const at::Tensor& seq_lens, const at::Tensor& page_table,
                       const at::Tensor* workspace, double scale, int64_t num_kv_splits, const char* what) {
  TORCH_CHECK(q_nope.is_cuda(), "q_nope must be on CUDA");
  TORCH_CHECK(q_nope.dim() == 3, "q_nope must be a 3D tensor");
  TORCH_CHECK(q_pe.dim() == 3, "q_pe must be a 3D tensor");
  TORCH_CHECK(cache.dim() == 3, "kv_c_and_k_pe_cache must be a 3D tensor");
  TORCH_CHECK(seq_lens.dim() == 1, "seq_lens must be a 1D tensor");
  TORCH_CHECK(page_table.dim() == 2, "page_table must be a 2D tensor");
  TORCH_CHECK(out.dim() == 3, "out must be a 3D tensor");
  const int64_t B = q_nope.size(0), H = q_nope.size(1);
  TORCH_CHECK(q_nope.size(2) == 512, "D_q_nope must be equal to 512");
  TORCH_CHECK(q_pe.size(2) == 64, "D_q_pe must be equal to 64");
  TORCH_CHECK(cache.size(2) == 576, "D_ckv must be equal to 576");
  TORCH_CHECK(out.size(2) == 512, "D_o must be equal to 512");
  TORCH_CHECK(H == q_pe.size(1) && H == out.size(1) && H >= 1 && H <= 128,
              "H_q_nope, H_q_pe, and H_o must be equal and at most 128");
  const int64_t page = cache.size(1);
  TORCH_CHECK(page > 0 && (page & (page - 1)) == 0, "PAGE_SIZE must be a power of 2");
  TORCH_CHECK(page >= 16, "PAGE_SIZE must be at least 16");
  TORCH_CHECK(B == q_pe.size(0) && B == page_table.size(0) && B == out.size(0) && B == seq_lens.size(0),
              "Batch dims must be same for page_table, q_nope and q_pe, and out");
  const bool fp8 = q_nope.scalar_type() == at::kFloat8_e4m3fn;
  TORCH_CHECK(q_nope.scalar_type() == at::kHalf || q_nope.scalar_type() == at::kBFloat16 || fp8,
              "q_nope must be a half, bfloat16, or float8_e4m3fn tensor");
  TORCH_CHECK(cache.scalar_type() == q_nope.scalar_type() && q_pe.scalar_type() == q_nope.scalar_type(),
              "kv_c_and_k_pe_cache, q_nope, and q_pe must be the same type");
  // float8 inputs (descale 1, as in tests/kernels/attention/test_cutlass_mla_decode.py:91-101): out is bf16 / half
  TORCH_CHECK(fp8 ? (out.scalar_type() == at::kHalf || out.scalar_type() == at::kBFloat16)
                  : out.scalar_type() == q_nope.scalar_type(),
              "out must have the query type (bfloat16 / half for float8 queries)");
  TORCH_CHECK(seq_lens.scalar_type() == at::kInt, "seq_lens must be a 32-bit integer tensor");
  TORCH_CHECK(page_table.scalar_type() == at::kInt, "page_table must be a 32-bit integer tensor");
  TORCH_CHECK(q_nope.stride(2) == 1 && q_pe.stride(2) == 1 && cache.stride(2) == 1 && out.stride(2) == 1 &&
                  page_table.stride(1) == 1 && seq_lens.is_contiguous(),
              "innermost dims must be contiguous");
  TORCH_CHECK(q_pe.is_cuda() && cache.is_cuda() && seq_lens.is_cuda() && page_table.is_cuda() && out.is_cuda(),
              "all tensors must be on the CUDA device");
  b200_mla_decode_params_t p;
  memset(&p, 0, sizeof(p));
  p.out = out.data_ptr();
  p.q_nope = q_nope.data_ptr();
  p.q_pe = q_pe.data_ptr();
  p.kv_cache = cache.data_ptr();
  p.seq_lens = seq_lens.data_ptr<int32_t>();
  p.page_table = page_table.data_ptr<int32_t>();
  p.out_seq_stride = out.stride(0), p.out_head_stride = out.stride(1);
  p.q_nope_seq_stride = q_nope.stride(0), p.q_nope_head_stride = q_nope.stride(1);
  p.q_pe_seq_stride = q_pe.stride(0), p.q_pe_head_stride = q_pe.stride(1);
  p.block_stride = cache.stride(0), p.entry_stride = cache.stride(1);
  p.page_table_stride = page_table.stride(0);
  p.num_blocks = cache.size(0);
  p.scale = static_cast<float>(scale);
  p.num_seqs = static_cast<int32_t>(B);
  p.num_heads = static_cast<int32_t>(H);
  p.page_size = static_cast<int32_t>(page);
  p.max_pages = static_cast<int32_t>(page_table.size(1));
  p.dtype = to_dtype(out.scalar_type());
  p.kv_dtype = fp8 ? B200_KV_FP8_E4M3 : B200_KV_AUTO;
  p.num_kv_splits = num_kv_splits <= 0 ? 0 : static_cast<int32_t>(num_kv_splits);
  if (lse != nullptr) {
    TORCH_CHECK(lse->scalar_type() == at::kFloat && lse->dim() == 2 && lse->size(0) == B && lse->size(1) >= H &&
                    lse->stride(1) == 1 && lse->is_cuda(),
                "lse must be a float32 [batch, heads] CUDA tensor");
    p.lse = lse->data_ptr<float>();
    p.lse_stride = lse->stride(0);
  }
  const at::cuda::OptionalCUDAGuard guard(device_of(q_nope));
  const uint64_t need = b200_mla_decode_workspace_bytes(p.num_seqs, p.num_heads, (int64_t)p.max_pages * p.page_size,
                                                        p.num_kv_splits);
  at::Tensor own;
  if (need > 0) {
    if (workspace != nullptr && workspace->is_cuda() && (uint64_t)workspace->nbytes() >= need &&
        reinterpret_cast<uintptr_t>(workspace->data_ptr()) % 16 == 0) {
      p.workspace = workspace->data_ptr();
      p.workspace_bytes = workspace->nbytes();
    } else {  // as the reference op does (cutlass_mla_kernels.cu:141-147)
      own = at::empty({static_cast<int64_t>(need)}, q_nope.options().dtype(at::kByte));
      p.workspace = own.data_ptr();
      p.workspace_bytes = need;
    }
  }
  check_status(b200_mla_decode(&p, at::cuda::getCurrentCUDAStream()), what);
}

// csrc/torch_bindings.cpp:287-290 ; csrc/attention/mla/cutlass_mla_entry.cu:28-38
void cutlass_mla_decode(at::Tensor& out, const at::Tensor& q_nope, const at::Tensor& q_pe, const at::Tensor& cache,
                        const at::Tensor& seq_lens, const at::Tensor& page_table, double scale) {
  mla_decode_common(out, nullptr, q_nope, q_pe, cache, seq_lens, page_table, nullptr, scale, 0, "cutlass_mla_decode");
}

// upstream 0.22's form of the same op (the one the reference's test calls, tests/kernels/attention/
// test_cutlass_mla_decode.py:121-123): LSE output, caller-provided workspace, split count (-1 = choose)
void sm100_cutlass_mla_decode(at::Tensor& out, at::Tensor& lse, const at::Tensor& q_nope, const at::Tensor& q_pe,
                              const at::Tensor& cache, const at::Tensor& seq_lens, const at::Tensor& page_table,
                              const at::Tensor& workspace, double scale, int64_t num_kv_splits) {
  mla_decode_common(out, &lse, q_nope, q_pe, cache, seq_lens, page_table, &workspace, scale, num_kv_splits,
                    "sm100_cutlass_mla_decode");
}

int64_t sm100_cutlass_mla_get_workspace_size(int64_t max_seq_len, int64_t num_batches, int64_t sm_count,
                                             int64_t num_kv_splits) {
  (void)sm_count;
  // the head count is not an argument of the stock op: size for whichever kernel the call will pick (<= 64 heads: flat
  // schedule; 65..128: pair kernel)
  const int32_t n = num_kv_splits <= 0 ? 0 : static_cast<int32_t>(num_kv_splits);
  uint64_t need = 0;
  for (int32_t heads : {16, 32, 64, 128})
    need = std::max<uint64_t>(need, b200_mla_decode_workspace_bytes(static_cast<int32_t>(num_batches), heads, max_seq_len, n));
  return static_cast<int64_t>(need);
}

void fill_gather_params(b200_gather_cache_params_t& p, const at::Tensor& src_cache, const at::Tensor& dst,
                        const at::Tensor& block_table, const at::Tensor& cu_seq_lens, int64_t batch_size,
                        const std::optional<at::Tensor>& seq_starts) {
  TORCH_CHECK(block_table.scalar_type() == at::kInt, "block_table must be int32");
  TORCH_CHECK(cu_seq_lens.scalar_type() == at::kInt, "cu_seq_lens must be int32");
  TORCH_CHECK(src_cache.is_cuda(), "src_cache must be on a GPU");
  TORCH_CHECK(src_cache.device() == dst.device(), "src_cache and dst must be on the same device");
  TORCH_CHECK(src_cache.device() == block_table.device(), "src_cache and block_table must be on the same device");
  TORCH_CHECK(src_cache.device() == cu_seq_lens.device(), "src_cache and cu_seq_lens must be on the same device");
  TORCH_CHECK(src_cache.dim() >= 3, "src_cache must be [num_blocks, block_size, entries...]");
  TORCH_CHECK(cu_seq_lens.is_contiguous() && cu_seq_lens.numel() >= batch_size + 1, "cu_seq_lens needs batch_size + 1 entries");
  TORCH_CHECK(block_table.dim() == 2 && block_table.stride(1) == 1 && block_table.size(0) >= batch_size,
              "block_table must be [batch, blocks] with contiguous rows");
  memset(&p, 0, sizeof(p));
  p.src_cache = src_cache.data_ptr();
  p.dst = dst.data_ptr();
  p.block_table = block_table.data_ptr<int32_t>();
  p.cu_seq_lens = cu_seq_lens.data_ptr<int32_t>();
  if (seq_starts.has_value() && seq_starts->defined()) {
    TORCH_CHECK(seq_starts->scalar_type() == at::kInt, "seq_starts must be int32");
    TORCH_CHECK(src_cache.device() == seq_starts->device(), "src_cache and seq_starts must be on the same device");
    TORCH_CHECK(seq_starts->is_contiguous() && seq_starts->numel() >= batch_size, "seq_starts needs batch_size entries");
    p.seq_starts = seq_starts->data_ptr<int32_t>();
  }
  int64_t entry = 1;
  for (int64_t i = 2; i < src_cache.dim(); ++i) entry *= src_cache.size(i);  // flatten(2, -1), cache_kernels.cu:720
  // the entry (and each dst row) must be one contiguous run
  int64_t expect = 1;
  for (int64_t i = src_cache.dim() - 1; i >= 2; --i) {
    TORCH_CHECK(src_cache.size(i) == 1 || src_cache.stride(i) == expect, "src_cache entries must be contiguous");
    expect *= src_cache.size(i);
  }
  TORCH_CHECK(dst.dim() >= 2 && dst.numel() == dst.size(0) * entry, "dst must be [tot_tokens, entries...]");
  p.block_table_stride = block_table.stride(0);
  p.cache_block_stride = src_cache.stride(0);
  p.cache_entry_stride = src_cache.stride(1);
  p.dst_entry_stride = dst.stride(0);
  p.batch_size = static_cast<int32_t>(batch_size);
  p.block_size = static_cast<int32_t>(src_cache.size(1));
  p.entry_size = static_cast<int32_t>(entry);
}

int width_dtype(int64_t bytes) {
  switch (bytes) {
    case 4: return B200_DTYPE_F32;
    case 2: return B200_DTYPE_F16;
    case 1: return B200_DTYPE_U8;
    default: return -1;
  }
}

// csrc/cache.h ; csrc/cache_kernels.cu:702-773
void gather_and_maybe_dequant_cache(const at::Tensor& src_cache, const at::Tensor& dst, const at::Tensor& block_table,
                                    const at::Tensor& cu_seq_lens, int64_t batch_size,
                                    const std::string& kv_cache_dtype, const at::Tensor& scale,
                                    std::optional<at::Tensor> seq_starts) {
  b200_gather_cache_params_t p;
  fill_gather_params(p, src_cache, dst, block_table, cu_seq_lens, batch_size, seq_starts);
  const int kv = kv_dtype_or_throw(kv_cache_dtype);
  p.kv_dtype = kv;
  p.token_granular_starts = 0;
  if (kv == B200_KV_AUTO) {
    TORCH_CHECK(src_cache.scalar_type() == dst.scalar_type(), "kv_cache_dtype 'auto' needs src_cache of the dst dtype");
    p.src_dtype = p.dst_dtype = width_dtype(dst.element_size());
    TORCH_CHECK(p.src_dtype >= 0, "Unsupported data type width: ", dst.element_size() * 8);
  } else {
    TORCH_CHECK(is_cache_byte_type(src_cache.scalar_type()), "fp8 kv cache must be stored as uint8");
    const int dt = to_dtype(dst.scalar_type());
    TORCH_CHECK(dt == B200_DTYPE_F32 || dt == B200_DTYPE_F16 || dt == B200_DTYPE_BF16, "Unsupported data type: ",
                dst.scalar_type());
    TORCH_CHECK(scale.is_cuda() && scale.scalar_type() == at::kFloat, "scale must be a float32 CUDA tensor");
    p.src_dtype = B200_DTYPE_U8;
    p.dst_dtype = dt;
    p.scale = scale.data_ptr<float>();
  }
  if (dst.numel() == 0) return;  // every sequence is empty: nothing to gather
  const at::cuda::OptionalCUDAGuard guard(device_of(src_cache));
  check_status(b200_gather_cache(&p, at::cuda::getCurrentCUDAStream()), "gather_and_maybe_dequant_cache");
}

// csrc/cache.h ; csrc/cache_kernels.cu:1075-1134
void cp_gather_cache(const at::Tensor& src_cache, const at::Tensor& dst, const at::Tensor& block_table,
                     const at::Tensor& cu_seq_lens, int64_t batch_size, std::optional<at::Tensor> seq_starts) {
  b200_gather_cache_params_t p;
  fill_gather_params(p, src_cache, dst, block_table, cu_seq_lens, batch_size, seq_starts);
  TORCH_CHECK(src_cache.scalar_type() == dst.scalar_type(), "src_cache and dst must have the same dtype");
  p.kv_dtype = B200_KV_AUTO;
  p.token_granular_starts = 1;
  p.src_dtype = p.dst_dtype = width_dtype(src_cache.element_size());
  TORCH_CHECK(p.src_dtype >= 0, "Unsupported data type width: ", src_cache.element_size() * 8);
  if (dst.numel() == 0) return;  // every sequence is empty: nothing to gather
  const at::cuda::OptionalCUDAGuard guard(device_of(src_cache));
  check_status(b200_gather_cache(&p, at::cuda::getCurrentCUDAStream()), "cp_gather_cache");
}

const char* kSchemaV1 =
    "paged_attention_v1("
    "    Tensor! out, Tensor query, Tensor key_cache,"
    "    Tensor value_cache, int num_kv_heads, float scale,"
    "    Tensor block_tables, Tensor seq_lens, int block_size,"
    "    int max_seq_len, Tensor? alibi_slopes,"
    "    str kv_cache_dtype, Tensor k_scale, Tensor v_scale,"
    "    int tp_rank, int blocksparse_local_blocks,"
    "    int blocksparse_vert_stride, int blocksparse_block_size,"
    "    int blocksparse_head_sliding_step) -> ()";
const char* kSchemaV2 =
    "paged_attention_v2("
    "    Tensor! out, Tensor! exp_sums, Tensor! max_logits,"
    "    Tensor! tmp_out, Tensor query, Tensor key_cache,"
    "    Tensor value_cache, int num_kv_heads, float scale,"
    "    Tensor block_tables, Tensor seq_lens, int block_size,"
    "    int max_seq_len, Tensor? alibi_slopes,"
    "    str kv_cache_dtype, Tensor k_scale, Tensor v_scale,"
    "    int tp_rank, int blocksparse_local_blocks,"
    "    int blocksparse_vert_stride, int blocksparse_block_size,"
    "    int blocksparse_head_sliding_step) -> ()";
const char* kSchemaReshape =
    "reshape_and_cache(Tensor key, Tensor value,"
    "                  Tensor! key_cache, Tensor! value_cache,"
    "                  Tensor slot_mapping,"
    "                  str kv_cache_dtype,"
    "                  Tensor k_scale, Tensor v_scale) -> ()";
const char* kSchemaReshapeFlash =
    "reshape_and_cache_flash(Tensor key, Tensor value,"
    "                        Tensor! key_cache,"
    "                        Tensor! value_cache,"
    "                        Tensor slot_mapping,"
    "                        str kv_cache_dtype,"
    "                        Tensor k_scale, Tensor v_scale) -> ()";
const char* kSchemaConvert =
    "convert_fp8(Tensor! dst_cache, Tensor src_cache, float scale, "
    "str kv_cache_dtype) -> ()";

// csrc/torch_bindings.cpp:74-82
const char* kSchemaMerge =
    "merge_attn_states("
    "    Tensor! output,"
    "    Tensor!? output_lse,"
    "    Tensor prefix_output,"
    "    Tensor prefix_lse,"
    "    Tensor suffix_output,"
    "    Tensor suffix_lse) -> ()";
// csrc/torch_bindings.cpp:380, :385-386, :431-435, :440-441
const char* kSchemaSwap = "swap_blocks(Tensor src, Tensor! dst, Tensor block_mapping) -> ()";
const char* kSchemaCopy =
    "copy_blocks(Tensor(a!)[] key_caches, Tensor[](b!) value_caches, "
    "Tensor block_mapping) -> ()";
// csrc/torch_bindings.cpp:389-390, :415-419
const char* kSchemaCopyMla = "copy_blocks_mla(Tensor(a!)[] kv_caches, Tensor block_mapping) -> ()";
const char* kSchemaConcatMla =
    "concat_and_cache_mla(Tensor kv_c, Tensor k_pe,"
    "                     Tensor! kv_cache,"
    "                     Tensor slot_mapping,"
    "                     str kv_cache_dtype,"
    "                     Tensor scale) -> ()";
const char* kSchemaGather =
    "gather_and_maybe_dequant_cache(Tensor src_cache, Tensor! dst, "
    "                               Tensor block_table, Tensor cu_seq_lens, "
    "                               int batch_size, "
    "                               str kv_cache_dtype, "
    "                               Tensor scale, Tensor? seq_starts) -> ()";
const char* kSchemaCpGather =
    "cp_gather_cache(Tensor src_cache, Tensor! dst, Tensor block_table, "
    "Tensor cu_seq_lens, int batch_size, Tensor? seq_starts) -> ()";
// ours (namespace _C_b200): decode attention that also returns the LSE
const char* kSchemaV1Lse =
    "paged_attention_v1_lse(Tensor! out, Tensor! lse, Tensor query, Tensor key_cache, Tensor value_cache, "
    "int num_kv_heads, float scale, Tensor block_tables, Tensor seq_lens, int block_size, int max_seq_len, "
    "Tensor? alibi_slopes, str kv_cache_dtype, Tensor k_scale, Tensor v_scale) -> ()";

// csrc/torch_bindings.cpp:216-218
const char* kSchemaRotary =
    "rotary_embedding(Tensor positions, Tensor! query,"
    "                 Tensor!? key, int head_size,"
    "                 Tensor cos_sin_cache, bool is_neox) -> ()";
const char* kSchemaRotaryCache =
    "rotary_embedding_and_cache_flash(Tensor positions, Tensor! query, Tensor! key, Tensor value, int head_size, "
    "Tensor cos_sin_cache, bool is_neox, Tensor! key_cache, Tensor! value_cache, Tensor slot_mapping, "
    "str kv_cache_dtype, Tensor k_scale, Tensor v_scale) -> ()";
const char* kSchemaReserve = "reserve_workspace(Tensor query_like, int max_seq_len, bool reset=False) -> ()";
const char* kSchemaFlash =
    "paged_attention_flash(Tensor! out, Tensor!? lse, Tensor query, Tensor key_cache, Tensor value_cache, "
    "float scale, Tensor block_tables, Tensor seq_lens, int max_seq_len, Tensor? alibi_slopes) -> ()";

// csrc/torch_bindings.cpp:287-290
const char* kSchemaCutlassMla =
    "cutlass_mla_decode(Tensor! out, Tensor q_nope, Tensor q_pe,"
    "                   Tensor kv_c_and_k_pe_cache, Tensor seq_lens,"
    "                   Tensor page_table, float scale) -> ()";
// upstream vLLM 0.22.0 (_C_stable_libtorch), string-identical
const char* kSchemaSm100Mla =
    "sm100_cutlass_mla_decode(Tensor! out, Tensor! lse, Tensor q_nope,"
    "                         Tensor q_pe, Tensor kv_c_and_k_pe_cache,"
    "                         Tensor seq_lens, Tensor page_table,"
    "                         Tensor workspace, float scale,"
    "                         int num_kv_splits) -> ()";
const char* kSchemaSm100MlaWs =
    "sm100_cutlass_mla_get_workspace_size(int max_seq_len, int num_batches,"
    "                                     int sm_count, int num_kv_splits) -> int";

bool op_exists(const char* qualified) {
  return c10::Dispatcher::singleton().findSchema({qualified, ""}).has_value();
}

template <typename F>
void register_op(std::vector<std::unique_ptr<torch::Library>>& keep, const char* ns, const char* name,
                 const char* schema, F* fn) {
  const std::string qualified = std::string(ns) + "::" + name;
  if (auto h = c10::Dispatcher::singleton().findSchema({qualified.c_str(), ""})) {
    // another library (upstream vllm._C) already defined this op.  Newer upstream releases changed some of these
    // signatures (swap_blocks, gather_and_maybe_dequant_cache, merge_attn_states); a kernel with the reference's
    // argument list must not be bound to a different schema, so such ops keep their upstream implementation.
    const size_t nargs = std::count(schema, schema + strlen(schema), ',') + 1;
    if (h->schema().arguments().size() != nargs) return;
  }
  if (!op_exists(qualified.c_str())) {
    auto frag = std::make_unique<torch::Library>(torch::Library::FRAGMENT, ns, std::nullopt, __FILE__, __LINE__);
    frag->def(schema);
    keep.push_back(std::move(frag));
  }
  auto impl = std::make_unique<torch::Library>(torch::Library::IMPL, ns, c10::DispatchKey::CUDA, __FILE__, __LINE__);
  impl->impl(name, fn);
  keep.push_back(std::move(impl));
}

void register_all() {
  static std::once_flag once;
  std::call_once(once, [] {
    static std::vector<std::unique_ptr<torch::Library>> keep;
    register_op(keep, "_C", "paged_attention_v1", kSchemaV1, &paged_attention_v1);
    register_op(keep, "_C", "paged_attention_v2", kSchemaV2, &paged_attention_v2);
    register_op(keep, "_C_cache_ops", "reshape_and_cache", kSchemaReshape, &reshape_and_cache);
    register_op(keep, "_C_cache_ops", "reshape_and_cache_flash", kSchemaReshapeFlash, &reshape_and_cache_flash);
    register_op(keep, "_C_cache_ops", "convert_fp8", kSchemaConvert, &convert_fp8);
    register_op(keep, "_C", "merge_attn_states", kSchemaMerge, &merge_attn_states);
    register_op(keep, "_C_cache_ops", "swap_blocks", kSchemaSwap, &swap_blocks);
    register_op(keep, "_C_cache_ops", "copy_blocks", kSchemaCopy, &copy_blocks);
    register_op(keep, "_C_cache_ops", "gather_and_maybe_dequant_cache", kSchemaGather, &gather_and_maybe_dequant_cache);
    register_op(keep, "_C_cache_ops", "cp_gather_cache", kSchemaCpGather, &cp_gather_cache);
    register_op(keep, "_C_cache_ops", "copy_blocks_mla", kSchemaCopyMla, &copy_blocks_mla);
    register_op(keep, "_C_cache_ops", "concat_and_cache_mla", kSchemaConcatMla, &concat_and_cache_mla);
    register_op(keep, "_C", "cutlass_mla_decode", kSchemaCutlassMla, &cutlass_mla_decode);
    register_op(keep, "_C", "sm100_cutlass_mla_decode", kSchemaSm100Mla, &sm100_cutlass_mla_decode);
    if (!op_exists("_C::sm100_cutlass_mla_get_workspace_size")) {  // no tensor arguments: a catch-all kernel
      auto frag = std::make_unique<torch::Library>(torch::Library::FRAGMENT, "_C", std::nullopt, __FILE__, __LINE__);
      frag->def(kSchemaSm100MlaWs);
      frag->impl("sm100_cutlass_mla_get_workspace_size", &sm100_cutlass_mla_get_workspace_size);
      keep.push_back(std::move(frag));
    }
    register_op(keep, "_C_b200", "paged_attention_v1_lse", kSchemaV1Lse, &paged_attention_v1_lse);
    register_op(keep, "_C_b200", "paged_attention_flash", kSchemaFlash, &paged_attention_flash);
    register_op(keep, "_C", "rotary_embedding", kSchemaRotary, &rotary_embedding);
    register_op(keep, "_C_b200", "reserve_workspace", kSchemaReserve, &reserve_workspace);
    register_op(keep, "_C_b200", "rotary_embedding_and_cache_flash", kSchemaRotaryCache, &rotary_embedding_and_cache_flash);
  });
}

struct StaticInit {
  StaticInit() { register_all(); }
} g_static_init;

}  // namespace

// `import vllm_metax_b200._C` only has to load the library (csrc/core/registration.h:22-27).
PyMODINIT_FUNC PyInit__C(void) {
  static struct PyModuleDef module = {PyModuleDef_HEAD_INIT, "_C", nullptr, 0, nullptr};
  register_all();
  return PyModule_Create(&module);
}
