"""C-ABI boundary checks that need no GPU: the library loads, exports every symbol the header declares,
the ctypes mirrors match the C struct layouts, host-side validation returns the documented errors, the
torch shim registers the reference's schemas, and there is no CPU fallback."""
import ctypes as C
import os
import re
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "b200_paged_attn.h")


@pytest.fixture(scope="module")
def native():
    from vllm_metax_b200 import build

    build.build_all()
    from vllm_metax_b200 import _native

    _native.load()
    return _native


def _declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", src)))


def test_every_declared_symbol_is_exported(native):
    names = _declared_functions()
    assert sorted(native.EXPORTED_SYMBOLS) == names
    lib = C.CDLL(os.fspath(native.LIB_PATH))
    for n in names:
        assert hasattr(lib, n), n
    assert native.load().b200_abi_version() == 3


def test_ctypes_structs_match_header_layout(native, tmp_path):
    def fields(struct):
        return [f[0] for f in struct._fields_]

    prog = ['#include <stdio.h>', '#include <stddef.h>', '#include "b200_paged_attn.h"', "int main(void){"]
    for cname, st in (("b200_paged_attn_params_t", native.PagedAttnParams),
                      ("b200_cache_write_params_t", native.CacheWriteParams),
                      ("b200_merge_attn_states_params_t", native.MergeAttnStatesParams),
                      ("b200_gather_cache_params_t", native.GatherCacheParams),
                      ("b200_rotary_params_t", native.RotaryParams),
                      ("b200_concat_mla_params_t", native.ConcatMlaParams),
                      ("b200_mla_decode_params_t", native.MlaDecodeParams),
                      ("b200_dcp_fold_params_t", native.DcpFoldParams),
                      ("b200_dcp_gather_params_t", native.DcpGatherParams)):
        prog.append('printf("%%zu\\n", sizeof(%s));' % cname)
        for f in fields(st):
            prog.append('printf("%%zu\\n", offsetof(%s, %s));' % (cname, f))
    prog.append("return 0;}")
    src = tmp_path / "layout.c"
    src.write_text("\n".join(prog))
    exe = tmp_path / "layout"
    subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    vals = [int(v) for v in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split()]
    it = iter(vals)
    for st in (native.PagedAttnParams, native.CacheWriteParams, native.MergeAttnStatesParams, native.GatherCacheParams,
               native.RotaryParams, native.ConcatMlaParams, native.MlaDecodeParams, native.DcpFoldParams,
               native.DcpGatherParams):
        assert next(it) == C.sizeof(st)
        for f in fields(st):
            assert next(it) == getattr(st, f).offset, f


def test_kv_dtype_strings(native):
    lib = native.load()
    assert lib.b200_kv_dtype_from_string(b"auto") == native.KV_AUTO
    assert lib.b200_kv_dtype_from_string(b"fp8") == native.KV_FP8_E4M3
    assert lib.b200_kv_dtype_from_string(b"fp8_e4m3") == native.KV_FP8_E4M3
    assert lib.b200_kv_dtype_from_string(b"fp8_e5m2") == -1  # listed by the reference enum, implemented nowhere
    assert lib.b200_kv_dtype_from_string(b"int8") == -1


def _params(native, **kw):
    p = native.PagedAttnParams()
    p.num_seqs, p.num_heads, p.num_kv_heads, p.head_size, p.block_size = 4, 32, 8, 128, 16
    p.max_seq_len, p.max_num_blocks_per_seq = 4096, 256
    p.dtype, p.kv_dtype = native.DTYPE_BF16, native.KV_AUTO
    p.kv_block_stride, p.kv_head_stride, p.q_stride = 8 * 128 * 16, 128 * 16, 32 * 128
    for k, v in kw.items():
        setattr(p, k, v)
    return p


@pytest.mark.parametrize("kw,code,msg", [
    (dict(head_size=48), 2, "Unsupported head size: 48"),
    (dict(block_size=64), 2, "Unsupported block size: 64"),
    (dict(kv_dtype=7), 2, "Unsupported data type of kv cache"),
    (dict(dtype=9), 2, "Unsupported data type"),
    (dict(num_heads=30), 1, "not a multiple"),
    (dict(blocksparse_vert_stride=8), 2, "block-sparse"),
])
def test_validation_errors_without_a_device(native, kw, code, msg):
    """Argument errors are reported before any CUDA call (reference: TORCH_CHECK(false, "Unsupported head
    size: ...") csrc/attention/paged_attention_v1.cu:121-123,155-157)."""
    p = _params(native, **kw)
    rc = native.load().b200_paged_attention_v1(C.byref(p), None)
    assert rc == code
    assert msg in native.load().b200_last_error().decode()


def test_null_pointers_are_rejected(native):
    p = _params(native)
    assert native.load().b200_paged_attention_v1(C.byref(p), None) == native.ERR_INVALID_ARGUMENT
    assert "NULL" in native.load().b200_last_error().decode()
    cp = native.CacheWriteParams()
    cp.num_tokens, cp.num_heads, cp.head_size, cp.block_size, cp.x = 4, 8, 128, 16, 8
    assert native.load().b200_reshape_and_cache(C.byref(cp), None) == native.ERR_INVALID_ARGUMENT


def test_empty_batches_are_noops(native):
    p = _params(native, num_seqs=0)
    assert native.load().b200_paged_attention_v1(C.byref(p), None) == 0
    cp = native.CacheWriteParams()
    cp.num_tokens, cp.num_heads, cp.head_size, cp.block_size, cp.x = 0, 8, 128, 16, 8
    assert native.load().b200_reshape_and_cache(C.byref(cp), None) == 0
    assert native.load().b200_reshape_and_cache_flash(C.byref(cp), None) == 0


def test_new_entry_points_validate_on_the_host(native):
    """merge_attn_states / cache movers: argument errors and empty calls need no device."""
    lib = native.load()
    mp = native.MergeAttnStatesParams()
    mp.num_tokens, mp.num_heads, mp.head_size, mp.dtype = 0, 4, 128, native.DTYPE_BF16
    assert lib.b200_merge_attn_states(C.byref(mp), None) == 0
    mp.num_tokens = 8
    assert lib.b200_merge_attn_states(C.byref(mp), None) == native.ERR_INVALID_ARGUMENT  # NULL tensors
    mp.head_size, mp.dtype = 36, native.DTYPE_F16  # not a multiple of the 8-element pack
    mp.output = mp.prefix_output = mp.suffix_output = mp.prefix_lse = mp.suffix_lse = 4096
    assert lib.b200_merge_attn_states(C.byref(mp), None) == native.ERR_UNSUPPORTED
    assert "pack_size" in lib.b200_last_error().decode()
    mp.dtype = 7
    assert lib.b200_merge_attn_states(C.byref(mp), None) == native.ERR_UNSUPPORTED
    assert "Unsupported data type of O" in lib.b200_last_error().decode()
    gp = native.GatherCacheParams()
    gp.batch_size, gp.block_size, gp.entry_size = 0, 16, 576
    assert lib.b200_gather_cache(C.byref(gp), None) == 0
    gp.batch_size = 2
    assert lib.b200_gather_cache(C.byref(gp), None) == native.ERR_INVALID_ARGUMENT
    gp.src_cache = gp.dst = gp.block_table = gp.cu_seq_lens = 4096
    gp.src_dtype, gp.dst_dtype, gp.kv_dtype = native.DTYPE_F32, native.DTYPE_F16, native.KV_AUTO
    assert lib.b200_gather_cache(C.byref(gp), None) == native.ERR_UNSUPPORTED
    assert "same dtype" in lib.b200_last_error().decode()
    rp = native.RotaryParams()
    rp.num_tokens, rp.num_heads, rp.num_kv_heads, rp.head_size, rp.rot_dim, rp.dtype = 0, 8, 2, 128, 128, native.DTYPE_BF16
    assert lib.b200_rotary_embedding(C.byref(rp), None) == 0
    rp.num_tokens = 4
    assert lib.b200_rotary_embedding(C.byref(rp), None) == native.ERR_INVALID_ARGUMENT  # NULL tensors
    rp.rot_dim = 130
    assert lib.b200_rotary_embedding(C.byref(rp), None) == native.ERR_INVALID_ARGUMENT
    assert "rot_dim" in lib.b200_last_error().decode()
    cm = native.ConcatMlaParams()
    cm.num_tokens, cm.kv_lora_rank, cm.pe_dim, cm.block_size, cm.dtype = 0, 512, 64, 16, native.DTYPE_BF16
    assert lib.b200_concat_and_cache_mla(C.byref(cm), None) == 0
    cm.num_tokens = 3
    assert lib.b200_concat_and_cache_mla(C.byref(cm), None) == native.ERR_INVALID_ARGUMENT
    assert lib.b200_copy_blocks_mla(None, 0, None, 0, 0, None) == 0
    assert lib.b200_mla_decode(None, None) == native.ERR_INVALID_ARGUMENT
    md = native.MlaDecodeParams()
    assert lib.b200_mla_decode(C.byref(md), None) == native.ERR_INVALID_ARGUMENT  # NULL tensors
    assert "null tensor" in lib.b200_last_error().decode()
    # split count left to the kernel: the flat schedule's two partial slots per worker (+ a header) -- workers are the
    # 148 CTAs with group-wide slots (<= 64 heads) or the 74 CTA pairs with 128-wide slots; an explicit count: [S][H][n][513]
    assert lib.b200_mla_decode_workspace_bytes(296, 64, 4096, 0) == 2560 + 148 * 2 * 64 * 513 * 4
    assert lib.b200_mla_decode_workspace_bytes(296, 128, 4096, 0) == 2560 + 74 * 2 * 128 * 513 * 4
    assert lib.b200_mla_decode_workspace_bytes(1, 128, 131072, 1) == 0
    assert lib.b200_mla_decode_workspace_bytes(2, 16, 4096, 4) == 2 * 16 * 4 * 513 * 4
    assert lib.b200_copy_blocks(None, None, 0, None, 0, 0, None) == 0
    assert lib.b200_swap_blocks(None, None, None, 0, 4096, native.COPY_D2D, None) == 0
    m = (C.c_int64 * 2)(0, 1)
    assert lib.b200_swap_blocks(4096, 8192, m, 1, 64, 9, None) == native.ERR_INVALID_ARGUMENT
    assert "Invalid device combination" in lib.b200_last_error().decode()


def test_dcp_exchange_fold_validates_on_the_host(native):
    import torch

    out = torch.empty(2, 4, 128, dtype=torch.bfloat16)
    with pytest.raises(native.B200Error, match="NULL buffer of rank 1") as e:
        native.dcp_exchange_fold([4096, 0], [4096, 0], [4096, 0], out, None, 0, 2, 4, 128, 1, 0)
    assert e.value.code == native.ERR_INVALID_ARGUMENT
    with pytest.raises(native.B200Error, match="epoch 0"):
        native.dcp_exchange_fold([4096], [4096], [4096], out, None, 0, 2, 4, 128, 0, 0)
    with pytest.raises(native.B200Error, match="rank 3 of 2"):
        native.dcp_exchange_fold([4096, 4096], [4096, 4096], [4096, 4096], out, None, 3, 2, 4, 128, 1, 0)
    with pytest.raises(native.B200Error, match="world size 17"):
        native.dcp_exchange_fold([4096] * 17, [4096] * 17, [4096] * 17, out, None, 0, 2, 4, 128, 1, 0)
    with pytest.raises(native.B200Error, match="not 16-byte aligned"):
        native.dcp_exchange_fold([4104], [4096], [4096], out, None, 0, 2, 4, 128, 1, 0)
    assert native.DCP_MAX_WORLD == 16
    q = torch.empty(2, 4, 128, dtype=torch.bfloat16)
    with pytest.raises(native.B200Error, match="NULL buffer of rank 1"):
        native.dcp_gather_query([4096, 0], [4096, 4096], q, 4096, 0, 1, 0)
    with pytest.raises(native.B200Error, match="epoch 0"):
        native.dcp_gather_query([4096], [4096], q, 4096, 0, 0, 0)
    with pytest.raises(native.B200Error, match="NULL pointer"):
        native.dcp_gather_query([4096], [4096], q, 0, 0, 1, 0)


def test_workspace_sizing_is_host_only_and_monotone(native):
    small = native.workspace_bytes(_params(native, max_seq_len=256, num_seqs=64))
    big = native.workspace_bytes(_params(native, max_seq_len=131072, num_seqs=8))
    assert small >= 0 and big > 0
    # layout: (D + 2) floats per (seq, head, partition) + one int per (seq, kv head)
    assert (big - 8 * 8 * 4) % ((128 + 2) * 4) == 0


def test_variant_hook(native):
    native.set_attention_variant(native.VARIANT_GENERIC)
    native.set_attention_variant(native.VARIANT_AUTO)
    with pytest.raises(native.B200Error):
        native.set_attention_variant(17)


def test_torch_shim_registers_reference_schemas_and_has_no_cpu_fallback(native):
    import vllm_metax_b200

    vllm_metax_b200.import_kernels()
    s = str(torch.ops._C.paged_attention_v1.default._schema)
    assert s.startswith("_C::paged_attention_v1(Tensor($0! -> ) out, Tensor query, Tensor key_cache, Tensor value_cache, "
                        "int num_kv_heads, float scale, Tensor block_tables, Tensor seq_lens, int block_size, "
                        "int max_seq_len, Tensor? alibi_slopes, str kv_cache_dtype, Tensor k_scale, Tensor v_scale, "
                        "int tp_rank, int blocksparse_local_blocks, int blocksparse_vert_stride, "
                        "int blocksparse_block_size, int blocksparse_head_sliding_step) -> ()")
    s2 = str(torch.ops._C.paged_attention_v2.default._schema)
    assert "Tensor($0! -> ) out, Tensor($1! -> ) exp_sums, Tensor($2! -> ) max_logits, Tensor($3! -> ) tmp_out" in s2
    for name in ("reshape_and_cache", "reshape_and_cache_flash"):
        s3 = str(getattr(torch.ops._C_cache_ops, name).default._schema)
        assert "Tensor key, Tensor value, Tensor($0! -> ) key_cache, Tensor($1! -> ) value_cache, Tensor slot_mapping, " \
               "str kv_cache_dtype, Tensor k_scale, Tensor v_scale) -> ()" in s3
    assert "convert_fp8(Tensor($0! -> ) dst_cache, Tensor src_cache, float scale, str kv_cache_dtype) -> ()" in str(
        torch.ops._C_cache_ops.convert_fp8.default._schema)
    assert str(torch.ops._C.merge_attn_states.default._schema) == (
        "_C::merge_attn_states(Tensor($0! -> ) output, Tensor($1! -> )? output_lse, Tensor prefix_output, "
        "Tensor prefix_lse, Tensor suffix_output, Tensor suffix_lse) -> ()")
    assert str(torch.ops._C_cache_ops.swap_blocks.default._schema) == (
        "_C_cache_ops::swap_blocks(Tensor src, Tensor($0! -> ) dst, Tensor block_mapping) -> ()")
    assert str(torch.ops._C_cache_ops.copy_blocks.default._schema) == (
        "_C_cache_ops::copy_blocks(Tensor(a!)[] key_caches, Tensor[](b!) value_caches, Tensor block_mapping) -> ()")
    assert str(torch.ops._C_cache_ops.gather_and_maybe_dequant_cache.default._schema) == (
        "_C_cache_ops::gather_and_maybe_dequant_cache(Tensor src_cache, Tensor($0! -> ) dst, Tensor block_table, "
        "Tensor cu_seq_lens, int batch_size, str kv_cache_dtype, Tensor scale, Tensor? seq_starts) -> ()")
    assert str(torch.ops._C_cache_ops.cp_gather_cache.default._schema) == (
        "_C_cache_ops::cp_gather_cache(Tensor src_cache, Tensor($0! -> ) dst, Tensor block_table, Tensor cu_seq_lens, "
        "int batch_size, Tensor? seq_starts) -> ()")
    assert str(torch.ops._C.rotary_embedding.default._schema) == (
        "_C::rotary_embedding(Tensor positions, Tensor($0! -> ) query, Tensor($1! -> )? key, int head_size, "
        "Tensor cos_sin_cache, bool is_neox) -> ()")
    # CPU tensors: the dispatcher must refuse (only the CUDA key is implemented, as in the reference)
    q = torch.zeros(1, 8, 128, dtype=torch.bfloat16)
    kc = torch.zeros(4, 2, 16, 16, 8, dtype=torch.bfloat16)
    vc = torch.zeros(4, 2, 128, 16, dtype=torch.bfloat16)
    one = torch.ones(1)
    with pytest.raises((NotImplementedError, RuntimeError)):
        torch.ops._C.paged_attention_v1(torch.empty_like(q), q, kc, vc, 2, 0.1, torch.zeros(1, 4, dtype=torch.int32),
                                        torch.ones(1, dtype=torch.int32), 16, 16, None, "auto", one, one, 0, 0, 0, 64, 0)


def test_product_code_never_touches_the_oracle():
    """No import, dlopen or path of oracle/ anywhere in the product package (comments may mention the word)."""
    pkg = os.path.join(ROOT, "vllm_metax_b200")
    pat = re.compile(r"(^\s*(from|import)\s+oracle\b|liboracle|oracle/|oracle\.(torch_oracle|c_oracle|e4m3_np))", re.M)
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")) and f != "build.py":
                txt = open(os.path.join(dirpath, f), errors="ignore").read()
                assert not pat.search(txt), os.path.join(dirpath, f)
