"""b200_paged_attention_v1_static: the v1 entry point whose split-KV scratch is a static device array inside the
library, for callers that cannot allocate (include/b200_paged_attn.h).  It has no counterpart in the reference; the
checker here is the existing entry point b200_paged_attention_v1 (caller workspace) on the same inputs, plus the CPU
oracle, at the tolerance the v1 tests use (tests/test_paged_attention_gpu.py).

Collected last on purpose.  Every test calls through the C ABI (ctypes, raw
pointers) and passes alone in a fresh process as well as after the rest of the suite: the array is shared, so each call
re-zeroes the arrival counters itself and nothing depends on what earlier tests left in it.
"""
import pytest
import torch

from oracle import torch_oracle as orc

pytestmark = pytest.mark.gpu

ATOL = 1e-3  # tests/test_paged_attention_gpu.py ATOL_AUTO (reference: tests/kernels/attention/test_attention.py:342)
ATOL_FP8 = 1e-2


def _setup(ops, S, H, Hkv, D, block_size, max_seq_len, kv_cache_dtype="auto", seed=0, num_blocks=2048):
    from vllm_metax_b200 import _native

    d = orc.make_decode_inputs(S, H, Hkv, D, block_size, max_seq_len=max_seq_len, num_blocks=num_blocks, seed=seed,
                               kv_cache_dtype=kv_cache_dtype)
    t = {k: (v.cuda() if isinstance(v, torch.Tensor) else v) for k, v in d.items()}
    one = torch.ones(1, dtype=torch.float32, device="cuda")
    t["_one"] = one

    def params(out, workspace=None):
        return _native.attn_params(out, t["query"], t["key_cache"], t["value_cache"], Hkv, d["scale"], t["block_tables"],
                                   t["seq_lens"], block_size, d["max_seq_len"], None, kv_cache_dtype, one, one,
                                   workspace=workspace)

    return d, t, params


def _existing(params, like):
    """The existing entry point with a caller-allocated, zero-filled workspace."""
    from vllm_metax_b200 import _native

    out = torch.full_like(like, float("nan"))
    p = params(out)
    ws = torch.zeros(max(16, _native.workspace_bytes(p)), dtype=torch.uint8, device="cuda")
    p = params(out, ws)
    _native.paged_attention_v1(p, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    return out


def _static(params, like):
    from vllm_metax_b200 import _native

    out = torch.full_like(like, float("nan"))
    p = params(out)  # workspace NULL / 0: ignored by the static entry point
    _native.paged_attention_v1_static(p, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    return out


def _oracle(d):
    return orc.paged_attention(d["query"], d["key_cache"], d["value_cache"], d["block_tables"], d["seq_lens"], d["scale"],
                               d["alibi_slopes"], d["kv_cache_dtype"], 1.0, 1.0)


SHAPES = [
    # S, H, Hkv, D, block, max_len, kv dtype          which kernel / what scratch
    (6, 32, 8, 128, 16, 2500, "auto"),   # streaming kernel, flat schedule: pairs split across CTAs, counters + slots
    (3, 32, 8, 128, 16, 9000, "fp8"),    # fp8 cache, long ragged rows
    (5, 40, 40, 80, 16, 1700, "auto"),   # generic kernel: 512-token partitions in the region behind the flat workspace
]


@pytest.mark.parametrize("shape", SHAPES, ids=lambda s: "S%d_H%d_%d_D%d_b%d_L%d_%s" % s)
def test_two_calls_in_a_row_match_existing_entry_point(ops, shape):
    S, H, Hkv, D, bs, L, kvd = shape
    atol = ATOL_FP8 if kvd == "fp8" else ATOL
    # two DIFFERENT batches back to back: the second call finds whatever the first left in the shared array
    for seed in (31, 32):
        d, t, params = _setup(ops, S, H, Hkv, D, bs, L, kvd, seed=seed)
        want = _existing(params, t["query"])
        got = _static(params, t["query"])
        assert torch.isfinite(got.float()).all()
        torch.testing.assert_close(got.float(), want.float(), atol=atol, rtol=1e-5)
        torch.testing.assert_close(got.float().cpu(), _oracle(d).float(), atol=atol, rtol=1e-5)
    # and the same batch again: identical bits to its own previous run (the kernels are deterministic)
    again = _static(params, t["query"])
    assert torch.equal(again, got)


def test_shapes_alternate_over_the_shared_array(ops):
    """Different kernels and shapes interleaved on one stream, no synchronisation between the launches."""
    from vllm_metax_b200 import _native

    stream = torch.cuda.current_stream().cuda_stream
    cases = []
    for i, (S, H, Hkv, D, bs, L, kvd) in enumerate(SHAPES):
        d, t, params = _setup(ops, S, H, Hkv, D, bs, L, kvd, seed=50 + i)
        cases.append((t, params, _existing(params, t["query"]), ATOL_FP8 if kvd == "fp8" else ATOL))
    outs = []
    for rep in range(3):
        for t, params, _, _ in cases:
            out = torch.full_like(t["query"], float("nan"))
            _native.paged_attention_v1_static(params(out), stream)  # stream order is the only ordering
            outs.append(out)
    torch.cuda.synchronize()
    for i, out in enumerate(outs):
        _, _, want, atol = cases[i % len(cases)]
        torch.testing.assert_close(out.float(), want.float(), atol=atol, rtol=1e-5)


def test_call_allocates_no_device_memory(ops):
    from vllm_metax_b200 import _native

    S, H, Hkv, D, bs, L, kvd = SHAPES[0]
    d, t, params = _setup(ops, S, H, Hkv, D, bs, L, kvd, seed=77)
    want = _existing(params, t["query"])
    out = torch.full_like(t["query"], float("nan"))
    p = params(out)
    stream = torch.cuda.current_stream().cuda_stream
    _native.paged_attention_v1_static(p, stream)  # first use on this device: module load, symbol look-up
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    for _ in range(4):
        _native.paged_attention_v1_static(p, stream)
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    # a workspace allocation would be ~26.6 MB; allow one 2 MiB driver page of noise
    assert free0 - free1 <= (2 << 20), (free0, free1)
    torch.testing.assert_close(out.float(), want.float(), atol=ATOL, rtol=1e-5)


def test_graph_capture_and_replay(ops):
    """The case the entry point exists for: captured with no workspace to reserve, replayed on new data."""
    from vllm_metax_b200 import _native

    S, H, Hkv, D, bs, L, kvd = SHAPES[0]
    d, t, params = _setup(ops, S, H, Hkv, D, bs, L, kvd, seed=91)
    out = torch.full_like(t["query"], float("nan"))
    p = params(out)
    _native.paged_attention_v1_static(p, torch.cuda.current_stream().cuda_stream)  # warm-up outside capture
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        _native.paged_attention_v1_static(p, torch.cuda.current_stream().cuda_stream)
    for seed in (92, 93):
        d2 = orc.make_decode_inputs(S, H, Hkv, D, bs, max_seq_len=L, num_blocks=2048, seed=seed, kv_cache_dtype=kvd)
        for k in ("query", "key_cache", "value_cache", "block_tables", "seq_lens"):
            t[k].copy_(d2[k])
        want = _existing(params, t["query"])
        out.fill_(float("nan"))
        g.replay()
        torch.cuda.synchronize()
        torch.testing.assert_close(out.float(), want.float(), atol=ATOL, rtol=1e-5)


def test_rejects_shapes_beyond_the_static_array(ops):
    from vllm_metax_b200 import _native

    limit = _native.static_scratch_bytes()
    assert limit == 32 << 20
    # generic kernel (head size 80) with enough (seq, head, partition) partials to pass the limit; nothing is launched,
    # so the tensors only have to carry the shapes
    S, H, D, L = 64, 40, 80, 8192
    q = torch.zeros(S, H, D, dtype=torch.bfloat16, device="cuda")
    kc = torch.zeros(4, H, D // 8, 16, 8, dtype=torch.bfloat16, device="cuda")
    vc = torch.zeros(4, H, D, 16, dtype=torch.bfloat16, device="cuda")
    bt = torch.zeros(S, L // 16, dtype=torch.int32, device="cuda")
    sl = torch.full((S,), L, dtype=torch.int32, device="cuda")
    one = torch.ones(1, dtype=torch.float32, device="cuda")
    out = torch.zeros_like(q)
    p = _native.attn_params(out, q, kc, vc, H, 0.1, bt, sl, 16, L, None, "auto", one, one)
    assert _native.workspace_bytes(p) > limit
    with pytest.raises(_native.B200Error, match="static array holds"):
        _native.paged_attention_v1_static(p, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert torch.count_nonzero(out).item() == 0
