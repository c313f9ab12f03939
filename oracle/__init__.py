"""CPU oracle of the paged-attention decode path — TEST INFRASTRUCTURE ONLY.

Nothing under oracle/ is product code.  Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / `--impl reference` leg may import it, and only as the checker or the timed CPU
baseline.  The product path (vllm_metax_b200/) never imports it and has no CPU fallback.

Parity status: PINNED for attention and the cache writers (auto dtype) against golden vectors
produced by the reference's own Python test oracles run in the build container
(tests/golden/make_golden.py imports /root/reference/tests/kernels/attention/test_attention.py and
test_cache.py's scatter loops).  FP8-e4m3 *bytes* are UNPINNED by the reference: its conversion is an
assert(false) stub (csrc/quantization/fp8/metax/quant_utils.cuh:13-23) and its tests only compare
dequantised values with rtol 0.1; the byte-level oracle here restates the semantics of the surviving
call sites (see torch_oracle.fp8_e4m3_quant) and is cross-checked against an independent integer
encoder (e4m3_np.py).
"""
