#!/bin/bash
# per-warp item mode (fp8 v2): parity + timing
timeout 900 python -m pytest tests/test_paged_attention_gpu.py tests/test_named_configs_gpu.py tests/test_boundary_gpu.py -q -m gpu -x \
  > gpurun_out/r2_pw_tests.log 2>&1
echo "tests rc=$?"; tail -5 gpurun_out/r2_pw_tests.log
run() { echo "== $*"; timeout 120 python tools/quick_bench.py "$@" 2>&1 | tail -1; }
{
run --fp8 --v2
B200_STREAM_PW=-1 run --fp8 --v2
run --fp8
run --S 8 --L 131072 --v2 --fp8
B200_STREAM_PW=-1 run --S 8 --L 131072 --v2 --fp8
run --S 128 --H 8 --Hkv 1 --L 8192 --fp8 --v2
run --B 32 --fp8 --v2
run --S 8 --L 131072 --v2
} > gpurun_out/r2_pw_bench.log 2>&1
cat gpurun_out/r2_pw_bench.log
