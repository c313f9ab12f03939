#!/bin/bash
# v1 backend under the reference's test template; partition-parallel split reduce
timeout 900 python -m pytest tests/test_v1_backend.py -q -m gpu > gpurun_out/r2_t6_backend.log 2>&1
echo "backend rc=$?"; tail -15 gpurun_out/r2_t6_backend.log
timeout 1500 python -m pytest tests/test_paged_attention_gpu.py tests/test_named_configs_gpu.py tests/test_dcp_gpu.py -q -m gpu -x > gpurun_out/r2_t6.log 2>&1
echo "tests rc=$?"; tail -4 gpurun_out/r2_t6.log
run() { echo "== $*"; timeout 120 python tools/quick_bench.py "$@" 2>&1 | tail -1; }
{
run --S 8 --L 131072 --v2
run --S 8 --L 131072 --v2 --fp8
run --v2
run --fp8 --v2
} > gpurun_out/r2_b6.log 2>&1
cat gpurun_out/r2_b6.log
