#!/bin/bash
# dynamic tail region of the flat schedule + per-warp-item producer lookahead: parity, then timing with / without
timeout 1500 python -m pytest tests/test_paged_attention_gpu.py tests/test_named_configs_gpu.py tests/test_boundary_gpu.py \
  tests/test_flash_decode_gpu.py tests/test_merge_movers_gpu.py tests/test_dcp_gpu.py -q -m gpu -x > gpurun_out/r2_dyn_tests.log 2>&1
echo "tests rc=$?"; tail -5 gpurun_out/r2_dyn_tests.log
run() { echo "== $*"; timeout 120 python tools/quick_bench.py "$@" 2>&1 | tail -1; }
{
for dyn in 0 -1 2 8 16; do
  echo "---- B200_STREAM_DYN=$dyn"; export B200_STREAM_DYN=$dyn
  run
  run --fp8
  run --S 8 --H 32 --Hkv 32 --L 1024
  run --S 128 --H 8 --Hkv 1 --L 8192
  run --S 128 --H 8 --Hkv 1 --L 8192 --fp8
  run --S 128 --H 64 --Hkv 8 --L 8192
  run --flash
done
unset B200_STREAM_DYN
echo "---- per-warp items"
run --fp8 --v2
run --S 8 --L 131072 --v2 --fp8
B200_STREAM_NSTAGES=16 run --S 8 --L 131072 --v2 --fp8
} > gpurun_out/r2_dyn_bench.log 2>&1
cat gpurun_out/r2_dyn_bench.log
