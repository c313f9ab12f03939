#!/bin/bash
# static default + dynamic test hook + strided per-warp items; ncu of the per-warp-item kernel; DCP guards
timeout 1500 python -m pytest tests/test_paged_attention_gpu.py tests/test_named_configs_gpu.py tests/test_boundary_gpu.py \
  tests/test_dcp_gpu.py tests/test_merge_movers_gpu.py -q -m gpu -x > gpurun_out/r2_t5.log 2>&1
echo "tests rc=$?"; tail -5 gpurun_out/r2_t5.log
run() { echo "== $*"; timeout 120 python tools/quick_bench.py "$@" 2>&1 | tail -1; }
{
run
run --fp8
run --fp8 --v2
run --S 8 --L 131072 --v2 --fp8
run --S 128 --H 8 --Hkv 1 --L 8192 --fp8 --v2
} > gpurun_out/r2_b5.log 2>&1
cat gpurun_out/r2_b5.log
timeout 300 ncu --set full --clock-control none --import-source on -k regex:paged_attention_stream -s 3 -c 1 \
  -o gpurun_out/r2_ncu_c5fp8_pw -f python tools/quick_bench.py --S 8 --L 131072 --v2 --fp8 --iters 3 > gpurun_out/r2_ncu_c5fp8_pw.log 2>&1
ls -la gpurun_out/r2_ncu_c5fp8_pw.ncu-rep
