#!/bin/bash
timeout 900 python -m pytest tests/test_mla_decode_gpu.py -q -m gpu --timeout 120 -x > gpurun_out/r2_mla_tests.log 2>&1
echo "tests rc=$?"; tail -3 gpurun_out/r2_mla_tests.log | cut -c1-300
timeout 300 python tools/mla_debug.py --timeline > gpurun_out/r2_mla_timeline.log 2>&1; grep -A8 "MMA lane" gpurun_out/r2_mla_timeline.log | cut -c1-200
timeout 600 python tools/mla_bench.py --impl ours --out gpurun_out/mla_ours > gpurun_out/r2_mla_bench_ours.log 2>&1; echo "ours rc=$?"; cut -c1-180 gpurun_out/r2_mla_bench_ours.log
