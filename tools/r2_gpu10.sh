#!/bin/bash
timeout 900 python -m pytest tests/test_mla_decode_gpu.py -q -m gpu --timeout 120 > gpurun_out/r2_mla_tests.log 2>&1
echo "tests rc=$?"; tail -5 gpurun_out/r2_mla_tests.log | cut -c1-300
timeout 600 python tools/mla_bench.py --impl ours --out gpurun_out/mla_ours > gpurun_out/r2_mla_bench_ours.log 2>&1; echo "ours rc=$?"; cat gpurun_out/r2_mla_bench_ours.log | cut -c1-300
timeout 100 python tools/mla_bench.py --compare gpurun_out/mla_ours.pt gpurun_out/mla_upstream.pt 2>&1 | grep out_ | head -4
