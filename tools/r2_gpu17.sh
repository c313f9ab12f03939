#!/bin/bash
timeout 900 python -m pytest tests/test_mla_decode_gpu.py tests/test_upstream_witness.py -q -m gpu --timeout 120 -x > gpurun_out/r2_mla_tests.log 2>&1
echo "tests rc=$?"; tail -4 gpurun_out/r2_mla_tests.log | cut -c1-300
for only in 0 1 2 3; do timeout 200 python tools/mla_bench.py --impl ours --only $only --out gpurun_out/mla_tmp 2>&1 | cut -c1-170; done
timeout 100 python tools/mla_bench.py --compare gpurun_out/mla_tmp.pt gpurun_out/mla_upstream.pt 2>&1 | tail -3
timeout 300 python tools/mla_debug.py --timeline 128 128 8192 > gpurun_out/r2_mla_timeline_pair.log 2>&1; head -14 gpurun_out/r2_mla_timeline_pair.log | cut -c1-120
