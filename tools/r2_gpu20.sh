#!/bin/bash
timeout 900 python -m pytest tests/test_rotary_gpu.py tests/test_merge_movers_gpu.py -q -m gpu > gpurun_out/r2_t20.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/r2_t20.log | cut -c1-200
timeout 300 python tools/movers_bench.py > gpurun_out/r2_movers_bench.log 2>&1; echo "movers rc=$?"; cat gpurun_out/r2_movers_bench.log | cut -c1-150 | tail -11
for c in 8 32; do echo "cover=$c"; B200_GATHER_COVER=$c timeout 300 python tools/movers_bench.py 2>&1 | grep gather | cut -c1-150; done
