#!/bin/bash
timeout 900 python -m pytest tests/test_mla_decode_gpu.py tests/test_upstream_witness.py -q -m gpu --timeout 120 > gpurun_out/r2_mla_tests.log 2>&1
echo "tests rc=$?"; tail -3 gpurun_out/r2_mla_tests.log | cut -c1-300
echo "flat:"; for only in 0 1 2 3 4 5 6 7; do timeout 200 python tools/mla_bench.py --impl ours --only $only --out gpurun_out/mla_tmp 2>&1 | cut -c1-150; done
echo "flat off:"; for only in 5 6; do B200_MLA_FLAT_OFF=1 timeout 200 python tools/mla_bench.py --impl ours --only $only --out gpurun_out/mla_tmp 2>&1 | cut -c1-150; done
echo "varlen:"; for only in 1 5 6; do timeout 200 python tools/mla_bench.py --impl ours --varlen --only $only --out gpurun_out/mla_tmp 2>&1 | cut -c1-150; done
timeout 200 python tools/mla_debug.py --cta-times 128 64 8192 2>&1 | tail -3 | cut -c1-200
