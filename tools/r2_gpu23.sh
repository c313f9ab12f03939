#!/bin/bash
timeout 120 python tools/mla_debug.py > gpurun_out/r2_mla_debug.log 2>&1; echo "debug rc=$?"; tail -8 gpurun_out/r2_mla_debug.log | cut -c1-170
timeout 900 python -m pytest tests/test_mla_decode_gpu.py tests/test_upstream_witness.py -q -m gpu --timeout 120 -x > gpurun_out/r2_mla_tests.log 2>&1
echo "tests rc=$?"; tail -5 gpurun_out/r2_mla_tests.log | cut -c1-300
for only in 4 5 6 7; do timeout 200 python tools/mla_bench.py --impl ours --only $only --out gpurun_out/mla_tmp 2>&1 | cut -c1-150; done
echo "flat off:"; for only in 5 6; do B200_MLA_FLAT_OFF=1 timeout 200 python tools/mla_bench.py --impl ours --only $only --out gpurun_out/mla_tmp 2>&1 | cut -c1-150; done
echo "varlen:"; for only in 1 5 6; do timeout 200 python tools/mla_bench.py --impl ours --varlen --only $only --out gpurun_out/mla_tmp 2>&1 | cut -c1-150; done
echo "varlen flat off:"; for only in 5 6; do B200_MLA_FLAT_OFF=1 timeout 200 python tools/mla_bench.py --impl ours --varlen --only $only --out gpurun_out/mla_tmp 2>&1 | cut -c1-150; done
echo "varlen upstream:"; timeout 300 python tools/mla_bench.py --impl upstream --varlen --only 1 --out gpurun_out/mla_tmp2 2>&1 | cut -c1-150
