#!/bin/bash
timeout 120 python tools/mla_debug.py > gpurun_out/r2_mla_debug.log 2>&1; echo "debug rc=$?"; grep "H128" gpurun_out/r2_mla_debug.log | cut -c1-170
timeout 900 python -m pytest tests/test_mla_decode_gpu.py tests/test_upstream_witness.py -q -m gpu --timeout 120 > gpurun_out/r2_mla_tests.log 2>&1
echo "tests rc=$?"; tail -5 gpurun_out/r2_mla_tests.log | cut -c1-300
echo "flat:"; for only in 0 1 2 3; do timeout 200 python tools/mla_bench.py --impl ours --only $only --out gpurun_out/mla_tmp 2>&1 | cut -c1-150; done
echo "flat off:"; for only in 1; do B200_MLA_FLAT_OFF=1 timeout 200 python tools/mla_bench.py --impl ours --only $only --out gpurun_out/mla_tmp 2>&1 | cut -c1-150; done
echo "varlen:"; for only in 1; do timeout 200 python tools/mla_bench.py --impl ours --varlen --only $only --out gpurun_out/mla_tmp 2>&1 | cut -c1-150; done
