#!/bin/bash
timeout 900 python -m pytest tests/test_mla_decode_gpu.py tests/test_upstream_witness.py -q -m gpu --timeout 120 > gpurun_out/r2_mla_tests.log 2>&1
echo "tests rc=$?"; tail -5 gpurun_out/r2_mla_tests.log | cut -c1-300
B200_MLA_FLAT_OFF=1 timeout 300 python tools/mla_debug.py --timeline 128 64 8192 2>&1 | head -12 | cut -c1-150
timeout 300 python tools/mla_debug.py --timeline 128 64 8192 2>&1 | head -12 | cut -c1-150
