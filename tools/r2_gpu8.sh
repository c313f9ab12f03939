#!/bin/bash
# MLA decode bring-up: stage check, then the parity tests
timeout 300 python tools/mla_debug.py > gpurun_out/r2_mla_debug.log 2>&1; echo "debug rc=$?"; tail -20 gpurun_out/r2_mla_debug.log
timeout 900 python -m pytest tests/test_mla_decode_gpu.py -q -m gpu -x --timeout 120 > gpurun_out/r2_mla_tests.log 2>&1
echo "tests rc=$?"; tail -15 gpurun_out/r2_mla_tests.log | cut -c1-300
