#!/bin/bash
timeout 120 python tools/mla_debug.py > gpurun_out/r2_mla_debug.log 2>&1; echo "debug rc=$?"; tail -12 gpurun_out/r2_mla_debug.log | cut -c1-220
