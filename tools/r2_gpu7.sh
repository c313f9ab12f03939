#!/bin/bash
# whole GPU suite (with the v1 backend tests), smoke, default bench line
timeout 1800 python -m pytest tests -q -m gpu > gpurun_out/r2_t7.log 2>&1
echo "tests rc=$?"; tail -12 gpurun_out/r2_t7.log | cut -c1-300
timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r2_smoke7.log 2>&1; tail -3 gpurun_out/r2_smoke7.log
timeout 600 python bench.py > gpurun_out/r2_bench7.json 2> gpurun_out/r2_bench7.err
echo "bench rc=$?"; cat gpurun_out/r2_bench7.json; tail -3 gpurun_out/r2_bench7.err
