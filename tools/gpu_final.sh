#!/bin/bash
# Round-end check on one B200: DCP launch breakdown at N=1, the whole GPU parity suite, smoke(), the default bench line.
mkdir -p gpurun_out
timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node 1 --master-addr 127.0.0.1 --master-port 29511 \
  tools/dcp_bench.py 2>gpurun_out/dcp_breakdown.err > gpurun_out/dcp_breakdown.log
cat gpurun_out/dcp_breakdown.log
timeout 500 python -m pytest tests -q -m gpu -x > gpurun_out/final_gpu_tests.log 2>&1
echo "rc=$?" >> gpurun_out/final_gpu_tests.log
tail -4 gpurun_out/final_gpu_tests.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/final_smoke.log 2>&1
tail -2 gpurun_out/final_smoke.log
timeout 200 python bench.py > gpurun_out/final_bench_c2.json 2>gpurun_out/final_bench_c2.err
cat gpurun_out/final_bench_c2.json
