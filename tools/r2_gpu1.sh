#!/bin/bash
# Round 2, first GPU call: upstream witness (own process), the new parity / boundary tests, the default bench line,
# launch list + ncu --set full of the C1 and C4-TP8 shapes.
mkdir -p gpurun_out/witness
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv > gpurun_out/r2_gpu.txt 2>&1
timeout 900 python tools/upstream_witness.py --out gpurun_out/witness > gpurun_out/witness/stdout.log 2> gpurun_out/witness/stderr.log
echo "witness rc=$?"; tail -c 1500 gpurun_out/witness/stdout.log
mkdir -p tests/golden && cp gpurun_out/witness/fp8_upstream_*.pt tests/golden/ 2>/dev/null
timeout 1500 python -m pytest tests/test_boundary_gpu.py tests/test_named_configs_gpu.py tests/test_upstream_witness.py -q -m gpu \
  > gpurun_out/r2_new_tests.log 2>&1
echo "new tests rc=$?"; tail -40 gpurun_out/r2_new_tests.log
timeout 600 python bench.py > gpurun_out/r2_bench_c2.json 2> gpurun_out/r2_bench_c2.err
echo "bench rc=$?"; cat gpurun_out/r2_bench_c2.json; tail -5 gpurun_out/r2_bench_c2.err
timeout 400 bash tools/sweep_r1.sh > gpurun_out/r2_sweep0.log 2>&1
cat gpurun_out/r2_sweep0.log
# ncu --set full of the two small-work shapes (one launch each)
timeout 300 ncu --set full --clock-control none --import-source on -k regex:paged_attention_stream -s 3 -c 1 \
  -o gpurun_out/r2_ncu_c1 -f python tools/quick_bench.py --S 8 --H 32 --Hkv 32 --L 1024 --iters 3 > gpurun_out/r2_ncu_c1.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:paged_attention_stream -s 3 -c 1 \
  -o gpurun_out/r2_ncu_c4tp8 -f python tools/quick_bench.py --S 128 --H 8 --Hkv 1 --L 8192 --iters 3 > gpurun_out/r2_ncu_c4tp8.log 2>&1
ls -la gpurun_out/*.ncu-rep
