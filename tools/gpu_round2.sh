#!/bin/bash
# Second GPU pass of round 1: full parity suite, flash-layout bench + ncu, movers bench.
mkdir -p gpurun_out
O=gpurun_out
echo "== pytest gpu"; timeout -s KILL 1200 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "rc=$?" | tee -a $O/pytest_gpu.log; tail -5 $O/pytest_gpu.log
echo "== smoke"; timeout -s KILL 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "rc=$?" | tee -a $O/smoke.log; tail -4 $O/smoke.log
echo "== bench c2_flash"; timeout -s KILL 600 python bench.py --workload c2_flash > $O/bench_c2_flash.json 2> $O/bench_c2_flash.err; echo "rc=$?"; cat $O/bench_c2_flash.json
echo "== movers"; timeout -s KILL 300 python tools/movers_bench.py > $O/movers_bench.log 2>&1; cat $O/movers_bench.log
echo "== ncu launches (flash)"
timeout -s KILL 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_c2_flash.csv \
   python bench.py --workload c2_flash --steps 20 --warmup 3 --no-cpu-baseline > $O/ncu_launch_flash.log 2>&1; echo "rc=$?"
echo "== ncu full (flash)"
timeout -s KILL 600 ncu --set full --clock-control none --import-source on -k regex:paged_attention_stream -s 6 -c 1 \
   -o $O/ncu_c2_flash_stream -f python tools/quick_bench.py --flash --iters 3 > $O/ncu_full_c2_flash.log 2>&1; echo "rc=$?"
ls -la $O
