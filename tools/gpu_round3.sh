#!/bin/bash
# Final GPU pass of round 1: full parity suite, smoke, the three bench workloads + reference arm.
mkdir -p gpurun_out
O=gpurun_out
echo "== pytest gpu"; timeout -s KILL 1500 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "rc=$?" | tee -a $O/pytest_gpu.log; tail -4 $O/pytest_gpu.log
echo "== smoke"; timeout -s KILL 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "rc=$?" | tee -a $O/smoke.log; tail -4 $O/smoke.log
for w in c2 c3 c2_flash; do
  echo "== bench $w"; timeout -s KILL 600 python bench.py --workload $w > $O/bench_$w.json 2> $O/bench_$w.err; echo "rc=$?"; cut -c1-250 $O/bench_$w.json
done
echo "== bench reference"; timeout -s KILL 600 python bench.py --impl reference --steps 5 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; echo "rc=$?"; cut -c1-200 $O/bench_ref.json
