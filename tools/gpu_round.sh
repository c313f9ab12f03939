#!/bin/bash
# One GPU-box pass: parity tests, smoke, bench (both arms), developer sweep, ncu launch list + full capture.
# Everything lands in gpurun_out/ (merged back by gpurun).
mkdir -p gpurun_out
O=gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.max.mem,power.limit --format=csv > $O/gpu.txt 2>&1
nproc > $O/nproc.txt; lscpu | grep -i "model name" >> $O/nproc.txt

echo "== pytest gpu"; timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "rc=$?" | tee -a $O/pytest_gpu.log; tail -3 $O/pytest_gpu.log
echo "== smoke"; timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "rc=$?" | tee -a $O/smoke.log; tail -3 $O/smoke.log
echo "== bench c2"; timeout 600 python bench.py > $O/bench_c2.json 2> $O/bench_c2.err; echo "rc=$?"; cat $O/bench_c2.json
echo "== bench c3"; timeout 600 python bench.py --workload c3 > $O/bench_c3.json 2> $O/bench_c3.err; echo "rc=$?"; cat $O/bench_c3.json
echo "== bench reference"; timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; echo "rc=$?"; cat $O/bench_ref.json
echo "== sweep"; timeout 900 bash tools/sweep_r1.sh > $O/sweep.log 2>&1; cat $O/sweep.log
echo "== ncu launches"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_c2.csv \
   python bench.py --steps 20 --warmup 3 --no-cpu-baseline > $O/ncu_launch.log 2>&1; echo "rc=$?"
echo "== ncu full"
for w in "" "--fp8"; do
  tag=c2; [ -n "$w" ] && tag=c3
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:paged_attention_stream -s 6 -c 1 \
     -o $O/ncu_${tag}_stream -f python tools/quick_bench.py $w --iters 3 > $O/ncu_full_$tag.log 2>&1; echo "rc=$?"
done
ls -la $O
