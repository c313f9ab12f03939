#!/bin/bash
# final MLA evidence: both implementations on the same box (uniform and ragged), ncu --set full of the two kernels
timeout 600 python tools/mla_bench.py --impl ours --out gpurun_out/mla_ours > gpurun_out/r2_mla_v4_bench_ours.log 2>&1; cut -c1-200 gpurun_out/r2_mla_v4_bench_ours.log
timeout 600 python tools/mla_bench.py --impl ours --varlen --out gpurun_out/mla_ours_varlen > gpurun_out/r2_mla_v4_bench_ours_varlen.log 2>&1; cut -c1-200 gpurun_out/r2_mla_v4_bench_ours_varlen.log
timeout 900 python tools/mla_bench.py --impl upstream --out gpurun_out/mla_upstream > gpurun_out/r2_mla_v4_bench_upstream.log 2>&1; cut -c1-200 gpurun_out/r2_mla_v4_bench_upstream.log
timeout 900 python tools/mla_bench.py --impl upstream --varlen --out gpurun_out/mla_upstream_varlen > gpurun_out/r2_mla_v4_bench_upstream_varlen.log 2>&1; cut -c1-200 gpurun_out/r2_mla_v4_bench_upstream_varlen.log
timeout 100 python tools/mla_bench.py --compare gpurun_out/mla_ours.pt gpurun_out/mla_upstream.pt 2>&1 | grep out_ | head -8
timeout 100 python tools/mla_bench.py --compare gpurun_out/mla_ours_varlen.pt gpurun_out/mla_upstream_varlen.pt 2>&1 | grep out_ | head -8
for only in 5 1; do
timeout 400 ncu --set full --clock-control none --import-source on -k regex:mla_decode --launch-skip 3 -c 1 -o gpurun_out/r2_mla_ncu_full_$only -f python tools/mla_bench.py --impl ours --only $only --reps 2 --out gpurun_out/mla_tmp > gpurun_out/r2_mla_ncu_full_$only.log 2>&1; echo "ncu $only rc=$?"
done
ls -la gpurun_out/*.ncu-rep | tail -3
