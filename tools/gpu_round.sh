#!/bin/bash
# One GPU-box visit, stage by stage (run through gpurun; every stage writes its log under gpurun_out/):
#   bash tools/gpu_round.sh tests        whole -m gpu suite + smoke()
#   bash tools/gpu_round.sh bench        default bench.py line (+ --impl reference)
#   bash tools/gpu_round.sh mla          MLA decode: this repository's kernels and the stock kernel, uniform and ragged
#   bash tools/gpu_round.sh movers       tools/movers_bench.py
#   bash tools/gpu_round.sh witness      tools/upstream_witness.py + the MLA witness (stock kernels' outputs)
#   bash tools/gpu_round.sh ncu          launch list of bench.py and ncu --set full of the headline kernel and the MLA kernels
#   bash tools/gpu_round.sh multi N      (gpurun --gpus N) bench.py under torchrun, DCP tests + soak
# Stages can be combined: bash tools/gpu_round.sh tests bench mla
set -u
mkdir -p gpurun_out
while [ $# -gt 0 ]; do
  stage=$1; shift
  case $stage in
    tests)
      timeout 2400 python -m pytest tests -q -m gpu > gpurun_out/gpu_tests.log 2>&1; echo "tests rc=$?"; tail -4 gpurun_out/gpu_tests.log | cut -c1-200
      timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1; tail -2 gpurun_out/smoke.log ;;
    bench)
      timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cut -c1-400 gpurun_out/bench.json
      timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_reference.json 2>&1; cut -c1-300 gpurun_out/bench_reference.json ;;
    mla)
      for v in "" "--varlen"; do
        timeout 600 python tools/mla_bench.py --impl ours $v --out gpurun_out/mla_ours$v > gpurun_out/mla_bench_ours$v.log 2>&1; cut -c1-160 gpurun_out/mla_bench_ours$v.log
        timeout 900 python tools/mla_bench.py --impl upstream $v --out gpurun_out/mla_upstream$v > gpurun_out/mla_bench_upstream$v.log 2>&1; cut -c1-160 gpurun_out/mla_bench_upstream$v.log
        timeout 100 python tools/mla_bench.py --compare gpurun_out/mla_ours$v.pt gpurun_out/mla_upstream$v.pt 2>&1 | grep out_
      done ;;
    movers)
      timeout 300 python tools/movers_bench.py > gpurun_out/movers_bench.log 2>&1; cat gpurun_out/movers_bench.log | cut -c1-160 ;;
    witness)
      timeout 600 python tools/upstream_witness.py --out gpurun_out/witness 2>&1 | tail -5
      timeout 300 python tools/mla_bench.py --impl upstream --witness gpurun_out/mla_upstream_bf16_h128.pt 2>&1 | tail -2 ;;
    ncu)
      timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 20 --warmup 3 --no-extras > gpurun_out/ncu_launches.log 2>&1
      timeout 600 ncu --set full --clock-control none --import-source on -k regex:paged_attention_stream -c 1 --launch-skip 5 -o gpurun_out/ncu_c2_stream -f python bench.py --steps 5 --warmup 3 --no-extras > gpurun_out/ncu_c2.log 2>&1
      for only in 5 1; do
        timeout 400 ncu --set full --clock-control none --import-source on -k regex:mla_decode --launch-skip 3 -c 1 -o gpurun_out/ncu_mla_$only -f python tools/mla_bench.py --impl ours --only $only --reps 2 --out gpurun_out/mla_tmp > gpurun_out/ncu_mla_$only.log 2>&1
      done
      ls -la gpurun_out/*.ncu-rep ;;
    multi)
      n=$1; shift
      timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $n --steps 20 --warmup 3 > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err; echo "bench rc=$?"; cut -c1-300 gpurun_out/bench_n$n.json
      B200_DCP_SOAK_STEPS=1000 timeout 900 python -m pytest tests/test_dcp_gpu.py -q -m gpu --timeout 800 > gpurun_out/dcp_n$n.log 2>&1; echo "dcp rc=$?"; tail -3 gpurun_out/dcp_n$n.log ;;
    *) echo "unknown stage $stage" ;;
  esac
done
