#!/bin/bash
# eight GPUs: bench.py under torchrun (tp block at TP8), DCP soak at world 8 / 4
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 8 --steps 20 --warmup 3 > gpurun_out/r2_bench_n8.json 2> gpurun_out/r2_bench_n8.err
echo "bench rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/r2_bench_n8.json')); print(d['value'], d['n_gpus'], d['e2e']['value']);
for k in ('c4','c4_fp8'): t=d['tp'][k]; print(k, t['shard_attention_us'], t['aggregate_gbs'], t['strong_scaling_efficiency'], t['parity_ok'])"
B200_DCP_SOAK_STEPS=1000 timeout 900 python -m pytest tests/test_dcp_gpu.py -q -m gpu -k "soak" --timeout 800 > gpurun_out/r2_dcp_soak_n8.log 2>&1
echo "dcp rc=$?"; tail -3 gpurun_out/r2_dcp_soak_n8.log | cut -c1-200
timeout 300 python tools/dcp_bench.py > gpurun_out/r2_dcp_bench_n8.log 2>&1; tail -6 gpurun_out/r2_dcp_bench_n8.log | cut -c1-200
