#!/bin/bash
# two GPUs: DCP soak with skewed ranks (2000 steps), DCP tests, bench.py under torchrun (tp block)
B200_DCP_SOAK_STEPS=2000 timeout 900 python -m pytest tests/test_dcp_gpu.py -q -m gpu -k "soak or nccl" --timeout 800 > gpurun_out/r2_dcp_soak_n2.log 2>&1
echo "dcp rc=$?"; tail -4 gpurun_out/r2_dcp_soak_n2.log | cut -c1-200
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/r2_bench_n2.json 2> gpurun_out/r2_bench_n2.err
echo "bench rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/r2_bench_n2.json')); print(d['value'], d['n_gpus'], d['e2e']['value']); print(json.dumps(d.get('tp'), indent=0)[:1500])"
