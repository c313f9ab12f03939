#!/bin/bash
for occ in 0 1; do for only in 4 5 7; do
B200_MLA_OCC2=$occ timeout 200 python tools/mla_bench.py --impl ours --only $only --out gpurun_out/mla_tmp 2>&1 | cut -c1-150 | sed "s/^/occ2=$occ /"
done; done
B200_MLA_OCC2=1 timeout 600 python -m pytest tests/test_mla_decode_gpu.py -q -m gpu --timeout 120 -x 2>&1 | tail -2
