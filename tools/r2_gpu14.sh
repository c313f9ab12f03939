#!/bin/bash
for sp in 1 2 3 4 8; do
timeout 200 python tools/mla_bench.py --impl ours --only 1 --splits $sp --out gpurun_out/mla_tmp 2>&1 | cut -c1-150 | sed "s/^/splits=$sp /"
done
for only in 3 7; do timeout 200 python tools/mla_bench.py --impl ours --only $only --out gpurun_out/mla_tmp 2>&1 | cut -c1-150; done
