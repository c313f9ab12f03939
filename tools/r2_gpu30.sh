#!/bin/bash
echo "flat:"; for only in 1; do timeout 200 python tools/mla_bench.py --impl ours --only $only --out gpurun_out/mla_tmp 2>&1 | cut -c1-150; done
echo "flat off:"; for only in 1; do B200_MLA_FLAT_OFF=1 timeout 200 python tools/mla_bench.py --impl ours --only $only --out gpurun_out/mla_tmp 2>&1 | cut -c1-150; done
echo "varlen:"; for only in 1; do timeout 200 python tools/mla_bench.py --impl ours --varlen --only $only --out gpurun_out/mla_tmp 2>&1 | cut -c1-150; done
echo "old lib:"; B200_PAGED_ATTN_LIB=$PWD/build/old_lib/libb200_paged_attn.so timeout 200 python tools/mla_bench.py --impl ours --only 1 --out gpurun_out/mla_tmp 2>&1 | cut -c1-150
