#!/bin/bash
# ncu: DRAM bytes and L2 hit rate of the MLA kernel, H=16 and H=128 at b128/L8192
for only in 5 1; do
timeout 300 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,lts__t_sectors_srcunit_tex_op_read.sum,lts__t_bytes.sum,sm__inst_executed_pipe_tensor.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,smsp__cycles_active.avg \
  --clock-control none -k regex:mla_decode -c 2 --csv --log-file gpurun_out/r2_mla_ncu_$only.csv python tools/mla_bench.py --impl ours --only $only --reps 1 --out gpurun_out/mla_tmp > gpurun_out/r2_mla_ncu_$only.log 2>&1
echo "ncu rc=$?"; grep -v "^==" gpurun_out/r2_mla_ncu_$only.csv | cut -d, -f5,13,15 | tail -12
done
