#!/bin/bash
for mode in 0 1; do
B200_MLA_FLAT_OFF=$mode timeout 300 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,lts__t_sectors_srcunit_tex_op_read.sum,sm__cycles_elapsed.avg.per_second,sm__cycles_active.avg,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active \
  --clock-control none -k regex:mla_decode_kernel -c 2 --csv --log-file gpurun_out/r2_mla_ncu_flat$mode.csv python tools/mla_bench.py --impl ours --only 6 --reps 1 --out gpurun_out/mla_tmp > /dev/null 2>&1
echo "flat_off=$mode"; python - <<PY
import csv
rows=[r for r in csv.reader(l for l in open('gpurun_out/r2_mla_ncu_flat$mode.csv') if not l.startswith('=='))]
h=rows[0]; i=h.index('ID'); n=h.index('Metric Name'); v=h.index('Metric Value')
for r in rows[1:]:
    if r[i]=='1': print('  ', r[n], r[v])
PY
done
