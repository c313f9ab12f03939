#!/bin/bash
timeout 120 python tools/mla_debug.py > gpurun_out/r2_mla_debug.log 2>&1; echo "debug rc=$?"; grep H128 gpurun_out/r2_mla_debug.log | cut -c1-200
timeout 900 python -m pytest tests/test_mla_decode_gpu.py tests/test_upstream_witness.py -q -m gpu --timeout 120 > gpurun_out/r2_mla_tests.log 2>&1
echo "tests rc=$?"; tail -6 gpurun_out/r2_mla_tests.log | cut -c1-300
for only in 0 1 2 3; do timeout 200 python tools/mla_bench.py --impl ours --only $only --out gpurun_out/mla_tmp 2>&1 | cut -c1-170; done
timeout 300 python tools/mla_debug.py --timeline 128 128 8192 > gpurun_out/r2_mla_timeline_pair.log 2>&1; python - <<'PY'
import re
rows=[l for l in open('gpurun_out/r2_mla_timeline_pair.log') if re.match(r'\s+\d+ \|',l)]
base=None
for l in rows[4:12]:
    parts=[x.split() for x in l.split('|')]
    v=[int(x) for x in parts[1]+parts[2]]
    if base is None: base=v[0]
    print(parts[0], [x-base for x in v])
PY
