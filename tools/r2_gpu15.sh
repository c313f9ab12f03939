#!/bin/bash
# MLA witness from the stock kernel, witness tests, sanitizer passes over the reduced grid, default bench line
timeout 300 python tools/mla_bench.py --impl upstream --witness gpurun_out/mla_upstream_bf16_h128.pt 2>&1 | tail -2
cp gpurun_out/mla_upstream_bf16_h128.pt tests/golden/ 2>/dev/null
timeout 300 python -m pytest tests/test_upstream_witness.py -q -k mla 2>&1 | tail -3
timeout 200 python tools/sanitizer_cases.py > gpurun_out/r2_sanitizer_plain.log 2>&1; echo "plain rc=$?"; tail -3 gpurun_out/r2_sanitizer_plain.log
for tool in memcheck racecheck; do
  timeout 1500 compute-sanitizer --tool $tool --print-limit 20 python tools/sanitizer_cases.py > gpurun_out/r2_sanitizer_$tool.log 2>&1
  echo "$tool rc=$?"; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|hazard|Invalid|sanitizer cases done" gpurun_out/r2_sanitizer_$tool.log | sort | uniq -c | head -12
done
timeout 600 python bench.py > gpurun_out/r2_bench15.json 2> gpurun_out/r2_bench15.err; echo "bench rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/r2_bench15.json')); print(d['value'], d['roofline']['frac'], d['e2e']['value']); print(json.dumps(d['extra'].get('mla'), indent=1)[:1500])"
