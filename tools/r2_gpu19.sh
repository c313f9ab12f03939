#!/bin/bash
# whole GPU suite, smoke, default bench line (round-2 tree with MLA decode)
timeout 2400 python -m pytest tests -q -m gpu > gpurun_out/r2_t19.log 2>&1
echo "tests rc=$?"; tail -6 gpurun_out/r2_t19.log | cut -c1-300
timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r2_smoke19.log 2>&1; tail -2 gpurun_out/r2_smoke19.log
timeout 600 python bench.py > gpurun_out/r2_bench19.json 2> gpurun_out/r2_bench19.err
echo "bench rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/r2_bench19.json')); print(d['value'], d['roofline']['frac'], d['e2e']['value'], d['parity_sample']['max_abs']);
print([(c['name'], round(c.get('frac',0),3)) for c in d['extra']['configs']]); print([(c['name'], round(c.get('attention_us',0),1), c.get('parity_ok')) for c in d['extra']['mla']])"
timeout 300 python tools/movers_bench.py > gpurun_out/r2_movers_bench.log 2>&1; echo "movers rc=$?"; cat gpurun_out/r2_movers_bench.log | cut -c1-200 | tail -14
