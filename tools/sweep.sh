#!/bin/bash
# run a list of quick_bench configs, each in its own process with a timeout
export CUDA_LAUNCH_BLOCKING=1
while read -r args; do
  [ -z "$args" ] && continue
  echo "== $args"
  timeout 60 python tools/quick_bench.py $args --iters 3 2>&1 | tail -2
  echo "rc=$?"
done
