#!/bin/bash
# developer sweep: tcgen05 fp8 kernel vs the mma.sync kernel on the fp8 configs
run() { echo "== UMMA_OFF=${B200_DISABLE_UMMA:-0} $*"; timeout 120 python tools/quick_bench.py "$@" 2>&1 | tail -1; }
for off in "" 1; do
  if [ -n "$off" ]; then export B200_DISABLE_UMMA=1; else unset B200_DISABLE_UMMA; fi
  run --fp8
  run --fp8 --v2
  run --S 128 --H 8 --Hkv 1 --L 8192 --fp8
  run --S 128 --H 16 --Hkv 2 --L 8192 --fp8
  run --S 8 --L 131072 --v2 --fp8
  run --S 8 --L 131072 --fp8
  run --S 8 --H 32 --Hkv 32 --L 1024 --fp8
  run --S 256 --L 1024 --fp8
done
