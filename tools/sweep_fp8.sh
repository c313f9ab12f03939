#!/bin/bash
# developer sweep: producer-warp count x ring depth for the streaming kernel
run() { echo "== NPROD=$B200_STREAM_NPROD NST=$B200_STREAM_NSTAGES $*"; timeout 120 python tools/quick_bench.py "$@" 2>&1 | tail -1; }
for np in 4 8; do for nst in 8 16 24; do
  export B200_STREAM_NPROD=$np B200_STREAM_NSTAGES=$nst
  run --fp8
  run --fp8 --B 32
  run --S 128 --H 8 --Hkv 1 --L 8192 --fp8
done; done
for np in 4 8; do
  export B200_STREAM_NPROD=$np; unset B200_STREAM_NSTAGES
  run
  run --B 32
  run --fp8 --v2
  run --S 8 --L 131072 --v2 --fp8
done
