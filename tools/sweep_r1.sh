#!/bin/bash
# developer sweep: named configs + v1 chunk sizes
run() { echo "== $*"; timeout 120 python tools/quick_bench.py "$@" 2>&1 | tail -1; }
run                                   # C2 v1
run --v2                              # C2 v2
run --fp8                             # C3 v1
run --fp8 --v2
run --S 8 --H 32 --Hkv 32 --L 1024    # C1
run --S 128 --H 8 --Hkv 1 --L 8192    # C4 TP8 shard
run --S 128 --H 16 --Hkv 2 --L 8192   # C4 TP4 shard
run --S 128 --H 8 --Hkv 1 --L 8192 --fp8
run --S 8 --L 131072 --v2             # C5
run --S 8 --L 131072 --v2 --fp8
run --S 8 --L 131072                  # C5 shape through v1
for c in 512 1024 2048 4096; do
  echo "-- B200_V1_CHUNK=$c"; B200_V1_CHUNK=$c run
  B200_V1_CHUNK=$c run --fp8
done
for n in 8 16; do echo "-- NSTAGES=$n"; B200_STREAM_NSTAGES=$n run --fp8; done
