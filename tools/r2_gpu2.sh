#!/bin/bash
# timeline probe of the streaming kernel (developer build), then restore the normal build
touch vllm_metax_b200/csrc/attention_stream.cu
B200_EXTRA_NVCC_FLAGS=-DB200_TIMELINE python -m vllm_metax_b200.build > gpurun_out/r2_tl_build.log 2>&1 || { tail gpurun_out/r2_tl_build.log; exit 1; }
{
for args in "--S 8 --H 32 --Hkv 32 --L 1024" "--S 64 --H 32 --Hkv 8 --L 4096" "--S 64 --H 32 --Hkv 8 --L 4096 --fp8" \
            "--S 128 --H 8 --Hkv 1 --L 8192" "--S 128 --H 64 --Hkv 8 --L 8192" "--S 64 --H 32 --Hkv 8 --L 4096 --fp8 --v2"; do
  echo "== $args"; timeout 200 python tools/timeline_probe.py $args 2>&1 | tail -8
done
} > gpurun_out/r2_timeline.log 2>&1
cat gpurun_out/r2_timeline.log
