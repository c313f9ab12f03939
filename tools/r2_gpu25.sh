#!/bin/bash
echo "old kernel:"; B200_PAGED_ATTN_LIB=$PWD/build/old_lib/libb200_paged_attn.so timeout 300 python tools/mla_debug.py --timeline 128 64 8192 2>&1 | sed -n 3,3p | cut -c1-170
echo "new kernel, split mode:"; B200_MLA_FLAT_OFF=1 timeout 300 python tools/mla_debug.py --timeline 128 64 8192 2>&1 | sed -n 3,3p | cut -c1-170
echo "old again:"; B200_PAGED_ATTN_LIB=$PWD/build/old_lib/libb200_paged_attn.so timeout 300 python tools/mla_debug.py --timeline 128 64 8192 2>&1 | sed -n 3,3p | cut -c1-170
