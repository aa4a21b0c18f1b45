#!/bin/bash
run() { echo -n "$* : "; env "$@" python scratch/prof_decode.py --only-full 2>&1 | tail -1; }
run IPEX_LLM_B200_DECODE_PREFETCH=0
run LOWBIT_GEMV_PREFETCH_MB=16
run LOWBIT_GEMV_PREFETCH_MB=8
run LOWBIT_GEMV_PREFETCH_MB=32
run LOWBIT_GEMV_PREFETCH_MB=64
run IPEX_LLM_B200_DECODE_PREFETCH=0
