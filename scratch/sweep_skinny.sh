#!/bin/bash
timeout 300 python scratch/test_skinny.py check 2>&1 | tail -2
timeout 200 python scratch/test_skinny.py time 2>&1 | grep -E "11008x4096|4096x4096 M=(16|32)"
LOWBIT_SK_DBG=8 timeout 200 python scratch/test_skinny.py time 2>&1 | grep -E "11008x4096 M=(16|32|128)"
LLAMA_CPP_LIB=$PWD/scratch/lib_ts.so timeout 120 python scratch/ts_skinny.py | sed -n 8,14p
timeout 300 python scratch/mb_gemm.py 2>&1 | tail -15
timeout 600 python -m pytest tests/test_gpu_parity.py -q -x -k "gemm or prefill or M512 or tc" 2>&1 | tail -3
