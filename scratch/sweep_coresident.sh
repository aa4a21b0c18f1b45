#!/bin/bash
# development: does leaving room for the NEXT launch's CTAs (PDL) hide the per-launch fixed cost?
export IPEX_LLM_B200_DECODE_FUSE_NORM=0
run() { echo "== $*"; env "$@" python scratch/prof_decode.py --only-full 2>&1 | tail -1; }
run A=base
run LOWBIT_GEMV_PAD_L=6 LOWBIT_GEMV_OCC=0.5
run LOWBIT_GEMV_PAD_L=6 LOWBIT_GEMV_OCC=0.67
run LOWBIT_GEMV_PAD_L=6 LOWBIT_GEMV_OCC=0.34
run LOWBIT_GEMV_PAD_L=4 LOWBIT_GEMV_OCC=0.5
run LOWBIT_GEMV_PAD_L=5 LOWBIT_GEMV_OCC=0.6
run LOWBIT_GEMV_PAD_L=8 LOWBIT_GEMV_OCC=0.5
run LOWBIT_GEMV_PAD_L=6 LOWBIT_GEMV_OCC=0.5 LOWBIT_GEMV_CLUSTER_POLICY=1
run LOWBIT_GEMV_PAD_L=6 LOWBIT_GEMV_OCC=0.5 LOWBIT_GEMV_CLUSTER_POLICY=0
