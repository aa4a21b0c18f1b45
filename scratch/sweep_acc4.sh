#!/bin/bash
for lib in "" scratch/lib_acc4_5.so scratch/lib_acc4_4.so; do
  echo "== lib ${lib:-default (6 CTAs/SM cap)}"
  if [ -n "$lib" ]; then export LLAMA_CPP_LIB=$PWD/$lib; else unset LLAMA_CPP_LIB; fi
  timeout 300 python scratch/sweep_batch.py sym_int4 2>&1 | grep -E "\| (1|2|4|8) \|"
done
