#!/bin/bash
for tag in base cur; do
  if [ $tag = base ]; then export LLAMA_CPP_LIB=$PWD/scratch/ab_base/ipex_llm_b200/libs/libb200lowbit.so; else unset LLAMA_CPP_LIB; fi
  timeout 500 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/b16_$tag.csv python scratch/prof_b16.py > gpurun_out/b16_$tag.log 2>&1
  tail -1 gpurun_out/b16_$tag.log
done
