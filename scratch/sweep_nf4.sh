#!/bin/bash
for shape in "15360 5120" "5120 5120" "13824 5120" "5120 13824" "12288 4096" "11008 4096" "4096 11008"; do
  for qt in nf4 sym_int4; do python scratch/prof_nf4.py $qt fp16 1 $shape 2>&1 | tail -1; done
  python scratch/prof_nf4.py nf4 bf16 1 $shape 2>&1 | tail -1
done
