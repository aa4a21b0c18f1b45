#!/bin/bash
# development: decode step time against the split count of the one-launch attention
for n in 1 2 3 4 6 8; do
  echo "== splits $n"; LOWBIT_DECODE_ATTN_SPLITS=$n timeout 200 python scratch/prof_decode.py --only-full 2>&1 | grep "full step" | head -1
done
