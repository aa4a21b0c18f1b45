#!/bin/bash
# development: the decode step of two builds on the same box
R=$PWD
for rep in 1 2; do
for t in scratch/ab_base .; do
  cd $R/$t
  echo -n "$t: "; python scratch/prof_decode.py --only-full 2>&1 | tail -1
done
cd $R
done
