#!/bin/bash
# development: the decode step of two builds on the same box (batch 1 and batch 16)
R=$PWD
for t in scratch/ab_base .; do
  cd $R/$t
  echo -n "$t: "; python scratch/prof_decode.py --only-full 2>&1 | tail -1
  python - <<'PY'
import sys, os
sys.path.insert(0, os.getcwd())
import torch, bench
dev = torch.device("cuda", 0)
model = bench.build_llama(dev, 32)
for B in (4, 16):
    ids = torch.randint(0, 32000, (B, 512), device=dev)
    with torch.no_grad():
        for _ in range(2):
            model.generate(ids, max_new_tokens=33, min_new_tokens=33, do_sample=False, pad_token_id=0)
    print(f"  batch {B}: {model.rest_cost_mean_device * 1e3:.4f} ms/step", flush=True)
PY
done
