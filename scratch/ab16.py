import sys, os
sys.path.insert(0, os.getcwd())
import torch, bench
dev = torch.device("cuda", 0)
model = bench.build_llama(dev, 32)
ids = torch.randint(0, 32000, (16, 512), device=dev)
with torch.no_grad():
    for i in range(5):
        model.generate(ids, max_new_tokens=33, min_new_tokens=33, do_sample=False, pad_token_id=0)
        print(f"  batch 16 run {i}: {model.rest_cost_mean_device * 1e3:.4f} ms/step", flush=True)
