import sys, os
sys.path.insert(0, os.getcwd())
import torch, bench
dev = torch.device("cuda", 0)
model = bench.build_llama(dev, 32)
ids = torch.randint(0, 32000, (16, 512), device=dev)
with torch.no_grad():
    model.generate(ids, max_new_tokens=4, min_new_tokens=4, do_sample=False, pad_token_id=0)
    torch.cuda.synchronize()
    torch.cuda.cudart().cudaProfilerStart()
    model.generate(ids, max_new_tokens=3, min_new_tokens=3, do_sample=False, pad_token_id=0)
    torch.cuda.synchronize()
    torch.cuda.cudart().cudaProfilerStop()
