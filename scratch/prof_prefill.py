"""development: kernel table of one 512-token prefill through the public forward (torch profiler)"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench
from torch.profiler import profile, ProfilerActivity
dev = torch.device("cuda", 0)
model = bench.build_llama(dev, 32)
ids = torch.randint(0, 32000, (1, int(sys.argv[1]) if len(sys.argv) > 1 else 512), device=dev)
with torch.no_grad():
    for _ in range(3):
        model(ids, use_cache=True)
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record(); model(ids, use_cache=True); e1.record(); torch.cuda.synchronize()
    print("prefill ms", e0.elapsed_time(e1))
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        model(ids, use_cache=True)
        torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=25, max_name_column_width=70))
print(prof.key_averages().table(sort_by="self_cpu_time_total", row_limit=22, max_name_column_width=60))
