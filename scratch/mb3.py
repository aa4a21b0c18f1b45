import sys, os; sys.path.insert(0,".")
import torch
dev=torch.device("cuda",0)
def timeit(fn, n=24, tag=""):
    g=torch.cuda.CUDAGraph(); s=torch.cuda.Stream()
    with torch.cuda.stream(s):
        for i in range(3): fn(i)
        torch.cuda.synchronize()
        with torch.cuda.graph(g,stream=s):
            for i in range(n): fn(i)
    torch.cuda.synchronize(); g.replay(); torch.cuda.synchronize()
    e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): g.replay()
    e1.record(); torch.cuda.synchronize()
    print(f"{tag:30s} {e0.elapsed_time(e1)/(5*n)*1e3:8.2f} us/kernel",flush=True)
a=torch.zeros(1024,device=dev); b=torch.zeros(1024,device=dev)
timeit(lambda i: torch.add(a,1,out=b), tag="tiny add")
big=[torch.empty(24_000_000,dtype=torch.uint8,device=dev) for _ in range(24)]
dst=torch.empty(24_000_000,dtype=torch.uint8,device=dev)
timeit(lambda i: dst.copy_(big[i]), tag="24MB copy (r+w)")
acc=torch.zeros(1,device=dev,dtype=torch.int64)
bigi=[b_.view(torch.int64) for b_ in big]
timeit(lambda i: torch.sum(bigi[i],dim=0,keepdim=True,out=acc), tag="24MB sum int64 (read)")
