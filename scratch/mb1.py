import sys; sys.path.insert(0,".")
import torch, time, ctypes
from ipex_llm_b200.transformers import low_bit_linear as lbl
import ipex_llm_b200.ggml.model.llama.llama_cpp as ggml
dev=torch.device("cuda",0)
torch.manual_seed(0)
def bench(N,K,M=1,qt=2,xdt=torch.float16,iters=50,nw=24):
    # nw distinct weight copies cycled so that the working set (>126MB L2) does not stay cached
    bs=ggml.ggml_type_size(qt); nbytes=N*(K//64)*bs
    nw=max(2,min(nw,int(600e6//nbytes)+1))
    ws=[torch.randint(0,255,(nbytes,),dtype=torch.uint8,device=dev) for _ in range(nw)]
    x=torch.randn(M,K,device=dev).to(xdt); out=torch.empty(M,N,device=dev,dtype=xdt)
    for i in range(5): lbl.lowbit_matmul(x,ws[i%nw],(N,K),qt,out=out)
    torch.cuda.synchronize()
    e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters): lbl.lowbit_matmul(x,ws[i%nw],(N,K),qt,out=out)
    e1.record(); torch.cuda.synchronize()
    ms=e0.elapsed_time(e1)/iters
    g=torch.cuda.CUDAGraph()
    s=torch.cuda.Stream()
    with torch.cuda.stream(s):
        for i in range(3): lbl.lowbit_matmul(x,ws[i%nw],(N,K),qt,out=out)
        torch.cuda.synchronize()
        with torch.cuda.graph(g,stream=s):
            for i in range(nw): lbl.lowbit_matmul(x,ws[i%nw],(N,K),qt,out=out)
    torch.cuda.synchronize()
    g.replay(); torch.cuda.synchronize()
    e0.record()
    for _ in range(5): g.replay()
    e1.record(); torch.cuda.synchronize()
    msg=e0.elapsed_time(e1)/(5*nw)
    print(f"N={N:6d} K={K:6d} M={M:2d} qt={qt:2d} {str(xdt)[6:]:9s} bytes={nbytes/1e6:7.2f}MB eager {ms*1e3:7.1f}us {nbytes/ms/1e6:7.0f}GB/s | graph {msg*1e3:7.1f}us {nbytes/msg/1e6:7.0f}GB/s",flush=True)
for (N,K) in [(12288,4096),(4096,4096),(11008,4096),(4096,11008),(32000,4096)]:
    bench(N,K)
bench(32000,4096,M=1,xdt=torch.bfloat16)
bench(32000,4096,M=1,xdt=torch.float32)
bench(11008,4096,M=4); bench(11008,4096,M=8); bench(11008,4096,M=16); bench(11008,4096,M=32)
for qt in (3,15,19,29): bench(11008,4096,qt=qt)
