import sys; sys.path.insert(0,".")
import torch
from ipex_llm_b200.transformers import low_bit_linear as lbl
import ipex_llm_b200.ggml.model.llama.llama_cpp as ggml
dev=torch.device("cuda",0)
def bench(N,K,M,qt=2,xdt=torch.float16,iters=20):
    nbytes=N*(K//64)*ggml.ggml_type_size(qt)
    w=torch.randint(0,255,(nbytes,),dtype=torch.uint8,device=dev)
    x=torch.randn(M,K,device=dev).to(xdt); out=torch.empty(M,N,device=dev,dtype=xdt)
    for i in range(3): lbl.lowbit_matmul(x,w,(N,K),qt,out=out)
    torch.cuda.synchronize()
    e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters): lbl.lowbit_matmul(x,w,(N,K),qt,out=out)
    e1.record(); torch.cuda.synchronize()
    ms=e0.elapsed_time(e1)/iters
    xf=x.half(); wd=torch.randn(N,K,device=dev).half()
    for i in range(3): torch.matmul(xf,wd.t())
    torch.cuda.synchronize(); e0.record()
    for i in range(iters): torch.matmul(xf,wd.t())
    e1.record(); torch.cuda.synchronize()
    msd=e0.elapsed_time(e1)/iters
    print(f"N={N:6d} K={K:6d} M={M:5d} qt={qt:2d} {str(xdt)[6:]:8s} {ms*1e3:8.1f}us {2*M*N*K/ms/1e9:8.1f} TFLOP/s | cuBLAS fp16 dense {msd*1e3:8.1f}us {2*M*N*K/msd/1e9:8.1f} TFLOP/s",flush=True)
for M in (64,128,512,2048):
    bench(4096,4096,M)
for M in (512,2048):
    bench(11008,4096,M); bench(4096,11008,M); bench(12288,4096,M)
bench(11008,4096,2048,qt=3); bench(11008,4096,2048,qt=10); bench(11008,4096,2048,qt=15); bench(11008,4096,2048,qt=29); bench(11008,4096,2048,xdt=torch.bfloat16)
