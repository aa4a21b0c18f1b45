import sys; sys.path.insert(0,".")
import torch
from ipex_llm_b200.transformers import low_bit_linear as lbl
import ipex_llm_b200.ggml.model.llama.llama_cpp as ggml
dev=torch.device("cuda",0)
def bench(N,K,M,qt=2,xdt=torch.float16):
    nbytes=N*(K//64)*ggml.ggml_type_size(qt)
    ws=[torch.randint(0,255,(nbytes,),dtype=torch.uint8,device=dev) for _ in range(12)]
    x=torch.randn(M,K,device=dev).to(xdt); out=torch.empty(M,N,device=dev,dtype=xdt)
    st=torch.cuda.Stream()
    with torch.cuda.stream(st):
        for w in ws: lbl.lowbit_matmul(x,w,(N,K),qt,out=out)
        torch.cuda.synchronize()
        g=torch.cuda.CUDAGraph()
        with torch.cuda.graph(g,stream=st):
            for w in ws: lbl.lowbit_matmul(x,w,(N,K),qt,out=out)
        e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True)
        g.replay(); e0.record()
        for _ in range(5): g.replay()
        e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1)/5/len(ws)*1e3
for (N,K) in ((11008,4096),(4096,11008),(12288,4096),(4096,4096),(14336,4096)):
    print(N,K," ".join(f"M={M}:{bench(N,K,M):6.1f}" for M in (9,16,17,24,32)),flush=True)
print("fp8", " ".join(f"M={M}:{bench(14336,4096,M,qt=15):6.1f}" for M in (16,32)),flush=True)
print("bf16", " ".join(f"M={M}:{bench(11008,4096,M,xdt=torch.bfloat16):6.1f}" for M in (16,32)),flush=True)
