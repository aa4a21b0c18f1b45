import sys, os; sys.path.insert(0,".")
import torch
from ipex_llm_b200.transformers import low_bit_linear as lbl
import ipex_llm_b200.ggml.model.llama.llama_cpp as ggml
dev=torch.device("cuda",0)
def bench(N,K,M=1,qt=2,xdt=torch.float16,nw=24,tag=""):
    bs=ggml.ggml_type_size(qt); nbytes=N*(K//64)*bs
    nw=max(2,min(nw,int(600e6//nbytes)+1))
    ws=[torch.randint(0,255,(nbytes,),dtype=torch.uint8,device=dev) for _ in range(nw)]
    x=torch.randn(M,K,device=dev).to(xdt); out=torch.empty(M,N,device=dev,dtype=xdt)
    g=torch.cuda.CUDAGraph(); s=torch.cuda.Stream()
    with torch.cuda.stream(s):
        for i in range(3): lbl.lowbit_matmul(x,ws[i%nw],(N,K),qt,out=out)
        torch.cuda.synchronize()
        with torch.cuda.graph(g,stream=s):
            for i in range(nw): lbl.lowbit_matmul(x,ws[i%nw],(N,K),qt,out=out)
    torch.cuda.synchronize(); g.replay(); torch.cuda.synchronize()
    e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): g.replay()
    e1.record(); torch.cuda.synchronize()
    msg=e0.elapsed_time(e1)/(5*nw)
    print(f"{tag:12s} N={N:6d} K={K:6d} M={M:2d} qt={qt:2d} bytes={nbytes/1e6:7.2f}MB graph {msg*1e3:7.1f}us {nbytes/msg/1e6:7.0f}GB/s",flush=True)
for dbg in (0,):
    os.environ["LOWBIT_GEMV_DBG"]=str(dbg)
    bench(11008,4096,tag=f"dbg={dbg}")
os.environ["LOWBIT_GEMV_DBG"]="0"
for N in (16,64,592,2368,4736,56832):
    bench(N,4096,tag="sizes")
ggml.lowbit_set_programmatic_dependent_launch(1)
for N in (16,4096,11008,32000):
    bench(N,4096,tag="pdl")
bench(4096,11008,tag="pdl")
