import sys; sys.path.insert(0,".")
import torch
from ipex_llm_b200.transformers import low_bit_linear as lbl
import ipex_llm_b200.ggml.model.llama.llama_cpp as ggml
dev=torch.device("cuda",0)
qt=2
N,K=int(sys.argv[1]),int(sys.argv[2])
L=8
ws=[torch.randint(0,255,(N*(K//64)*34,),dtype=torch.uint8,device=dev) for _ in range(L)]
x=torch.randn(1,K,device=dev).half()*0.01; o=torch.empty(1,N,device=dev,dtype=torch.float16)
ggml.lowbit_set_programmatic_dependent_launch(1)
st=torch.cuda.Stream()
with torch.cuda.stream(st):
    for w in ws: lbl.lowbit_matmul(x,w,(N,K),qt,out=o)
    torch.cuda.synchronize()
    g=torch.cuda.CUDAGraph()
    with torch.cuda.graph(g,stream=st):
        for _ in range(4):
            for w in ws: lbl.lowbit_matmul(x,w,(N,K),qt,out=o)
    torch.cuda.synchronize()
    e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True)
    g.replay(); e0.record()
    for _ in range(5): g.replay()
    e1.record(); torch.cuda.synchronize()
us=e0.elapsed_time(e1)/5/(4*L)*1e3
print(f"{N}x{K} {us:7.2f} us  {N*(K//64)*34/us/1e6:6.2f} TB/s", sys.argv[3:] )
