import sys; sys.path.insert(0,".")
import torch
from ipex_llm_b200.transformers import low_bit_linear as lbl
import ipex_llm_b200.ggml.model.llama.llama_cpp as ggml
dev=torch.device("cuda",0)
N,K,qt=int(sys.argv[1]),int(sys.argv[2]),int(sys.argv[3]) if len(sys.argv)>3 else 2
nbytes=N*(K//64)*ggml.ggml_type_size(qt)
ws=[torch.randint(0,255,(nbytes,),dtype=torch.uint8,device=dev) for _ in range(6)]
x=torch.randn(1,K,device=dev).half(); out=torch.empty(1,N,device=dev,dtype=torch.float16)
for i in range(6): lbl.lowbit_matmul(x,ws[i],(N,K),qt,out=out)
torch.cuda.synchronize()
# tiny fixed-overhead probe
e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True)
for (n,k) in [(16,4096),(256,4096),(2368*16//64,4096)]:
    w=torch.randint(0,255,(n*(k//64)*34,),dtype=torch.uint8,device=dev); xx=torch.randn(1,k,device=dev).half(); o=torch.empty(1,n,device=dev,dtype=torch.float16)
    for i in range(3): lbl.lowbit_matmul(xx,w,(n,k),2,out=o)
    torch.cuda.synchronize(); e0.record()
    for i in range(20): lbl.lowbit_matmul(xx,w,(n,k),2,out=o)
    e1.record(); torch.cuda.synchronize(); print("fixed probe",n,k,e0.elapsed_time(e1)/20*1e3,"us")
