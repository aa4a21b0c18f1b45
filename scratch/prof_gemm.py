import sys; sys.path.insert(0,".")
import torch
from ipex_llm_b200.transformers import low_bit_linear as lbl
import ipex_llm_b200.ggml.model.llama.llama_cpp as ggml
dev=torch.device("cuda",0)
N,K,M=11008,4096,2048
w=torch.randint(0,255,(N*(K//64)*34,),dtype=torch.uint8,device=dev)
x=torch.randn(M,K,device=dev).half(); out=torch.empty(M,N,device=dev,dtype=torch.float16)
for i in range(4): lbl.lowbit_matmul(x,w,(N,K),2,out=out)
torch.cuda.synchronize()
