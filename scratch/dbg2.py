import sys; sys.path.insert(0,".")
import torch
from ipex_llm_b200.transformers import low_bit_linear as lbl
dev=torch.device("cuda",0)
for (N,K) in [(11008,4096),(12288,4096),(4096,11008),(4096,4096),(32000,4096)]:
    w=torch.randint(0,255,(N*(K//64)*34,),dtype=torch.uint8,device=dev)
    x=torch.randn(1,K,device=dev).half(); o=torch.empty(1,N,device=dev,dtype=torch.float16)
    lbl.lowbit_matmul(x,w,(N,K),2,out=o)
w=torch.randint(0,255,(11008*64*34,),dtype=torch.uint8,device=dev)
x=torch.randn(1,4096,device=dev).half()
lbl.mlp_forward(x,w,w,1,4096,11008,0,2)
torch.cuda.synchronize()
