import sys; sys.path.insert(0,".")
import torch
from ipex_llm_b200.transformers import low_bit_linear as lbl
import ipex_llm_b200.ggml.model.llama.llama_cpp as ggml
dev=torch.device("cuda",0)
qt=int(sys.argv[1]); M=int(sys.argv[2]); N=int(sys.argv[3]); K=int(sys.argv[4])
nbytes=N*(K//64)*ggml.ggml_type_size(qt)
w=torch.randint(0,255,(nbytes,),dtype=torch.uint8,device=dev)
x=torch.randn(M,K,device=dev).half(); out=torch.empty(M,N,device=dev,dtype=torch.float16)
for i in range(3): lbl.lowbit_matmul(x,w,(N,K),qt,out=out)
torch.cuda.synchronize()
print("ok", qt, M, N, K, float(out.float().abs().mean()))
