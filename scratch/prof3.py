import sys; sys.path.insert(0,".")
import torch
from ipex_llm_b200.transformers import low_bit_linear as lbl
import ipex_llm_b200.ggml.model.llama.llama_cpp as ggml
dev=torch.device("cuda",0)
qt=2; H,I=4096,11008
w=torch.randint(0,255,(I*(H//64)*34,),dtype=torch.uint8,device=dev)
xh=torch.randn(1,H,device=dev).half()*0.01
oi=torch.empty(1,I,device=dev,dtype=torch.float16)
ggml.lowbit_set_programmatic_dependent_launch(0)
for _ in range(12): lbl.lowbit_matmul(xh,w,(I,H),qt,out=oi)
torch.cuda.synchronize()
e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20): lbl.lowbit_matmul(xh,w,(I,H),qt,out=oi)
e1.record(); torch.cuda.synchronize(); print("L2-hot gate no-pdl us/iter", e0.elapsed_time(e1)/20*1e3)
