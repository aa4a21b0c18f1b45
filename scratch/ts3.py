import sys, ctypes; sys.path.insert(0,".")
import torch, numpy as np
from ipex_llm_b200.transformers import low_bit_linear as lbl
import ipex_llm_b200.ggml.model.llama.llama_cpp as ggml
dev=torch.device("cuda",0)
N,K=int(sys.argv[1]),int(sys.argv[2]); pdl=int(sys.argv[3]); HINT=int(sys.argv[4])
ggml.lowbit_set_programmatic_dependent_launch(pdl)
nbytes=N*(K//64)*34
NK=8
ws=[torch.randint(0,255,(nbytes,),dtype=torch.uint8,device=dev) for _ in range(NK)]
x=torch.randn(1,K,device=dev).half(); out=torch.empty(1,N,device=dev,dtype=torch.float16)
ts=torch.zeros(NK,4096*8,dtype=torch.int64,device=dev)
s=torch.cuda.Stream()
with torch.cuda.stream(s):
    for i in range(3): lbl.lowbit_matmul(x,ws[i],(N,K),2,out=out,next_weight=(ws[(i+1)%NK] if HINT else None))
    torch.cuda.synchronize()
    g=torch.cuda.CUDAGraph()
    with torch.cuda.graph(g,stream=s):
        for i in range(NK):
            ggml._lib.lowbit_debug_set_timestamps(ctypes.c_void_p(ts[i].data_ptr()))
            lbl.lowbit_matmul(x,ws[i],(N,K),2,out=out,next_weight=(ws[(i+1)%NK] if HINT else None))
    ggml._lib.lowbit_debug_set_timestamps(None)
torch.cuda.synchronize()
for _ in range(3): g.replay()
torch.cuda.synchronize()
T=ts.cpu().numpy().reshape(NK,4096,8)
base=None
names=["start","issued","afterwait","firstdata","loopdone","flushed"]
for i in range(NK):
    t=T[i]; t=t[t[:,0]>0]
    if base is None: base=t[:,0].min()
    line=f"k{i}: " + "  ".join(f"{n}[{(t[:,j].min()-base)/1e3:6.2f},{(np.median(t[:,j])-base)/1e3:6.2f},{(t[:,j].max()-base)/1e3:6.2f}]" for j,n in enumerate(names))
    print(line)
