import sys, ctypes; sys.path.insert(0,".")
import torch, numpy as np
from ipex_llm_b200.transformers import low_bit_linear as lbl
import ipex_llm_b200.ggml.model.llama.llama_cpp as ggml
dev=torch.device("cuda",0)
N,K=int(sys.argv[1]),int(sys.argv[2]); pdl=int(sys.argv[3]) if len(sys.argv)>3 else 0
ggml.lowbit_set_programmatic_dependent_launch(pdl)
nbytes=N*(K//64)*34
ws=[torch.randint(0,255,(nbytes,),dtype=torch.uint8,device=dev) for _ in range(8)]
x=torch.randn(1,K,device=dev).half(); out=torch.empty(1,N,device=dev,dtype=torch.float16)
ts=torch.zeros(8,4096*8,dtype=torch.int64,device=dev)
for i in range(4): lbl.lowbit_matmul(x,ws[i],(N,K),2,out=out)
torch.cuda.synchronize()
for i in range(8):
    ggml._lib.lowbit_debug_set_timestamps(ctypes.c_void_p(ts[i].data_ptr()))
    lbl.lowbit_matmul(x,ws[i],(N,K),2,out=out)
ggml._lib.lowbit_debug_set_timestamps(None)
torch.cuda.synchronize()
T=ts.cpu().numpy().reshape(8,4096,8)
base=None
for i in range(2,7):
    t=T[i]; valid=t[:,0]>0; t=t[valid]
    t0=t[:,0].min()
    if base is None: base=t0
    names=["start","issued","afterwait","firstdata","loopdone","flushed"]
    print(f"kernel {i}: warps={valid.sum()} kernel_start@{(t0-base)/1e3:.2f}us  end@{(t[:,5].max()-base)/1e3:.2f}us")
    for j,n in enumerate(names):
        d=(t[:,j]-t0)/1e3
        print(f"   {n:10s} min {d.min():6.2f} med {np.median(d):6.2f} p90 {np.percentile(d,90):6.2f} max {d.max():6.2f} us")
