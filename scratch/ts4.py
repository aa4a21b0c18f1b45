import sys, ctypes; sys.path.insert(0,".")
import torch, numpy as np
from ipex_llm_b200.transformers import low_bit_linear as lbl
import ipex_llm_b200.ggml.model.llama.llama_cpp as ggml
dev=torch.device("cuda",0)
N,K=int(sys.argv[1]),int(sys.argv[2])
ggml.lowbit_set_programmatic_dependent_launch(1)
nbytes=N*(K//64)*34
NK=8
ws=[torch.randint(0,255,(nbytes,),dtype=torch.uint8,device=dev) for _ in range(NK)]
x=torch.randn(1,K,device=dev).half(); out=torch.empty(1,N,device=dev,dtype=torch.float16)
ts=torch.zeros(NK,8192*8,dtype=torch.int64,device=dev)
s=torch.cuda.Stream()
with torch.cuda.stream(s):
    for i in range(3): lbl.lowbit_matmul(x,ws[i],(N,K),2,out=out)
    torch.cuda.synchronize()
    g=torch.cuda.CUDAGraph()
    with torch.cuda.graph(g,stream=s):
        for i in range(NK):
            ggml._lib.lowbit_debug_set_timestamps(ctypes.c_void_p(ts[i].data_ptr()))
            lbl.lowbit_matmul(x,ws[i],(N,K),2,out=out)
    ggml._lib.lowbit_debug_set_timestamps(None)
torch.cuda.synchronize()
for _ in range(3): g.replay()
torch.cuda.synchronize()
T=ts.cpu().numpy().reshape(NK,8192,8)
i=5
t=T[i]; t=t[t[:,0]>0].astype(np.float64)
base=t[:,2].min()   # release
start=(t[:,0]-base)/1e3; wait=(t[:,2]-base)/1e3; first=(t[:,3]-base)/1e3; loop=(t[:,4]-base)/1e3; fl=(t[:,5]-base)/1e3
print("warps",len(t))
print("corr(start, loopdone)",np.corrcoef(start,loop)[0,1])
for lo,hi in [(-99,-4),(-4,-2),(-2,-1),(-1,-0.5),(-0.5,0),(0,9)]:
    m=(start>=lo)&(start<hi)
    if m.sum(): print(f"start in [{lo},{hi}): n={m.sum():5d} firstdata med {np.median(first[m]):5.2f} max {first[m].max():5.2f}  loopdone med {np.median(loop[m]):5.2f} max {loop[m].max():5.2f}  compute(loop-first) med {np.median((loop-first)[m]):5.2f}")
q=np.percentile(loop,[5,25,50,75,95,99,100]); print("loopdone pct 5/25/50/75/95/99/100",np.round(q,2))
q=np.percentile(loop-first,[5,50,95,100]); print("compute pct",np.round(q,2))
# per CTA (4 consecutive warps) and per cluster
sm=t[:,6].astype(int)-1
cnt=np.bincount(sm[::1],minlength=148)//1
print("warps per SM: min",cnt.min(),"max",cnt.max(),"hist",np.bincount(cnt//4))
# per-SM finish time
fin=np.array([loop[sm==k].max() if (sm==k).any() else 0 for k in range(148)])
for c in sorted(set(cnt)):
    m=cnt==c
    print(f"SMs with {c//4} CTAs: {m.sum():3d}  finish med {np.median(fin[m]):5.2f} max {fin[m].max():5.2f}")
