import sys, math; sys.path.insert(0,".")
import torch
from ipex_llm_b200.transformers import kv
dev=torch.device("cuda",0)
def run(B,Hq,Hkv,L,D=128,S=1,dt=torch.float16):
    q=torch.randn(B,Hq,S,D,device=dev).to(dt)
    c=kv.DynamicFp8Cache(); k,v=c.update(torch.randn(B,Hkv,L,D,device=dev).to(dt),torch.randn(B,Hkv,L,D,device=dev).to(dt),0)
    for _ in range(2): kv.sdp_fp8(q,k,v,None,1/math.sqrt(D))
    torch.cuda.synchronize()
run(1,32,32,32768); run(16,32,8,4096); run(1,32,8,8192)
