import sys, math; sys.path.insert(0,".")
import torch
from ipex_llm_b200.transformers import kv
dev=torch.device("cuda",0)
def bench(B,Hq,Hkv,L,D=128,S=1,dt=torch.float16):
    q=torch.randn(B,Hq,S,D,device=dev).to(dt)
    caches=[]
    for i in range(8):
        c=kv.DynamicFp8Cache(); k,v=c.update(torch.randn(B,Hkv,L,D,device=dev).to(dt),torch.randn(B,Hkv,L,D,device=dev).to(dt),0); caches.append((k,v))
    st=torch.cuda.Stream()
    with torch.cuda.stream(st):
        for k,v in caches: kv.sdp_fp8(q,k,v,None,1/math.sqrt(D))
        torch.cuda.synchronize()
        g=torch.cuda.CUDAGraph()
        with torch.cuda.graph(g,stream=st):
            for _ in range(4):
                for k,v in caches: kv.sdp_fp8(q,k,v,None,1/math.sqrt(D))
        torch.cuda.synchronize()
        e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True)
        g.replay(); e0.record()
        for _ in range(5): g.replay()
        e1.record(); torch.cuda.synchronize()
    us=e0.elapsed_time(e1)/5/32*1e3
    byts=2*B*Hkv*L*D
    # fp16 torch sdpa on restored cache for comparison
    k16,v16=kv.restore_fp8_kv_cache(caches[0][0],caches[0][1],dt)
    if Hq!=Hkv: k16=k16.repeat_interleave(Hq//Hkv,1); v16=v16.repeat_interleave(Hq//Hkv,1)
    for _ in range(3): torch.nn.functional.scaled_dot_product_attention(q,k16,v16)
    torch.cuda.synchronize(); e0.record()
    for _ in range(20): torch.nn.functional.scaled_dot_product_attention(q,k16,v16)
    e1.record(); torch.cuda.synchronize()
    print(f"B={B} Hq={Hq} Hkv={Hkv} L={L} S={S}: {us:7.2f} us  {byts/us/1e6:5.2f} TB/s of fp8 KV | torch sdpa fp16 (dense KV, eager) {e0.elapsed_time(e1)/20*1e3:7.1f} us",flush=True)
import os
if os.environ.get("MB_SHORT"):
    bench(1,32,32,2048); bench(1,32,32,32768); bench(1,32,8,8192); bench(16,32,8,4096)
else:
    for L in (512,2048,8192,32768): bench(1,32,32,L)
    for L in (2048,8192,32768): bench(1,32,8,L)
    bench(16,32,8,4096); bench(4,32,32,4096); bench(1,32,8,8192,S=4); bench(8,8,8,4096,D=64)
