import sys; sys.path.insert(0,".")
import torch
from ipex_llm_b200.transformers import low_bit_linear as lbl
import ipex_llm_b200.ggml.model.llama.llama_cpp as ggml
dev=torch.device("cuda",0)
qt=2
H,I=4096,11008
def rw(n,k): return torch.randint(0,255,(n*(k//64)*34,),dtype=torch.uint8,device=dev)
L=8
gate=[rw(I,H) for _ in range(L)]; up=[rw(I,H) for _ in range(L)]; down=[rw(H,I) for _ in range(L)]
qkv=[rw(3*H,H) for _ in range(L)]; o=[rw(H,H) for _ in range(L)]
xh=torch.randn(1,H,device=dev).half()*0.01; xi=torch.randn(1,I,device=dev).half()*0.01
oh=torch.empty(1,H,device=dev,dtype=torch.float16); oi=torch.empty(1,I,device=dev,dtype=torch.float16); oq=torch.empty(1,3*H,device=dev,dtype=torch.float16)
pdl=int(sys.argv[1]) if len(sys.argv)>1 else 1
ggml.lowbit_set_programmatic_dependent_launch(pdl)
def seq(kind):
    for l in range(L):
        if kind=="mlp": lbl.mlp_forward(xh,gate[l],up[l],1,H,I,0,qt,out=oi)
        elif kind=="gate": lbl.lowbit_matmul(xh,gate[l],(I,H),qt,out=oi)
        elif kind=="down": lbl.lowbit_matmul(xi,down[l],(H,I),qt,out=oh)
        elif kind=="qkv": lbl.lowbit_matmul(xh,qkv[l],(3*H,H),qt,out=oq)
        elif kind=="o": lbl.lowbit_matmul(xh,o[l],(H,H),qt,out=oh)
        elif kind=="layer":
            lbl.lowbit_matmul(xh,qkv[l],(3*H,H),qt,out=oq); lbl.lowbit_matmul(xh,o[l],(H,H),qt,out=oh)
            lbl.mlp_forward(xh,gate[l],up[l],1,H,I,0,qt,out=oi); lbl.lowbit_matmul(xi,down[l],(H,I),qt,out=oh)
bytes_={"mlp":2*I*H//64*34,"gate":I*H//64*34,"down":I*H//64*34,"qkv":3*H*H//64*34,"o":H*H//64*34}
bytes_["layer"]=bytes_["mlp"]+bytes_["down"]+bytes_["qkv"]+bytes_["o"]
st=torch.cuda.Stream()
for kind in ["gate","mlp","down","qkv","o","layer"]:
    with torch.cuda.stream(st):
        seq(kind); torch.cuda.synchronize()
        g=torch.cuda.CUDAGraph()
        with torch.cuda.graph(g,stream=st):
            for _ in range(4): seq(kind)
        torch.cuda.synchronize()
        e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True)
        g.replay(); e0.record()
        for _ in range(5): g.replay()
        e1.record(); torch.cuda.synchronize()
        us=e0.elapsed_time(e1)/5/(4*L)*1e3
        print(f"{kind:6s} pdl={pdl} {us:7.2f} us/iter  {bytes_[kind]/us/1e6:6.2f} TB/s")
