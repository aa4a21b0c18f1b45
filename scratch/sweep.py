"""GEMV sweep: per (qtype, shape, M) the period of PDL-chained launches in a CUDA graph (distinct weights)."""
import sys; sys.path.insert(0,".")
import torch
from ipex_llm_b200.transformers import low_bit_linear as lbl
import ipex_llm_b200.ggml.model.llama.llama_cpp as ggml
from ipex_llm_b200.ggml.quantize import ggml_tensor_qtype as QT
dev=torch.device("cuda",0)
ggml.lowbit_set_programmatic_dependent_launch(1)
qtypes=sys.argv[1].split(",") if len(sys.argv)>1 else ["sym_int4"]
Ms=[int(m) for m in sys.argv[2].split(",")] if len(sys.argv)>2 else [1]
shapes=[(11008,4096),(4096,11008),(12288,4096),(4096,4096),(32000,4096),(14336,4096),(4096,14336),(6144,4096),(28672,8192),(8192,28672)]
if len(sys.argv)>3: shapes=[tuple(int(v) for v in s.split("x")) for s in sys.argv[3].split(",")]
dt=torch.float16 if len(sys.argv)<=4 else {"f16":torch.float16,"bf16":torch.bfloat16,"f32":torch.float32}[sys.argv[4]]
st=torch.cuda.Stream()
for q in qtypes:
    qt=QT[q]; bs=ggml.ggml_type_size(qt)
    for (N,K) in shapes:
        nbytes=N*(K//64)*bs
        L=max(2,min(8,int(3e8//nbytes)))
        ws=[torch.randint(0,255,(nbytes,),dtype=torch.uint8,device=dev) for _ in range(L)]
        for M in Ms:
            x=(torch.randn(M,K,device=dev)*0.01).to(dt); o=torch.empty(M,N,device=dev,dtype=dt)
            with torch.cuda.stream(st):
                for w in ws: lbl.lowbit_matmul(x,w,(N,K),qt,out=o)
                torch.cuda.synchronize()
                g=torch.cuda.CUDAGraph()
                with torch.cuda.graph(g,stream=st):
                    for _ in range(4):
                        for w in ws: lbl.lowbit_matmul(x,w,(N,K),qt,out=o)
                torch.cuda.synchronize()
                e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True)
                g.replay(); e0.record()
                for _ in range(5): g.replay()
                e1.record(); torch.cuda.synchronize()
            us=e0.elapsed_time(e1)/5/(4*L)*1e3
            print(f"{q:10s} {N:6d}x{K:<6d} M={M:<3d} {us:8.2f} us  {nbytes/us/1e6:6.2f} TB/s", flush=True)
        del ws
