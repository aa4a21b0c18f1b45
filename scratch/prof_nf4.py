"""development: nf4 decode GEMV under ncu / timing.  python scratch/prof_nf4.py [qtype]"""
import sys; sys.path.insert(0, ".")
import torch
from ipex_llm_b200.transformers import low_bit_linear as lbl
import ipex_llm_b200.ggml.model.llama.llama_cpp as ggml
qt = sys.argv[1] if len(sys.argv) > 1 else "nf4"
dt = {"fp16": torch.float16, "bf16": torch.bfloat16}[sys.argv[2] if len(sys.argv) > 2 else "fp16"]
M = int(sys.argv[3]) if len(sys.argv) > 3 else 1
dev = torch.device("cuda", 0)
ggml.lowbit_set_programmatic_dependent_launch(1)
N = int(sys.argv[4]) if len(sys.argv) > 4 else 14336
K = int(sys.argv[5]) if len(sys.argv) > 5 else 4096
q = lbl.ggml_tensor_qtype[qt]
gen = torch.Generator(device=dev).manual_seed(0)
ws = [lbl.FP4Params(torch.empty(N, K, device=dev).normal_(0.0, 0.02, generator=gen), qtype=q).to(dev).data for _ in range(10)]
x = torch.randn(M, K, device=dev, generator=gen).to(dt)
out = torch.empty(M, N, device=dev, dtype=dt)
st = torch.cuda.Stream()
with torch.cuda.stream(st):
    for w in ws: lbl.lowbit_matmul(x, w, (N, K), q, out=out)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=st):
        for w in ws: lbl.lowbit_matmul(x, w, (N, K), q, out=out)
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    g.replay(); e0.record()
    for _ in range(5): g.replay()
    e1.record(); torch.cuda.synchronize()
us = e0.elapsed_time(e1) / 5 / len(ws) * 1e3
nbytes = N * (K // 64) * ggml.ggml_type_size(q)
print(f"{qt} {dt} {N}x{K} M={M}: {us:.2f} us per launch in a PDL chain, {nbytes / us / 1e6:.2f} TB/s")
