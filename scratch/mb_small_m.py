"""development: decode GEMV / fused MLP at M = 1..8 (PDL chain of 12 matrices), us per launch"""
import sys; sys.path.insert(0, ".")
import torch
from ipex_llm_b200.transformers import low_bit_linear as lbl
import ipex_llm_b200.ggml.model.llama.llama_cpp as ggml
dev = torch.device("cuda", 0)
ggml.lowbit_set_programmatic_dependent_launch(1)


def bench(N, K, M, qt=2, mlp=False):
    nbytes = N * (K // 64) * ggml.ggml_type_size(qt)
    ws = [torch.randint(0, 255, (nbytes,), dtype=torch.uint8, device=dev) for _ in range(12)]
    x = torch.randn(M, K, device=dev).half(); out = torch.empty(M, N, device=dev, dtype=torch.float16)
    def run():
        for i, w in enumerate(ws):
            if mlp: lbl.mlp_forward(x, w, ws[(i + 1) % 12], M, K, N, lbl.SILU, qt, out=out)
            else: lbl.lowbit_matmul(x, w, (N, K), qt, out=out)
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        run(); torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=st):
            run()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        g.replay(); e0.record()
        for _ in range(5): g.replay()
        e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / 5 / len(ws) * 1e3
for (N, K) in ((12288, 4096), (4096, 4096), (4096, 11008)):
    print(f"{N}x{K}", " ".join(f"M={M}:{bench(N, K, M):5.1f}" for M in (1, 2, 3, 4, 5, 8)), flush=True)
print("mlp 11008x4096", " ".join(f"M={M}:{bench(11008, 4096, M, mlp=True):5.1f}" for M in (1, 2, 4, 8)), flush=True)
