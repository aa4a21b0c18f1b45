"""development: one prefill GEMM (M = 2048) and one skinny GEMM (M = 32) for ncu"""
import sys; sys.path.insert(0, ".")
import torch
from ipex_llm_b200.transformers import low_bit_linear as lbl
dev = torch.device("cuda", 0)
gen = torch.Generator(device=dev).manual_seed(0)
N, K = 11008, 4096
q = lbl.ggml_tensor_qtype["sym_int4"]
w = torch.empty(N, K, device=dev).normal_(0.0, 0.02, generator=gen)
tiled = lbl.ggml_q_format_convet_cpu2xpu(lbl.ggml_convert_qtype(w, q, device=dev), N * K, q, shape=(N, K))
for M in (2048, 32):
    x = torch.randn(M, K, device=dev, generator=gen).half()
    out = torch.empty(M, N, device=dev, dtype=torch.float16)
    for _ in range(3):
        lbl.lowbit_matmul(x, tiled, (N, K), q, out=out)
torch.cuda.synchronize()
