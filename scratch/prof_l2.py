"""development: the M = 1 GEMV with L2-resident weights for ncu (--cache-control none)"""
import sys; sys.path.insert(0, ".")
import torch
from ipex_llm_b200.transformers import low_bit_linear as lbl
dev = torch.device("cuda", 0)
q = lbl.ggml_tensor_qtype["sym_int4"]
N, K = 12288, 4096
w = torch.randint(0, 255, (N * (K // 64) * 34,), dtype=torch.uint8, device=dev)
x = torch.randn(1, K, device=dev).half()
out = torch.empty(1, N, device=dev, dtype=torch.float16)
for _ in range(40):
    lbl.lowbit_matmul(x, w, (N, K), q, out=out)
torch.cuda.synchronize()
