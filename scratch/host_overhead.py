"""development: host time per call of the Python wrappers (tiny shapes, the GPU is never the limit)"""
import sys, time; sys.path.insert(0, ".")
import torch
from ipex_llm_b200.transformers import low_bit_linear as lbl
from ipex_llm_b200.transformers.models import common
import ipex_llm_b200.ggml.model.llama.llama_cpp as ggml
dev = torch.device("cuda", 0)
q = lbl.ggml_tensor_qtype["sym_int4"]
N, K, M = 256, 256, 512
w = lbl.FP4Params(torch.randn(N, K, device=dev) * 0.02, qtype=q).to(dev).data
x = torch.randn(M, K, device=dev).half()
out = torch.empty(M, N, device=dev, dtype=torch.float16)
lin = lbl.LowBitLinear(K, N, q, bias=False)
lin.weight = lbl.FP4Params(torch.randn(N, K, device=dev) * 0.02, qtype=q).to(dev)
lin = lin.to(dev)
norm = torch.nn.Module(); norm.weight = torch.ones(K, device=dev, dtype=torch.float16); norm.variance_epsilon = 1e-5


def t(name, fn, n=2000):
    for _ in range(50): fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n): fn()
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    print(f"{name:34s} {(t1 - t0) / n * 1e6:7.2f} us per call (host)", flush=True)


t("lowbit_matmul(out=)", lambda: lbl.lowbit_matmul(x, w, (N, K), q, out=out))
t("lowbit_matmul()", lambda: lbl.lowbit_matmul(x, w, (N, K), q))
t("LowBitLinear.forward", lambda: lin(x))
t("rms_norm_forward", lambda: common.rms_norm_forward(norm, x))
t("torch add", lambda: x + x)
t("torch.empty", lambda: torch.empty(M, N, device=dev, dtype=torch.float16))
t("with torch.cuda.device", lambda: torch.cuda.device(dev).__enter__())
