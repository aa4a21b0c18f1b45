"""development: per-item timestamps of CTA 0 of the skinny GEMM (library built with -DLOWBIT_SK_TIMESTAMPS)"""
import sys, ctypes; sys.path.insert(0, ".")
import numpy as np, torch
from ipex_llm_b200.transformers import low_bit_linear as lbl
import ipex_llm_b200.ggml.model.llama.llama_cpp as ggml
dev = torch.device("cuda", 0)
gen = torch.Generator(device=dev).manual_seed(0)
N, K, M = 11008, 4096, 32
q = lbl.ggml_tensor_qtype["sym_int4"]
w = torch.empty(N, K, device=dev).normal_(0.0, 0.02, generator=gen)
tiled = lbl.ggml_q_format_convet_cpu2xpu(lbl.ggml_convert_qtype(w, q, device=dev), N * K, q, shape=(N, K))
x = torch.randn(M, K, device=dev, generator=gen).half()
out = torch.empty(M, N, device=dev, dtype=torch.float16)
for _ in range(5): lbl.lowbit_matmul(x, tiled, (N, K), q, out=out)
torch.cuda.synchronize()
buf = (ctypes.c_ulonglong * (6 * 64))()
ggml._lib.lowbit_debug_skinny_timestamps(buf)
ts = np.array(buf[:], dtype=np.int64).reshape(6, 64)
t0 = ts[ts > 0].min()
names = ["prod issue", "mma wake", "mma commit", "deq wake", "deq pre-fence", "deq arrived"]
n = int((ts[1] > 0).sum())
print("items", n)
for it in range(min(n, 20)):
    print(it, " ".join(f"{names[r]}={ts[r, it] - t0 if ts[r, it] else -1:6d}" for r in range(6)))
