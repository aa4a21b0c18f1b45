"""development: relative L2 error of the matmul paths against float64 on Llama-2-7B shapes (random N(0, 0.02^2) weights)"""
import sys; sys.path.insert(0, ".")
import torch
from ipex_llm_b200.transformers import low_bit_linear as lbl
dev = torch.device("cuda", 0)
gen = torch.Generator(device=dev).manual_seed(0)
for qt in ("sym_int4", "asym_int4", "nf4", "fp8_e4m3"):
    q = lbl.ggml_tensor_qtype[qt]
    for (N, K) in ((4096, 4096), (11008, 4096), (4096, 11008)):
        w = torch.empty(N, K, device=dev).normal_(0.0, 0.02, generator=gen)
        p = lbl.FP4Params(w, qtype=q).to(dev)
        wd = lbl.dequantize_device_layout(p.data, (N, K), q).double()
        out = []
        for M, dt in ((1, torch.float16), (32, torch.float16), (64, torch.float16), (512, torch.float16), (512, torch.bfloat16), (512, torch.float32)):
            x = torch.randn(M, K, device=dev, generator=gen).to(dt)
            y = lbl.lowbit_matmul(x, p.data, (N, K), q).double()
            ref = x.double() @ wd.t()
            out.append(f"M={M} {str(dt)[6:]}: {float((y - ref).norm() / ref.norm()):.2e}")
        print(qt, N, K, " | ".join(out), flush=True)
