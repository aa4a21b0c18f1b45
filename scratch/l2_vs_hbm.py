"""development: is the GEMV chain bound by HBM?  Ten launches over ten distinct matrices vs ten launches over ONE matrix (L2 hits)"""
import sys; sys.path.insert(0, ".")
import torch
from ipex_llm_b200.transformers import low_bit_linear as lbl
import ipex_llm_b200.ggml.model.llama.llama_cpp as ggml
dev = torch.device("cuda", 0)
ggml.lowbit_set_programmatic_dependent_launch(1)
q = lbl.ggml_tensor_qtype["sym_int4"]
gen = torch.Generator(device=dev).manual_seed(0)
for N, K in ((12288, 4096), (4096, 4096), (4096, 11008), (32000, 4096)):
    nbytes = N * (K // 64) * 34
    distinct = [torch.randint(0, 255, (nbytes,), dtype=torch.uint8, device=dev) for _ in range(12)]
    for name, ws in (("distinct", distinct), ("same", [distinct[0]] * 12)):
        x = torch.randn(1, K, device=dev, generator=gen).half()
        out = torch.empty(1, N, device=dev, dtype=torch.float16)
        st = torch.cuda.Stream()
        with torch.cuda.stream(st):
            for w in ws: lbl.lowbit_matmul(x, w, (N, K), q, out=out)
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=st):
                for w in ws: lbl.lowbit_matmul(x, w, (N, K), q, out=out)
            e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
            g.replay(); e0.record()
            for _ in range(10): g.replay()
            e1.record(); torch.cuda.synchronize()
        us = e0.elapsed_time(e1) / 10 / len(ws) * 1e3
        print(f"{N}x{K} {name:8s}: {us:.2f} us per launch ({nbytes / us / 1e6:.2f} TB/s)", flush=True)
