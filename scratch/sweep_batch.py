"""development: config 3 (Mistral-7B shapes, fp8_e4m3 / fp6 / sym_int4): one decoder layer's linears at batch 1..32,
PDL chain in a CUDA graph over 6 layers of distinct weights.  python scratch/sweep_batch.py"""
import sys; sys.path.insert(0, ".")
import torch
from ipex_llm_b200.transformers import low_bit_linear as lbl
import ipex_llm_b200.ggml.model.llama.llama_cpp as ggml
dev = torch.device("cuda", 0)
gen = torch.Generator(device=dev).manual_seed(0)
H, I, QKV = 4096, 14336, 6144
ggml.lowbit_set_programmatic_dependent_launch(1)
LAYERS = 6


def tiled(N, K, q):
    w = torch.empty(N, K, device=dev).normal_(0.0, 0.02, generator=gen)
    return lbl.ggml_q_format_convet_cpu2xpu(lbl.ggml_convert_qtype(w, q, device=dev), N * K, q, shape=(N, K))


print("| qtype | M | layer us | weights TB/s | us per token |")
print("|---|---|---|---|---|")
for qt in (sys.argv[1:] or ["fp8_e4m3", "fp6", "sym_int4"]):
    q = lbl.ggml_tensor_qtype[qt]
    ws = [dict(qkv=tiled(QKV, H, q), o=tiled(H, H, q), gate=tiled(I, H, q), up=tiled(I, H, q), down=tiled(H, I, q)) for _ in range(LAYERS)]
    nbytes = sum(t.numel() * t.element_size() for t in ws[0].values())
    for M in (1, 2, 4, 8, 9, 12, 16, 17, 24, 32):
        x = torch.randn(M, H, device=dev, generator=gen).half()
        qkv = torch.empty(M, QKV, device=dev, dtype=torch.float16)
        h = torch.zeros(M, H, device=dev, dtype=torch.float16)
        act = torch.empty(M, I, device=dev, dtype=torch.float16)

        def layer(w):
            lbl.lowbit_matmul(x, w["qkv"], (QKV, H), q, out=qkv)
            lbl.lowbit_matmul(x, w["o"], (H, H), q, out=h)
            lbl.mlp_forward(x, w["gate"], w["up"], M, H, I, 0, q, out=act)
            lbl.lowbit_matmul(act, w["down"], (H, I), q, out=h)

        st = torch.cuda.Stream()
        with torch.cuda.stream(st):
            for w in ws: layer(w)
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=st):
                for w in ws: layer(w)
            e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
            g.replay(); e0.record()
            for _ in range(5): g.replay()
            e1.record(); torch.cuda.synchronize()
        us = e0.elapsed_time(e1) / 5 / LAYERS * 1e3
        print(f"| {qt} | {M} | {us:.1f} | {nbytes / us / 1e6:.2f} | {us / M:.2f} |", flush=True)
    del ws
