"""One small call of every kernel family, for compute-sanitizer (memcheck / racecheck / synccheck):
    compute-sanitizer --tool racecheck python scratch/sanitize_driver.py
Sizes are small (the tools slow kernels down 10-100x) but cover every code path the tests name: GEMV M = 1, 8, 16, 32
(cluster push / ring receive / pull reductions), fused MLP, one-token MoE, tensor-core GEMM incl. the K-split path,
quantise, relayout, decode attention (separate kernels and the one-launch cluster kernel), RMSNorm, rotary + cache write,
greedy pick."""
import math
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import ipex_llm_b200.ggml.model.llama.llama_cpp as ggml
from ipex_llm_b200.transformers import low_bit_linear as lbl, kv

dev = torch.device("cuda", 0)
st = lambda: torch.cuda.current_stream(dev).cuda_stream     # noqa: E731
gen = torch.Generator(device=dev).manual_seed(0)
ggml.lowbit_set_programmatic_dependent_launch(int(os.environ.get("SAN_PDL", "1")))


def weight(n, k, qt):
    w = torch.empty(n, k, device=dev).normal_(0.0, 0.02, generator=gen)
    return lbl.FP4Params(w, qtype=lbl.ggml_tensor_qtype[qt]).to(dev).data


for qt in ("sym_int4", "asym_int4", "nf4", "fp8_e4m3", "fp6"):
    N, K = 512, 1024
    w = weight(N, K, qt)
    q = lbl.ggml_tensor_qtype[qt]
    for M in (1, 8, 16, 32, 96):
        for dt in (torch.float16,) if M != 1 else (torch.float16, torch.bfloat16, torch.float32):
            x = torch.randn(M, K, device=dev, generator=gen).to(dt)
            y = lbl.lowbit_matmul(x, w, (N, K), q)
    w2 = weight(N, K, qt)
    lbl.mlp_forward(torch.randn(1, K, device=dev, generator=gen).half(), w, w2, 1, K, N, lbl.SILU, q)
    print("matmul", qt, "ok", flush=True)
# one-token MoE
H, I, E = 256, 512, 4
gates = [weight(I, H, "sym_int4") for _ in range(E)]
ups = [weight(I, H, "sym_int4") for _ in range(E)]
downs = [weight(H, I, "sym_int4") for _ in range(E)]
tab = lambda ws: torch.tensor([w.data_ptr() for w in ws], dtype=torch.int64, device=dev)     # noqa: E731
lbl.moe_forward_vec(torch.randn(1, H, device=dev, generator=gen).half(), torch.tensor([[2, 0]], device=dev),
                    torch.tensor([[0.6, 0.4]], device=dev), tab(gates), tab(ups), tab(downs), H, I, 2)
print("moe ok", flush=True)
# decode attention: separate kernels and the one-launch cluster kernel
B, Hq, Hkv, D, cap = 1, 8, 2, 128, 512
tot = (Hq + 2 * Hkv) * D
k = torch.randn(B, Hkv, cap, D, device=dev, generator=gen).to(torch.float8_e5m2).view(torch.uint8)
v = torch.randn(B, Hkv, cap, D, device=dev, generator=gen).to(torch.float8_e5m2).view(torch.uint8)
ang = torch.arange(cap, device=dev, dtype=torch.float32)[:, None] * (1.0 / (10000.0 ** (torch.arange(0, D, 2, device=dev) / D)))[None, :]
emb = torch.cat([ang, ang], -1)
cos, sin = emb.cos().half(), emb.sin().half()
ws = torch.empty(max(256, ggml.lowbit_decode_attention_workspace_size(B, Hq, Hkv, cap, D)), dtype=torch.uint8, device=dev)
for P in (3, 200, 400):
    state = torch.tensor([P, 0, 0, 0, 0, 0, 0, 0], dtype=torch.int32, device=dev)
    qkv = torch.randn(B, 1, tot, device=dev, generator=gen).half()
    out = torch.empty(B, Hq, 1, D, dtype=torch.float16, device=dev)
    ggml.lowbit_decode_attention(qkv, Hq, Hkv, D, cos, sin, state, k, v, 1 / math.sqrt(D), out, ws, st())
    ggml.lowbit_rope_kv_append(qkv, Hq, Hkv, D, cos, sin, state, 0, k, v, st())
    qv = qkv.view(B, 1, Hq + 2 * Hkv, D).transpose(1, 2)[:, :Hq]
    ggml.lowbit_sdp_fp8_devlen(qv, k, v, False, 1 / math.sqrt(D), state, 1, out, ws, st())
    kv.sdp_fp8(qv, k[:, :, :P + 1], v[:, :, :P + 1], None, 1 / math.sqrt(D))
print("attention ok", flush=True)
x = torch.randn(3, 1024, device=dev, generator=gen).half()
ggml.lowbit_rms_norm(x, torch.ones(1024, device=dev).half(), 1e-5, torch.empty_like(x), st())
ggml.lowbit_rms_norm(x[:, :1000], torch.ones(1000, device=dev).half(), 1e-5, torch.empty(3, 1000, device=dev, dtype=torch.float16), st())
state = torch.zeros(8, dtype=torch.int32, device=dev)
ggml.lowbit_greedy_advance(torch.randn(2, 32000, device=dev, generator=gen).half(), torch.zeros(2, dtype=torch.int64, device=dev),
                           torch.zeros(2, 8, dtype=torch.int64, device=dev), state, st())
print("small ops ok", flush=True)
torch.cuda.synchronize()
print("all ok")
