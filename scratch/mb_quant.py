"""development: quantise kernel throughput (12288x4096 fp32 -> each 4-bit format), CUDA events over 10 launches on rotating inputs"""
import sys; sys.path.insert(0, ".")
import ctypes, torch
from ipex_llm_b200.transformers import low_bit_linear as lbl
import ipex_llm_b200.ggml.model.llama.llama_cpp as ggml
dev = torch.device("cuda", 0)
N, K = 12288, 4096
ws = [torch.randn(N, K, device=dev) * 0.02 for _ in range(6)]
st = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
for qt in sys.argv[1:] or ["sym_int4", "asym_int4", "nf4", "fp8_e4m3", "fp6"]:
    q = lbl.ggml_tensor_qtype[qt]
    nbytes = (N * K // 64) * ggml.ggml_type_size(q)
    dst = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    def run(w):
        ggml.ggml_quantize_tensor_cuda(ctypes.c_void_p(w.data_ptr()), ctypes.c_void_p(dst.data_ptr()), q, N * K, K, None, False, st)
    for w in ws: run(w)
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(2):
        for w in ws: run(w)
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / 12 * 1e3
    print(f"{qt}: {us:.1f} us, {(N * K * 4 + nbytes) / us / 1e6:.2f} TB/s of algorithmic bytes", flush=True)
