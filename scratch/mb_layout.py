"""development: relayout (canonical <-> tiled) and dequantise kernel throughput, 12288x4096, CUDA events over rotating buffers"""
import sys; sys.path.insert(0, ".")
import ctypes, torch
from ipex_llm_b200.transformers import low_bit_linear as lbl
import ipex_llm_b200.ggml.model.llama.llama_cpp as ggml
dev = torch.device("cuda", 0)
N, K = 12288, 4096
st = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
P = lambda t: ctypes.c_void_p(t.data_ptr())
def timeit(fn, n=6, reps=5):
    """a CUDA graph of the n calls (no host launch overhead between them), replayed reps times"""
    global st
    for i in range(n): fn(i)
    torch.cuda.synchronize()
    side = torch.cuda.Stream(dev)
    g = torch.cuda.CUDAGraph()
    keep = st
    with torch.cuda.stream(side):
        st = ctypes.c_void_p(side.cuda_stream)
        with torch.cuda.graph(g, stream=side):
            for i in range(n): fn(i)
    st = keep
    g.replay(); torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): g.replay()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / (n * reps) * 1e3
for qt in sys.argv[1:] or ["sym_int4", "asym_int4", "nf4", "fp8_e4m3", "fp6"]:
    q = lbl.ggml_tensor_qtype[qt]
    nbytes = (N * K // 64) * ggml.ggml_type_size(q)
    canon = [torch.randint(0, 256, (nbytes,), dtype=torch.uint8, device=dev) for _ in range(6)]
    w = torch.randn(N, K, device=dev) * 0.02
    for c in canon: ggml.ggml_quantize_tensor_cuda(P(w), P(c), q, N * K, K, None, False, st)
    tiled = [torch.empty_like(c) for c in canon]
    back = torch.empty_like(canon[0])
    us = timeit(lambda i: ggml.lowbit_relayout_cuda(P(canon[i]), P(tiled[i]), N, K, q, True, st))
    print(f"{qt}: to tiled {us:.1f} us, {2 * nbytes / us / 1e6:.2f} TB/s", flush=True)
    us = timeit(lambda i: ggml.lowbit_relayout_cuda(P(tiled[i]), P(back), N, K, q, False, st))
    assert torch.equal(back, canon[5])
    print(f"{qt}: to canonical {us:.1f} us, {2 * nbytes / us / 1e6:.2f} TB/s", flush=True)
    for dt, tdt in ((1, torch.float16), (0, torch.float32)):
        outs = [torch.empty(N, K, dtype=tdt, device=dev) for _ in range(3)]
        us = timeit(lambda i: ggml.lowbit_dequantize_device_layout_cuda(P(tiled[i]), P(outs[i % 3]), N, K, q, dt, st))
        print(f"{qt}: dequantise tiled -> {tdt}: {us:.1f} us, {(nbytes + outs[0].numel() * outs[0].element_size()) / us / 1e6:.2f} TB/s", flush=True)
        us = timeit(lambda i: ggml.ggml_dequantize_cuda(P(canon[i]), P(outs[i % 3]), N * K, q, dt, st))
        print(f"{qt}: dequantise canonical -> {tdt}: {us:.1f} us, {(nbytes + outs[0].numel() * outs[0].element_size()) / us / 1e6:.2f} TB/s", flush=True)
