"""development: the skinny tensor-core GEMM (9 <= M <= 128) against float64 on the dequantised weights, and its time
   python scratch/test_skinny.py [check|time]"""
import sys; sys.path.insert(0, ".")
import torch
from ipex_llm_b200.transformers import low_bit_linear as lbl
import ipex_llm_b200.ggml.model.llama.llama_cpp as ggml
dev = torch.device("cuda", 0)
mode = sys.argv[1] if len(sys.argv) > 1 else "check"
gen = torch.Generator(device=dev).manual_seed(0)


def weights(N, K, qt):
    q = lbl.ggml_tensor_qtype[qt]
    w = torch.empty(N, K, device=dev).normal_(0.0, 0.02, generator=gen)
    packed = lbl.ggml_convert_qtype(w, q, device=dev)
    tiled = lbl.ggml_q_format_convet_cpu2xpu(packed, N * K, q, shape=(N, K))
    return packed, tiled, q


if mode == "check":
    worst = 0.0
    for qt in ("sym_int4", "asym_int4", "nf4", "fp8_e4m3", "fp8_e5m2", "fp6"):
        for N, K in ((256, 1024), (4096, 4096), (1280, 11008), (11008, 4096)):
            packed, tiled, q = weights(N, K, qt)
            wd = lbl.ggml_convert_fp32(packed, (N, K), N * K, q).double()
            for dt in (torch.float16, torch.bfloat16):
                for M in (9, 16, 17, 32, 33, 64, 100, 128):
                    x = torch.randn(M, K, device=dev, generator=gen).to(dt)
                    bias = torch.randn(N, device=dev, generator=gen).to(dt)
                    base = torch.randn(M, N, device=dev, generator=gen).to(dt)
                    out = base.clone()
                    lbl.lowbit_matmul(x, tiled, (N, K), q, out=out, bias=bias, accumulate=True)
                    want = x.double() @ wd.T + bias.double() + base.double()
                    err = float((out.double() - want).norm() / want.norm())
                    lim = 4e-3 if dt == torch.bfloat16 else 6e-4
                    worst = max(worst, err / lim)
                    if err > lim:
                        print("FAIL", qt, N, K, dt, M, err); sys.exit(1)
                    o32 = lbl.lowbit_matmul(x, tiled, (N, K), q, out_dtype=torch.float32)
                    e32 = float((o32.double() - x.double() @ wd.T).norm() / (x.double() @ wd.T).norm())
                    if e32 > (3e-3 if dt == torch.bfloat16 else 5e-4):
                        print("FAIL32", qt, N, K, dt, M, e32); sys.exit(1)
        print(qt, "ok", flush=True)
    print("all ok, worst err/limit", worst)
else:
    import os
    ggml.lowbit_set_programmatic_dependent_launch(int(os.environ.get('PDL', '1')))
    for N, K in ((11008, 4096), (4096, 4096)):
        ws = [weights(N, K, "sym_int4")[1] for _ in range(10)]
        q = lbl.ggml_tensor_qtype["sym_int4"]
        for M in (16, 17, 32, 64):
            x = torch.randn(M, K, device=dev, generator=gen).half()
            out = torch.empty(M, N, device=dev, dtype=torch.float16)
            st = torch.cuda.Stream()
            with torch.cuda.stream(st):
                for w in ws: lbl.lowbit_matmul(x, w, (N, K), q, out=out)
                torch.cuda.synchronize()
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=st):
                    for w in ws: lbl.lowbit_matmul(x, w, (N, K), q, out=out)
                e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
                g.replay(); e0.record()
                for _ in range(5): g.replay()
                e1.record(); torch.cuda.synchronize()
            us = e0.elapsed_time(e1) / 5 / len(ws) * 1e3
            print(f"{N}x{K} M={M}: {us:.2f} us ({N * K * 0.53125 / us / 1e6:.2f} TB/s), launches {ggml.last_launch_count()}", flush=True)
