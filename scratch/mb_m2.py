"""development: decode GEMV at M = 9..32 against cluster size (LOWBIT_GEMV_SPLIT) and activation prefetch distance
(LOWBIT_GEMV_XPF); one process per setting (both are read once)."""
import os, subprocess, sys
if len(sys.argv) > 1 and sys.argv[1] == "child":
    sys.path.insert(0, ".")
    import torch
    from ipex_llm_b200.transformers import low_bit_linear as lbl
    import ipex_llm_b200.ggml.model.llama.llama_cpp as ggml
    dev = torch.device("cuda", 0)
    ggml.lowbit_set_programmatic_dependent_launch(1)

    def bench(N, K, M, qt=2, xdt=torch.float16):
        nbytes = N * (K // 64) * ggml.ggml_type_size(qt)
        ws = [torch.randint(0, 255, (nbytes,), dtype=torch.uint8, device=dev) for _ in range(12)]
        x = torch.randn(M, K, device=dev).to(xdt); out = torch.empty(M, N, device=dev, dtype=xdt)
        st = torch.cuda.Stream()
        with torch.cuda.stream(st):
            for w in ws: lbl.lowbit_matmul(x, w, (N, K), qt, out=out)
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=st):
                for w in ws: lbl.lowbit_matmul(x, w, (N, K), qt, out=out)
            e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
            g.replay(); e0.record()
            for _ in range(5): g.replay()
            e1.record(); torch.cuda.synchronize()
        return e0.elapsed_time(e1) / 5 / len(ws) * 1e3
    for (N, K) in ((11008, 4096), (4096, 11008), (4096, 4096)):
        print(f"  {N}x{K}", " ".join(f"M={M}:{bench(N, K, M):6.1f}" for M in (9, 16, 24, 32)), flush=True)
    sys.exit(0)
for xpf in (0, 4):
    for split in (0, 1, 2, 3, 4, 5, 6, 8):
        env = dict(os.environ, LOWBIT_GEMV_XPF=str(xpf))
        if split:
            env["LOWBIT_GEMV_SPLIT"] = str(split)
        print(f"== xpf {xpf} split {split or 'auto'}", flush=True)
        subprocess.run([sys.executable, __file__, "child"], env=env)
