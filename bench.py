#!/usr/bin/env python
"""bench.py -- sym_int4 decode tok/s (Llama-2-7B, bs1) through generate(), and the LowBitLinear GEMV GB/s behind it.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Workload (BASELINE.json configs[0] / metric): Llama-2-7B architecture (hidden 4096, inter 11008, 32 layers, vocab
32000), random-init, converted by ipex_llm.transformers.AutoModelForCausalLM.from_model(load sym_int4: merge_qkv,
ggml_quantize_tensor on the GPU, optimised forwards), batch 1, greedy, 512-token prompt.  One STEP = one decoded token
of generate() after the first (the reference's "2+ avg latency", utils/benchmark_util_4_47.py:3391-3393): every linear
(3.51 GB of packed weights read once), RMSNorm, rotary, FP8 KV-cache write, decode attention over 512+ keys, lm_head and
the greedy pick, replayed from the recorded decode step.  Default K = 127 = tokens 2..128 of the 512-in / 128-out case.

value : K / (device time of the K decode steps of one generate() call, CUDA events), mean of 3 trials, x N replicas.
e2e   : the same call with the prompt in pinned host memory and a streamer: every token goes to the host as it is
        produced and is fed back from the host (8 B each way per step, copy nodes of the step's graph), host clock.
roofline : dominant kernel = the dequant-GEMV; achieved = algorithmic bytes of the token's 129 LowBitLinear launches /
        their time when replayed alone (parsed.linear_stream), the kernel's own PDL chain.
parsed.prefill : the tensor-core GEMM on the layer shapes at M = 512 / 2048; parsed.tp (N > 1): Llama-2-70B tensor parallel.
cpu_baseline / --impl reference : the CPU oracle in ggml-faithful mode A (int8 activations, AVX512-VNNI when the host
        has it, OpenMP) on a bounded sample (one decoder layer + lm_head scaled to the token's 161 matrices) -- the
        reference's own CPU binary is closed source and absent.  It covers the linears only (every weight byte of a
        token); attention, norms and rotary are not in the CPU sample.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

HIDDEN, INTER, LAYERS, VOCAB = 4096, 11008, 32, 32000
LAYER_SHAPES = [("qkv_proj", 3 * HIDDEN, HIDDEN), ("o_proj", HIDDEN, HIDDEN), ("gate_proj", INTER, HIDDEN),
                ("up_proj", INTER, HIDDEN), ("down_proj", HIDDEN, INTER)]
LM_HEAD = ("lm_head", VOCAB, HIDDEN)
QTYPE_NAME = "sym_int4"
BLK = 34  # bytes per 64-element block


def algorithmic_bytes():
    per_layer = sum(n * (k // 64) * BLK for _, n, k in LAYER_SHAPES)
    return LAYERS * per_layer + LM_HEAD[1] * (LM_HEAD[2] // 64) * BLK


FUSED_MLP = True   # gate_proj + up_proj + SiLU*mul in one launch (the reference's mlp_forward_xpu decode path)


def n_launches():
    return LAYERS * (len(LAYER_SHAPES) - (1 if FUSED_MLP else 0)) + 1


def workload_config(k_steps):
    """the workload both arms (ours and --impl reference) measure"""
    return {"workload": f"Llama-2-7B architecture, random-init, sym_int4 (QK=64, 34 B blocks), batch 1, greedy, 512-token "
                        f"prompt, {k_steps + 1} new tokens through ipex_llm.transformers.AutoModelForCausalLM.from_model(...)"
                        ".generate(); a step = one decoded token (tokens 2..): every linear, RMSNorm, rotary, FP8 KV cache "
                        "write, decode attention, lm_head and the greedy pick",
            "block": "QK=64, 34 B"}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, index=0):
        self.samples, self.stop_flag, self.index = [], False, index
        self.thread = threading.Thread(target=self.run, daemon=True)

    def run(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={q}",
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                parts = [p.strip() for p in out.strip().split(",")]
                if len(parts) >= 6:
                    self.samples.append(parts)
            except Exception:  # noqa: BLE001
                pass
            time.sleep(0.1)

    def __enter__(self):
        self.thread.start()
        return self

    def __exit__(self, *a):
        self.stop_flag = True
        self.thread.join(timeout=6)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        sm = sorted(int(float(s[0])) for s in self.samples)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(s[2 + i].lower().startswith("active") for s in self.samples)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": int(float(self.samples[0][1])), "reasons": reasons}


# --------------------------------------------------------------------------- CPU reference arm
def cpu_reference(steps, warmup):
    """Time the CPU oracle (mode A) on one decoder layer + lm_head, scale to a full token."""
    import numpy as np
    # torchrun exports OMP_NUM_THREADS=1 to its workers: the CPU arm uses every host core it can schedule on
    # (libgomp reads the variable when the oracle library is loaded)
    try:
        ncores = len(os.sched_getaffinity(0))
    except AttributeError:
        ncores = os.cpu_count() or 1
    for quota_file, period_file in (("/sys/fs/cgroup/cpu.max", None),           # a container's CPU quota can be far
                                    ("/sys/fs/cgroup/cpu/cpu.cfs_quota_us", "/sys/fs/cgroup/cpu/cpu.cfs_period_us")):   # below its visible cores
        try:
            if period_file is None:
                quota, period = open(quota_file).read().split()[:2]
            else:
                quota, period = open(quota_file).read().strip(), open(period_file).read().strip()
            if quota not in ("max", "-1"):
                ncores = max(1, min(ncores, -(-int(quota) // int(period))))
            break
        except (OSError, ValueError):
            continue
    if os.environ.get("LOCAL_RANK") is not None or "OMP_NUM_THREADS" not in os.environ:
        os.environ["OMP_NUM_THREADS"] = str(ncores)
    import oracle
    rng = np.random.default_rng(0)
    cores = int(os.environ["OMP_NUM_THREADS"])
    # CPU_LAYERS distinct decoder layers are walked in turn (3 x 107 MB + the 70 MB lm_head: beyond a server's L3), so
    # that a pass streams its weights from DRAM the way a real 3.5 GB token does instead of re-reading a cached layer
    CPU_LAYERS = 3
    base = [(rng.standard_normal((n, k)) * 0.02).astype(np.float32) for _, n, k in LAYER_SHAPES]
    layers = []
    for li in range(CPU_LAYERS):
        layers.append([(name, n, k, oracle.quantize(np.roll(w, 977 * li, axis=0) if li else w, QTYPE_NAME))
                       for (name, n, k), w in zip(LAYER_SHAPES, base)])
    del base
    name, n, k = LM_HEAD
    head = (name, n, k, oracle.quantize((rng.standard_normal((n, k)) * 0.02).astype(np.float32), QTYPE_NAME))
    mats = layers[0] + [head]
    x = {HIDDEN: rng.standard_normal((1, HIDDEN)).astype(np.float32), INTER: rng.standard_normal((1, INTER)).astype(np.float32)}
    turn = [0]

    def layer_pass():
        layer = layers[turn[0] % CPU_LAYERS]
        turn[0] += 1
        t0 = time.perf_counter()
        for name, n, k, p in layer:
            oracle.matmul_mode_a(p, (n, k), QTYPE_NAME, x[k])
        t1 = time.perf_counter()
        name, n, k, p = head
        oracle.matmul_mode_a(p, (n, k), QTYPE_NAME, x[k])
        t2 = time.perf_counter()
        return t1 - t0, t2 - t1

    for _ in range(max(1, warmup)):
        layer_pass()
    tl, th = [], []
    t_begin = time.perf_counter()
    for _ in range(max(1, steps)):
        a, b = layer_pass()
        tl.append(a); th.append(b)
        if time.perf_counter() - t_begin > 60.0:            # bounded sample: a slow or busy host stops after a minute
            break
    t_layer, t_head = sum(tl) / len(tl), sum(th) / len(th)
    t_token = LAYERS * t_layer + t_head
    return {"value": 1.0 / t_token, "ms_per_step": t_token * 1e3, "cores": cores,
            "sample": f"one decoder layer per pass (5 GEMVs, {sum(len(m[3]) for m in mats[:-1]) / 1e6:.1f} MB packed; {CPU_LAYERS} distinct "
                      f"layers in turn) x{LAYERS} + lm_head, mode-A int8-activation kernel, vnni={oracle.uses_vnni()}, {len(tl)} passes"}


# --------------------------------------------------------------------------- our arm
PROMPT_LEN, NEW_TOKENS = 512, 128      # BASELINE.json configs[0]: 512-in / 128-out


def build_llama(dev, layers=LAYERS):
    """random-init Llama-2-7B (N(0, 0.02^2), fp16) built on the GPU and converted through the public loader API:
    AutoModelForCausalLM.from_model -> merge_qkv, ggml_quantize_tensor on the GPU, optimised forwards"""
    import torch
    import transformers
    from ipex_llm_b200.optimize import low_memory_init
    from ipex_llm.transformers import AutoModelForCausalLM
    cfg = transformers.LlamaConfig(hidden_size=HIDDEN, intermediate_size=INTER, num_hidden_layers=layers,
                                   num_attention_heads=32, num_key_value_heads=32, vocab_size=VOCAB,
                                   max_position_embeddings=4096, rms_norm_eps=1e-5, tie_word_embeddings=False)
    with low_memory_init():
        model = transformers.LlamaForCausalLM(cfg)
    gen = torch.Generator(device=dev).manual_seed(0)
    for mod in model.modules():
        for pname, p in list(mod._parameters.items()):
            if p is None:
                continue
            if pname == "weight" and mod.__class__.__name__.endswith("RMSNorm"):
                t = torch.ones(p.shape, device=dev, dtype=torch.float16)
            else:
                t = torch.empty(p.shape, device=dev, dtype=torch.float16).normal_(0.0, 0.02, generator=gen)
            mod._parameters[pname] = torch.nn.Parameter(t, requires_grad=False)
    for mod in model.modules():                                  # non-persistent buffers (rotary inv_freq) follow
        for bname, b in list(mod._buffers.items()):
            if b is not None and b.device.type != "cuda":
                mod._buffers[bname] = b.to(dev)
    model = AutoModelForCausalLM.from_model(model.eval(), low_bit=QTYPE_NAME)
    return model.to(dev).half()


class TokenSink:
    """a transformers-style streamer: the host receives every token as it is produced"""

    def __init__(self):
        self.tokens = []

    def put(self, value):
        self.tokens.append(value)

    def end(self):
        pass


def linear_stream(model, dev, steps, warmup):
    """the LowBitLinear launches of one token (qkv, o, fused gate+up, down per layer, lm_head) on constant inputs,
    replayed from a CUDA graph: the GEMV kernel in its PDL chain with nothing else in between -- the number the
    roofline is quoted on.  Returns (ms per token, launches per token)."""
    import torch
    from ipex_llm_b200.transformers import low_bit_linear as lbl
    core = model.model
    x_h = torch.randn(1, HIDDEN, device=dev).half()
    x_i = torch.randn(1, INTER, device=dev).half()
    outs = {}

    def out_for(n):
        if n not in outs:
            outs[n] = torch.empty(1, n, device=dev, dtype=torch.float16)
        return outs[n]

    launches = [0]

    def step():
        launches[0] = 0
        for layer in core.layers:
            a, m = layer.self_attn, layer.mlp
            for lin, x in ((a.qkv_proj, x_h), (a.o_proj, x_h)):
                lbl.lowbit_matmul(x, lin.weight.data, lin.weight_shape, lin.qtype, out=out_for(lin.out_len))
            lbl.mlp_forward(x_h, m.gate_proj.weight.data, m.up_proj.weight.data, 1, HIDDEN, INTER, lbl.SILU, m.gate_proj.qtype,
                            out=out_for(INTER))
            lbl.lowbit_matmul(x_i, m.down_proj.weight.data, m.down_proj.weight_shape, m.down_proj.qtype, out=out_for(HIDDEN))
            launches[0] += 4
        h = model.lm_head
        lbl.lowbit_matmul(x_h, h.weight.data, h.weight_shape, h.qtype, out=out_for(h.out_len))
        launches[0] += 1

    stream = torch.cuda.Stream()
    with torch.cuda.stream(stream):
        for _ in range(2):
            step()
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=stream):
            step()
        torch.cuda.synchronize()
        for _ in range(max(3, warmup)):
            graph.replay()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(steps):
            graph.replay()                                      # 3.5 GB of distinct weights per step >> 126 MB L2
        e1.record()
        torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps, launches[0]


def prefill_numbers(model, dev, peaks):
    """prefill GEMM (kernel (c)) on the layer shapes at config 1's and config 2's prompt lengths, timed alone with CUDA
    events (L2 flushed between launches by the 200+ MB of other weights in the rotation), plus the whole model's
    512-token prefill through the public forward"""
    import torch
    from ipex_llm_b200.transformers import low_bit_linear as lbl
    core = model.model
    burst = float(peaks.get("bf16_tflops", 1644.0))
    out = {"peak_tflops": burst, "peak_source": "MEASURED_PEAKS.json bf16_tflops (burst: kernels timed alone)" if peaks else "fallback"}
    for M in (512, 2048):
        lins, mlps = [], []
        for layer in core.layers[:8]:                            # rotate over 8 layers' weights: each launch reads cold weights
            a, m = layer.self_attn, layer.mlp
            lins += [a.qkv_proj, a.o_proj, m.gate_proj, m.up_proj, m.down_proj]
            mlps.append((a.qkv_proj, a.o_proj, m.gate_proj, m.up_proj, m.down_proj))
        xs = {HIDDEN: torch.randn(M, HIDDEN, device=dev).half(), INTER: torch.randn(M, INTER, device=dev).half()}
        ys = {n: torch.empty(M, n, device=dev, dtype=torch.float16) for n in {l.out_len for l in lins}}

        def run_gemms():
            # kernel (c) alone: every linear of a layer as its own GEMM launch
            for lin in lins:
                lbl.lowbit_matmul(xs[lin.in_len], lin.weight.data, lin.weight_shape, lin.qtype, out=ys[lin.out_len])

        def run_model_path():
            # the linears as the model's forward issues them: qkv, o, gate + up through mlp_forward (ONE grid of the
            # tensor-core kernel for both matrices, then the activation kernel -- its time is inside the figure), down
            for qkv, o, gate, up, down in mlps:
                lbl.lowbit_matmul(xs[HIDDEN], qkv.weight.data, qkv.weight_shape, qkv.qtype, out=ys[qkv.out_len])
                lbl.lowbit_matmul(xs[HIDDEN], o.weight.data, o.weight_shape, o.qtype, out=ys[o.out_len])
                lbl.mlp_forward(xs[HIDDEN], gate.weight.data, up.weight.data, M, HIDDEN, INTER, lbl.SILU, gate.qtype, out=ys[INTER])
                lbl.lowbit_matmul(xs[INTER], down.weight.data, down.weight_shape, down.qtype, out=ys[down.out_len])

        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

        def timed(fn, reps=3):
            for _ in range(2):
                fn()
            torch.cuda.synchronize()
            e0.record()
            for _ in range(reps):
                fn()
            e1.record()
            torch.cuda.synchronize()
            return e0.elapsed_time(e1) / reps

        ms = timed(run_gemms)
        ms_model = timed(run_model_path)
        flops = sum(2.0 * M * l.in_len * l.out_len for l in lins)
        tf = flops / (ms * 1e-3) / 1e12
        out[f"M{M}"] = {"ms_per_layer": ms / 8, "tflops": tf, "frac_of_peak": tf / burst,
                        "tok_per_s_linears_only": M / (ms / 8 * LAYERS * 1e-3),
                        "model_path_ms_per_layer": ms_model / 8,
                        "model_path_note": "qkv, o, mlp_forward (gate + up in one grid + the activation kernel), down"}
        if peaks.get("bf16_tflops_sustained"):          # the loop is 120 GEMMs back to back: also against the sustained figure
            out[f"M{M}"]["frac_of_sustained_peak"] = tf / float(peaks["bf16_tflops_sustained"])
    ids = torch.randint(0, VOCAB, (1, PROMPT_LEN), device=dev)
    with torch.no_grad():
        for _ in range(2):
            model(ids, use_cache=True)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        model(ids, use_cache=True)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
    out["model_prefill_512"] = {"ms": dt * 1e3, "tok_per_s": PROMPT_LEN / dt,
                                "note": "whole forward through the public API, eager launches (FP8 cache write, attention, norms included)"}
    return out


def quantize_numbers(dev, peaks):
    """kernel (a): ggml_quantize_tensor on the GPU, 12288 x 4096 fp32 (the merged qkv projection) -> sym_int4; algorithmic
    bytes = 4 B read + 34/64 B written per element; rotating over 4 source tensors (805 MB > L2)"""
    import ctypes
    import torch
    import ipex_llm_b200.ggml.model.llama.llama_cpp as ggml
    n_rows, k = 3 * HIDDEN, HIDDEN
    srcs = [torch.randn(n_rows, k, device=dev) * 0.02 for _ in range(4)]
    dst = torch.empty(n_rows * (k // 64) * BLK, dtype=torch.uint8, device=dev)
    st = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)

    def run():
        for src in srcs:
            ggml.ggml_quantize_tensor_cuda(ctypes.c_void_p(src.data_ptr()), ctypes.c_void_p(dst.data_ptr()), 2, n_rows * k, k, None, False, st)

    run()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(5):
        run()
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / 20
    gbs = n_rows * k * (4 + BLK / 64) / (us * 1e-6) / 1e9
    peak = float(peaks.get("hbm_gbs", 6650.0))
    out = {"shape": [n_rows, k], "us": us, "GBps": gbs, "frac_of_hbm_peak": gbs / peak}

    # the other load-time kernels of row (a): canonical records -> tiled device layout (2 x packed bytes) and the tiled
    # layout -> fp16 (packed bytes + 2 B per element); four buffers in turn (107 MB of records, 403 MB of fp16)
    packed = [torch.empty_like(dst) for _ in range(4)]
    for src, p_ in zip(srcs, packed):
        ggml.ggml_quantize_tensor_cuda(ctypes.c_void_p(src.data_ptr()), ctypes.c_void_p(p_.data_ptr()), 2, n_rows * k, k, None, False, st)
    del srcs
    tiled = [torch.empty_like(dst) for _ in range(4)]
    dense = [torch.empty(n_rows, k, dtype=torch.float16, device=dev) for _ in range(4)]

    def timed(fn):
        fn()
        torch.cuda.synchronize()
        e0.record()
        for _ in range(5):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) * 1e3 / 20

    def relayout():
        for p_, t_ in zip(packed, tiled):
            ggml.lowbit_relayout_cuda(ctypes.c_void_p(p_.data_ptr()), ctypes.c_void_p(t_.data_ptr()), n_rows, k, 2, True, st)

    def dequant():
        for t_, d_ in zip(tiled, dense):
            ggml.lowbit_dequantize_device_layout_cuda(ctypes.c_void_p(t_.data_ptr()), ctypes.c_void_p(d_.data_ptr()), n_rows, k, 2, 1, st)

    us_r = timed(relayout)
    us_d = timed(dequant)
    nbytes = dst.numel()
    out["relayout_to_tiled"] = {"us": us_r, "GBps": 2 * nbytes / (us_r * 1e-6) / 1e9, "frac_of_hbm_peak": 2 * nbytes / (us_r * 1e-6) / 1e9 / peak}
    out["dequantize_tiled_f16"] = {"us": us_d, "GBps": (nbytes + 2 * n_rows * k) / (us_d * 1e-6) / 1e9,
                                   "frac_of_hbm_peak": (nbytes + 2 * n_rows * k) / (us_d * 1e-6) / 1e9 / peak}
    return out


def tp_numbers(args, dev, world, rank):
    """Llama-2-70B sym_int4 bs1 LowBitLinear stream, column/row sharded over the ranks with the all-reduce after o_proj and
    down_proj (BASELINE.json configs[3]; the reference's AutoTP recipe), and the same stream on ONE GPU measured on rank 0
    in the same run.  CUDA events, max over ranks."""
    import torch
    import torch.distributed as dist
    from ipex_llm_b200.transformers import low_bit_linear as lbl
    from ipex_llm_b200.transformers.tensor_parallel import PeerAllReduce
    H70, I70, L70, KV70 = 8192, 28672, 80, 1024
    qt = lbl.ggml_tensor_qtype["sym_int4"]
    gen = torch.Generator(device=dev).manual_seed(1234 + rank)

    def packed(n, k):
        w = torch.empty(n, k, device=dev, dtype=torch.float32).normal_(0.0, 0.02, generator=gen)
        return lbl.FP4Params(w, qtype=qt).to(dev).data

    def stream_for(tp, peer):
        shapes = {"qkv": ((H70 + 2 * KV70) // tp, H70), "o": (H70, H70 // tp), "gate": (I70 // tp, H70),
                  "up": (I70 // tp, H70), "down": (H70, I70 // tp)}
        layers = [{k: packed(*s) for k, s in shapes.items()} for _ in range(L70)]
        lm_head = packed(VOCAB, H70)
        x_h = (torch.randn(1, H70, device=dev, generator=gen) * 0.1).half()
        x_os = x_h[:, :H70 // tp].contiguous()
        qkv_out = torch.empty(1, shapes["qkv"][0], device=dev, dtype=torch.float16)
        mlp_out = torch.empty(1, I70 // tp, device=dev, dtype=torch.float16)
        hid = torch.empty(1, H70, device=dev, dtype=torch.float16)
        logits = torch.empty(1, VOCAB, device=dev, dtype=torch.float16)

        def row_parallel(x, w, shape):
            if peer is not None:
                peer.matmul_all_reduce(x, w, shape, qt, out=hid)
            else:
                lbl.lowbit_matmul(x, w, shape, qt, out=hid)

        def step():
            for L in layers:
                lbl.lowbit_matmul(x_h, L["qkv"], shapes["qkv"], qt, out=qkv_out)
                row_parallel(x_os, L["o"], shapes["o"])
                lbl.mlp_forward(x_h, L["gate"], L["up"], 1, H70, I70 // tp, lbl.SILU, qt, out=mlp_out)
                row_parallel(mlp_out, L["down"], shapes["down"])
            lbl.lowbit_matmul(x_h, lm_head, (VOCAB, H70), qt, out=logits)

        return step

    def timed(step, collective):
        stream = torch.cuda.Stream()
        with torch.cuda.stream(stream):
            for _ in range(2):
                step()
            torch.cuda.synchronize()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=stream, capture_error_mode="thread_local"):
                step()
            torch.cuda.synchronize()
            if collective:
                dist.barrier()
            for _ in range(max(3, args.warmup)):
                graph.replay()
            torch.cuda.synchronize()
            if collective:
                dist.barrier()
            best = None
            for _ in range(2):                                   # two timed passes, the better one (clock ramps, stragglers)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                if collective:
                    dist.barrier()
                e0.record()
                for _ in range(args.steps):
                    graph.replay()
                e1.record()
                torch.cuda.synchronize()
                ms = e0.elapsed_time(e1) / args.steps
                best = ms if best is None else min(best, ms)
        return best

    peer = PeerAllReduce()
    ms_tp = timed(stream_for(world, peer), True)
    timed_out = peer.timed_out()
    ms_nocomm = timed(stream_for(world, None), False)
    torch.cuda.empty_cache()
    t = torch.tensor([ms_tp, ms_nocomm], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_tp, ms_nocomm = float(t[0]), float(t[1])
    ms_tp1 = None
    if rank == 0:
        ms_tp1 = timed(stream_for(1, None), False)
    dist.barrier()
    peer.close()
    if rank != 0:
        return None
    n_ar = 2 * L70
    return {"model": "Llama-2-70B sym_int4 bs1 LowBitLinear stream (80 layers: qkv, o, fused gate+up, down; lm_head replicated)",
            "nranks": world, "ms_per_token": ms_tp, "tok_per_s": 1e3 / ms_tp, "ms_per_token_no_collectives": ms_nocomm,
            "allreduces_per_token": n_ar, "ms_per_allreduce": (ms_tp - ms_nocomm) / n_ar,
            "allreduce": "one kernel per row-parallel linear: the GEMV's leader CTAs send their fp32 partial sums to the peers' inboxes, "
                         "wait for the peers' words of the same rows and store the rank-ordered sum (lowbit_matmul_allreduce_cuda)",
            "tp1_ms_per_token": ms_tp1, "speedup_vs_tp1": ms_tp1 / ms_tp, "strong_efficiency": ms_tp1 / ms_tp / world,
            "wait_ran_out": bool(timed_out), "timing": "CUDA events over a replayed graph, max over ranks"}


def run_ours(args):
    import torch
    import torch.distributed as dist
    import ipex_llm_b200.ggml.model.llama.llama_cpp as ggml

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ["NCCL_DEBUG"] = os.environ.get("BENCH_NCCL_DEBUG", "WARN")     # no version banner on stdout: one JSON line
        dist.init_process_group("nccl", device_id=dev)
    ggml.lowbit_set_programmatic_dependent_launch(1 if args.pdl else 0)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    model = build_llama(dev)
    K = args.steps                                             # timed decode steps = tokens 2 .. K+1 of one generate() call
    W = max(3, args.warmup)
    prompt_host = torch.randint(0, VOCAB, (1, PROMPT_LEN), generator=torch.Generator().manual_seed(1)).pin_memory()
    prompt_dev = prompt_host.to(dev)
    gen_kwargs = dict(do_sample=False, pad_token_id=0)
    with torch.no_grad():
        # warm-up: one untimed generate() of W decode steps records the step's graph; a second one of the timed shape
        model.generate(prompt_dev, max_new_tokens=W + 1, min_new_tokens=W + 1, **gen_kwargs)
        model.generate(prompt_dev, max_new_tokens=K + 1, min_new_tokens=K + 1, **gen_kwargs)
        sink = TokenSink()
        model.generate(prompt_dev, max_new_tokens=W + 1, min_new_tokens=W + 1, streamer=sink, **gen_kwargs)
        barrier()
        with ClockSampler(local_rank) as clocks:
            # value: inputs resident in HBM; the K decode steps of generate() timed on the device (CUDA events around the
            # K graph replays) between synchronisations -- 1000 / "2+ avg latency" of the reference's BenchmarkWrapper
            trials = []
            for _ in range(3):                                 # num_trials 3 (all-in-one/run.py)
                out_dev = model.generate(prompt_dev, max_new_tokens=K + 1, min_new_tokens=K + 1, **gen_kwargs)
                trials.append((model.rest_cost_mean_device, model.rest_cost_mean, model.first_cost))
            barrier()
            ms_dev = 1e3 * sum(t[0] for t in trials) / len(trials)
            ms_wall = 1e3 * sum(t[1] for t in trials) / len(trials)
            first_ms = 1e3 * sum(t[2] for t in trials) / len(trials)
            launches_per_token = model.last_decode_launches_per_token
            # e2e: the same call with the prompt in pinned HOST memory, every token delivered to the host as it is produced
            # (streamer) and fed back from the host: per step 8 B H2D + 8 B D2H inside the step's graph, host clock
            e2e_trials = []
            for _ in range(3):
                sink = TokenSink()
                t0 = time.perf_counter()
                out = model.generate(prompt_host.to(dev, non_blocking=True), max_new_tokens=K + 1, min_new_tokens=K + 1,
                                     streamer=sink, **gen_kwargs)
                out_host = out.cpu()
                t_total = time.perf_counter() - t0
                e2e_trials.append((model.rest_cost_mean, t_total))
                assert len(sink.tokens) == K + 2 and torch.equal(out_host, out_dev.cpu()), "streamed and resident runs disagree"
            barrier()
            ms_e2e = 1e3 * sum(t[0] for t in e2e_trials) / len(e2e_trials)
            e2e_total_ms = 1e3 * sum(t[1] for t in e2e_trials) / len(e2e_trials)
            # batch 4 and 16 decode (BASELINE.json configs[1]'s batch sizes), same prompt length, 32 decode steps
            batch_rates = {}
            for bsz in (4, 16):
                pb = torch.randint(0, VOCAB, (bsz, PROMPT_LEN), device=dev)
                model.generate(pb, max_new_tokens=9, min_new_tokens=9, **gen_kwargs)
                model.generate(pb, max_new_tokens=33, min_new_tokens=33, **gen_kwargs)
                batch_rates[f"batch{bsz}_tok_per_s"] = bsz / model.rest_cost_mean_device
                batch_rates[f"batch{bsz}_ms_per_step"] = model.rest_cost_mean_device * 1e3
            # the LowBitLinear launches alone (roofline of the GEMV kernel in its PDL chain)
            ms_stream, stream_launches = linear_stream(model, dev, max(K, 20), W)
            barrier()
    if world > 1:
        t = torch.tensor([ms_dev, ms_e2e, ms_stream], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_dev, ms_e2e, ms_stream = float(t[0]), float(t[1]), float(t[2])
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:  # noqa: BLE001
        pass
    prefill = prefill_numbers(model, dev, peaks) if (rank == 0 and not args.no_prefill) else None
    quant = quantize_numbers(dev, peaks) if (rank == 0 and not args.no_prefill) else None
    del model
    torch.cuda.empty_cache()
    tp = None
    if world > 1 and not args.no_tp:
        tp = tp_numbers(args, dev, world, rank)
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    peak = float(peaks.get("hbm_gbs", 6650.0))
    achieved = algorithmic_bytes() / (ms_stream * 1e-3) / 1e9
    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        c = cpu_reference(steps=3, warmup=1)
        cpu = {"value": c["value"], "unit": "tok/s", "cores": c["cores"], "kind": "port", "sample": c["sample"]}
    traffic = None
    try:
        prof = json.load(open(os.path.join(ROOT, "profiles", "ncu_full_gemv_r1_final_summary.json")))[0]
        mb = float(prof["dram__bytes_read.sum"]) * 1e6 + float(prof["dram__bytes_write.sum"])
        traffic = mb / (INTER * (HIDDEN // 64) * BLK) * algorithmic_bytes() / n_launches()
    except Exception:  # noqa: BLE001
        pass
    line = {
        "metric": "sym_int4 decode tok/s (Llama-2-7B, bs1); LowBitLinear GEMV GB/s vs HBM peak",
        "value": world * 1e3 / ms_dev, "unit": "tok/s", "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": ms_dev, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u4 codes x f16 activations, f32 accumulate; e5m2 KV cache", "data": "synthetic",
        "config": dict(workload_config(K), **{"cache": "inputs larger than L2 (3.51 GB of weights per step)", "cuda_graph": True,
                   "pdl": bool(args.pdl), "trials": 3,
                   "parallelism": (f"replicas x{world} (value: independent copies of the 7B model, no data-path collective); "
                                   f"parsed.tp: Llama-2-70B tensor parallel tp{world}") if world > 1 else "single"}),
        "e2e": {"value": world * 1e3 / ms_e2e, "unit": "tok/s", "h2d_bytes_per_step": 8, "d2h_bytes_per_step": 8,
                "how": "generate() with the prompt in pinned host memory and a streamer: every step's token goes to the host and is "
                       "fed back from it (copy nodes of the step's graph), host clock over the decode steps",
                "whole_call_ms": e2e_total_ms, "prompt_h2d_bytes": PROMPT_LEN * 8, "result_d2h_bytes": (PROMPT_LEN + K + 1) * 8},
        "gpu_launches": int(launches_per_token * K * 6 + stream_launches * max(K, 20)),
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": traffic, "traffic_source": "profiles/ncu_full_gemv_r1_final_summary.json: dram read+write of one "
                                                            "gate_proj launch = 1.0011 x its algorithmic bytes, scaled to the average launch",
                     "peak_source": "MEASURED_PEAKS.json hbm_gbs (copy)" if peaks else "fallback 6650",
                     "kernel": "lowbit::gemv_kernel<sym_int4,f16,M=1>", "bytes_per_launch_avg": algorithmic_bytes() / n_launches(),
                     "measured_on": "parsed.linear_stream: the token's 129 LowBitLinear launches replayed alone from a CUDA graph "
                                    "(the kernel's PDL chain without the attention / norm launches in between)",
                     "launch_ms_avg": ms_stream / n_launches()},
        "parsed": {
            "generate": {"first_token_ms": first_ms, "rest_token_ms_device": ms_dev, "rest_token_ms_wall": ms_wall,
                         "kernels_per_token": launches_per_token, **batch_rates,
                         "weights_GBps_in_generate": algorithmic_bytes() / (ms_dev * 1e-3) / 1e9,
                         "frac_of_hbm_peak_in_generate": algorithmic_bytes() / (ms_dev * 1e-3) / 1e9 / peak},
            "linear_stream": {"ms_per_token": ms_stream, "tok_per_s": 1e3 / ms_stream, "launches_per_token": stream_launches,
                              "GBps": achieved, "note": "round 1's headline: LowBitLinear launches only, constant activations"},
        },
        "clocks": clocks.summary(),
    }
    if prefill:
        line["parsed"]["prefill"] = prefill
    if quant:
        line["parsed"]["quantize"] = quant
    if tp:
        line["parsed"]["tp"] = tp
    if cpu:
        line["cpu_baseline"] = cpu
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    c = cpu_reference(steps=args.steps, warmup=args.warmup)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    line = {
        "impl": "reference",
        "metric": "sym_int4 decode tok/s (Llama-2-7B, bs1); LowBitLinear GEMV GB/s vs HBM peak",
        "value": c["value"], "unit": "tok/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": c["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u4 codes x int8 activations (ggml mode A), f32 accumulate", "data": "synthetic",
        "config": dict(workload_config(args.steps),
                       arm="CPU restatement of the reference's ggml path (closed binary absent), all host threads; each step = a "
                           "bounded sample of the token: one decoder layer's five GEMVs x32 + lm_head = the 161 LowBitLinear "
                           "matrices (every weight byte of the token); attention over the KV cache, norms and rotary are NOT in "
                           "the CPU sample (the GPU arm's step includes them)"),
        "cpu_baseline": {"value": c["value"], "unit": "tok/s", "cores": c["cores"], "kind": "port", "sample": c["sample"]},
        "e2e": {"value": c["value"], "unit": "tok/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=127)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--pdl", type=int, default=1)
    ap.add_argument("--l2-hint", type=int, default=0, dest="l2_hint")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-prefill", action="store_true")
    ap.add_argument("--no-tp", action="store_true")
    ap.add_argument("--fused-mlp", type=int, default=1, dest="fused_mlp")
    args = ap.parse_args()
    global FUSED_MLP
    FUSED_MLP = bool(args.fused_mlp)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
