#!/usr/bin/env python
"""bench.py -- sym_int4 decode tok/s (Llama-2-7B, bs1) on the LowBitLinear GEMV stream.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Workload (BASELINE.json configs[0]/metric): Llama-2-7B architecture (hidden 4096, inter 11008,
32 layers, vocab 32000), random-init fp32 weights quantised to sym_int4 by our GPU quantiser,
batch 1 decode.  One STEP = one decoded token's worth of the hot path: the 161 LowBitLinear
forward calls of the reference's call stack (per layer qkv_proj 12288x4096, o_proj
4096x4096, gate_proj/up_proj 11008x4096, down_proj 4096x11008; then lm_head 32000x4096), 3.51 GB
of packed weights read once.  Attention / RMSNorm / RoPE kernels are outside the hot path this
repo replaces and are not part of the step; the number is therefore the decode rate
the low-bit linear path alone supports and is labelled so in `config.workload`.

value : steps/s with activations and weights resident in HBM, the step replayed from a CUDA graph.
e2e   : the same step driven through the public Python API (LowBitLinear.forward per call), the
        token's activation copied from pinned host memory at the start of every step and the
        lm_head logits' argmax read back to the host at the end (what a generate() loop does).
roofline : dominant kernel = gemv_kernel; achieved = algorithmic bytes of the 161 launches / step time.
cpu_baseline / --impl reference : the CPU oracle in ggml-faithful mode A (int8 activations,
        AVX512-VNNI when the host has it, OpenMP) on a bounded sample (one decoder layer + lm_head
        scaled to the full step) -- the reference's own CPU binary is closed source and absent.
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


def workload_config():
    """the workload both arms (ours and --impl reference) measure"""
    return {"workload": "Llama-2-7B sym_int4 bs1 decode, LowBitLinear stream only: 161 matrices/token "
                        "(32x[qkv 12288x4096, o 4096x4096, gate 11008x4096, up 11008x4096, down 4096x11008] + lm_head "
                        "32000x4096); attention/norm/rope not included",
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
    mats = []
    for name, n, k in LAYER_SHAPES + [LM_HEAD]:
        w = (rng.standard_normal((n, k)) * 0.02).astype(np.float32)
        mats.append((name, n, k, oracle.quantize(w, QTYPE_NAME)))
    x = {HIDDEN: rng.standard_normal((1, HIDDEN)).astype(np.float32), INTER: rng.standard_normal((1, INTER)).astype(np.float32)}

    def layer_pass():
        t0 = time.perf_counter()
        for name, n, k, p in mats[:-1]:
            oracle.matmul_mode_a(p, (n, k), QTYPE_NAME, x[k])
        t1 = time.perf_counter()
        name, n, k, p = mats[-1]
        oracle.matmul_mode_a(p, (n, k), QTYPE_NAME, x[k])
        t2 = time.perf_counter()
        return t1 - t0, t2 - t1

    for _ in range(max(1, warmup)):
        layer_pass()
    tl, th = [], []
    for _ in range(max(1, steps)):
        a, b = layer_pass()
        tl.append(a); th.append(b)
    t_layer, t_head = sum(tl) / len(tl), sum(th) / len(th)
    t_token = LAYERS * t_layer + t_head
    return {"value": 1.0 / t_token, "ms_per_step": t_token * 1e3, "cores": cores,
            "sample": f"one decoder layer (5 GEMVs, {sum(len(m[3]) for m in mats[:-1]) / 1e6:.1f} MB packed) x{LAYERS} + "
                      f"lm_head, mode-A int8-activation kernel, vnni={oracle.uses_vnni()}, {max(1, steps)} passes"}


# --------------------------------------------------------------------------- our arm
def build_model(dev):
    """Quantise random-init Llama-2-7B linears on the GPU into the tiled device layout."""
    import torch
    from ipex_llm_b200.transformers import low_bit_linear as lbl
    qt = lbl.ggml_tensor_qtype[QTYPE_NAME]
    gen = torch.Generator(device=dev).manual_seed(0)
    mods = []
    shapes = [s for _ in range(LAYERS) for s in LAYER_SHAPES] + [LM_HEAD]
    for name, n, k in shapes:
        lin = lbl.LowBitLinear(k, n, qt, bias=False)
        w = torch.empty(n, k, device=dev, dtype=torch.float32).normal_(0.0, 0.02, generator=gen)
        lin.weight = lbl.FP4Params(w, qtype=qt).to(dev)          # quantise + relayout on the GPU
        del w
        lin.to(dev)
        mods.append((name, lin))
    return mods


def run_ours(args):
    import torch
    import torch.distributed as dist
    from ipex_llm_b200.transformers import low_bit_linear as lbl
    import ipex_llm_b200.ggml.model.llama.llama_cpp as ggml

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    ggml.lowbit_set_programmatic_dependent_launch(1 if args.pdl else 0)

    mods = build_model(dev)
    torch.manual_seed(1)
    x_h = torch.randn(1, HIDDEN, device=dev).half()
    x_i = torch.randn(1, INTER, device=dev).half()
    outs = {n: torch.empty(1, n, device=dev, dtype=torch.float16) for n in {m[1].out_len for m in mods}}
    x_host = torch.randn(1, HIDDEN).half().pin_memory()
    tok_host = torch.zeros(1, dtype=torch.int64).pin_memory()

    # the order of the linears in a decoder is static: each one names its successor so that the kernel can
    # ask the L2 to prefetch the next weights (the last one wraps to the first: next token)
    if args.l2_hint:
        for i, (name, lin) in enumerate(mods):
            lin.next_linear = mods[(i + 1) % len(mods)][1]

    mlp_out = torch.empty(1, INTER, device=dev, dtype=torch.float16)

    def step_kernels():
        # resident-input step: raw kernel entry point, outputs into preallocated buffers
        for i, (name, lin) in enumerate(mods):
            if FUSED_MLP and name == "gate_proj":
                up = mods[i + 1][1]
                lbl.mlp_forward(x_h, lin.weight.data, up.weight.data, 1, HIDDEN, INTER, lbl.SILU, lin.qtype, out=mlp_out)
                continue
            if FUSED_MLP and name == "up_proj":
                continue
            x = x_i if lin.in_len == INTER else x_h
            nxt = lin.next_linear.weight.data if lin.next_linear is not None else None
            lbl.lowbit_matmul(x, lin.weight.data, lin.weight_shape, lin.qtype, out=outs[lin.out_len], next_weight=nxt)

    # e2e: the public API a generate() loop uses -- LowBitLinear.forward per linear, recorded once by
    # ipex_llm_b200.transformers.cuda_graph.capture_decode_step and replayed per token.  Every step
    # copies the token's activation from pinned host memory and reads the argmax token back.
    from ipex_llm_b200.transformers.cuda_graph import capture_decode_step
    x_e2e = torch.zeros(1, HIDDEN, device=dev, dtype=torch.float16)
    tok_dev = torch.zeros(1, dtype=torch.int64, device=dev)

    def forward_api():
        # the token's activation arrives from pinned host memory and the chosen token goes back to it: both copies
        # are recorded with the step (memcpy nodes of the graph), so one launch per token covers them
        x_e2e.copy_(x_host, non_blocking=True)              # H2D, 8 KB
        y = None
        for i, (name, lin) in enumerate(mods):
            if FUSED_MLP and name == "gate_proj":
                lbl.mlp_forward(x_e2e, lin.weight.data, mods[i + 1][1].weight.data, 1, HIDDEN, INTER, lbl.SILU, lin.qtype)
                continue
            if FUSED_MLP and name == "up_proj":
                continue
            y = lin(x_i if lin.in_len == INTER else x_e2e)
        tok_dev.copy_(torch.argmax(y, dim=-1))
        tok_host.copy_(tok_dev, non_blocking=True)          # D2H, 8 B
        return tok_dev

    graphed = capture_decode_step(forward_api)

    def step_e2e():
        graphed.replay()                                    # H2D copy, 129 launches, argmax, D2H copy
        graphed.stream.synchronize()
        return int(tok_host[0])

    stream = torch.cuda.Stream()
    with torch.cuda.stream(stream):
        for _ in range(2):
            step_kernels()
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=stream):
            step_kernels()
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(3, args.warmup)):
        graph.replay()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local_rank) as clocks:
        e0.record()
        for _ in range(args.steps):
            graph.replay()                      # 3.5 GB of distinct weights per step >> 126 MB L2
        e1.record()
        barrier()
        ms_dev = e0.elapsed_time(e1)
        # e2e leg
        for _ in range(max(3, args.warmup)):
            step_e2e()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            step_e2e()
        barrier()
        ms_e2e = (time.perf_counter() - t0) * 1e3
    if world > 1:
        t = torch.tensor([ms_dev, ms_e2e], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_dev, ms_e2e = float(t[0]), float(t[1])
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    ms_step = ms_dev / args.steps
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:  # noqa: BLE001
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    achieved = algorithmic_bytes() / (ms_step * 1e-3) / 1e9
    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        c = cpu_reference(steps=3, warmup=1)
        cpu = {"value": c["value"], "unit": "tok/s", "cores": c["cores"], "kind": "port", "sample": c["sample"]}
    # DRAM traffic of the dominant kernel from the committed `ncu --set full` capture (profiles/): the capture is
    # of one gate_proj launch (11008x4096, 23,953,408 algorithmic bytes); scaled to the average launch of the step
    traffic = None
    try:
        prof = json.load(open(os.path.join(ROOT, "profiles", "ncu_full_gemv_r1_final_summary.json")))[0]
        mb = float(prof["dram__bytes_read.sum"]) * 1e6 + float(prof["dram__bytes_write.sum"])
        traffic = mb / (INTER * (HIDDEN // 64) * BLK) * algorithmic_bytes() / n_launches()
    except Exception:  # noqa: BLE001
        pass
    line = {
        "metric": "sym_int4 decode tok/s (Llama-2-7B, bs1); LowBitLinear GEMV GB/s vs HBM peak",
        "value": world * 1e3 / ms_step, "unit": "tok/s", "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup),
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u4 codes x f16 activations, f32 accumulate", "data": "synthetic",
        "config": dict(workload_config(), cache="inputs larger than L2 (3.51 GB of weights per step)", cuda_graph=True,
                       pdl=bool(args.pdl), fused_mlp=FUSED_MLP, l2_next_weight_hint=bool(args.l2_hint),
                       parallelism="replicas" if world > 1 else "single"),
        "e2e": {"value": world * 1e3 / (ms_e2e / args.steps), "unit": "tok/s", "h2d_bytes_per_step": HIDDEN * 2,
                "d2h_bytes_per_step": 8},
        "gpu_launches": n_launches() * args.steps * 2,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": traffic, "traffic_source": "profiles/ncu_full_gemv_r1_final_summary.json: dram read+write of one "
                                                            "gate_proj launch = 1.0011 x its algorithmic bytes, scaled to the average launch",
                     "peak_source": "MEASURED_PEAKS.json hbm_gbs (copy)" if peaks else "fallback 6650",
                     "kernel": "lowbit::gemv_kernel<sym_int4,f16,M=1>", "bytes_per_launch_avg": algorithmic_bytes() / n_launches()},
        "clocks": clocks.summary(),
    }
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
        "config": dict(workload_config(), arm="CPU restatement of the reference's ggml path (closed binary absent), "
                                              "all host threads, bounded sample per step"),
        "cpu_baseline": {"value": c["value"], "unit": "tok/s", "cores": c["cores"], "kind": "port", "sample": c["sample"]},
        "e2e": {"value": c["value"], "unit": "tok/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--pdl", type=int, default=1)
    ap.add_argument("--l2-hint", type=int, default=0, dest="l2_hint")
    ap.add_argument("--no-cpu-baseline", action="store_true")
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
