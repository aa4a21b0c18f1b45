"""Tensor-parallel decode of the Llama-2-70B LowBitLinear stream (BASELINE.json config 4), one process per GPU:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P bench_tp.py

Sharding as in ipex_llm_b200/transformers/tensor_parallel.py (the reference's DeepSpeed-AutoTP recipe): qkv / gate / up
column-parallel (rows / N), o / down row-parallel (K / N) followed by ONE NCCL all-reduce of the [1, 8192] output
(low_bit_linear.py:715-725); lm_head replicated.  A step = one decoded token = 80 x [qkv, o + all-reduce, fused gate+up
MLP, down + all-reduce] + lm_head, captured in a CUDA graph.  Weights: random-init N(0, 0.02^2) quantised on the GPU.
Timing: CUDA events, max over ranks.  Prints one JSON line on rank 0.  (bench.py stays the driver's contract; this
script documents the sharded configuration.)
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

HIDDEN, INTER, LAYERS, VOCAB, KV = 8192, 28672, 80, 32000, 1024   # 64 heads, 8 kv heads x 128


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--layers", type=int, default=LAYERS)
    ap.add_argument("--no-allreduce", action="store_true", help="skip the collectives (shows their cost)")
    ap.add_argument("--no-graph", action="store_true", help="eager launches (no CUDA graph)")
    ap.add_argument("--pdl", type=int, default=1)
    ap.add_argument("--allreduce", choices=["fused", "peer", "nccl"], default="fused",
                    help="fused: the row-parallel GEMVs push their partial sums to the peers from their epilogue; peer: the library's one-shot kernel over IPC-mapped peer buffers (graph-capturable); nccl: torch.distributed")
    args = ap.parse_args()
    import torch
    import torch.distributed as dist
    from ipex_llm_b200.transformers import low_bit_linear as lbl
    import ipex_llm_b200.ggml.model.llama.llama_cpp as ggml

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    ggml.lowbit_set_programmatic_dependent_launch(args.pdl)
    qt = lbl.ggml_tensor_qtype["sym_int4"]
    gen = torch.Generator(device=dev).manual_seed(1234 + rank)

    def packed(n, k):
        w = torch.empty(n, k, device=dev, dtype=torch.float32).normal_(0.0, 0.02, generator=gen)
        return lbl.FP4Params(w, qtype=qt).to(dev).data

    tp = world
    shapes = {"qkv": ((HIDDEN + 2 * KV) // tp, HIDDEN), "o": (HIDDEN, HIDDEN // tp), "gate": (INTER // tp, HIDDEN),
              "up": (INTER // tp, HIDDEN), "down": (HIDDEN, INTER // tp)}
    assert (INTER // tp) % 64 == 0 and (HIDDEN // tp) % 64 == 0
    layers = [{k: packed(*s) for k, s in shapes.items()} for _ in range(args.layers)]
    lm_head = packed(VOCAB, HIDDEN)
    bytes_per_rank = sum(s[0] * (s[1] // 64) * 34 for s in shapes.values()) * args.layers + VOCAB * (HIDDEN // 64) * 34

    x_h = (torch.randn(1, HIDDEN, device=dev, generator=gen) * 0.1).half()
    x_os = x_h[:, :HIDDEN // tp].contiguous()          # this rank's slice of the attention output (stand-in)
    qkv_out = torch.empty(1, shapes["qkv"][0], device=dev, dtype=torch.float16)
    mlp_out = torch.empty(1, INTER // tp, device=dev, dtype=torch.float16)
    hid = torch.empty(1, HIDDEN, device=dev, dtype=torch.float16)
    logits = torch.empty(1, VOCAB, device=dev, dtype=torch.float16)

    peer = None
    if world > 1 and not args.no_allreduce and args.allreduce in ("peer", "fused"):
        from ipex_llm_b200.transformers.tensor_parallel import PeerAllReduce
        peer = PeerAllReduce()

    def allreduce(t):
        if world > 1 and not args.no_allreduce:
            if peer is not None:
                peer.all_reduce(t)
            else:
                dist.all_reduce(t)

    fused = peer is not None and args.allreduce == "fused"

    def step():
        for L in layers:
            lbl.lowbit_matmul(x_h, L["qkv"], shapes["qkv"], qt, out=qkv_out)
            if fused:
                peer.matmul_all_reduce(x_os, L["o"], shapes["o"], qt, out=hid)
            else:
                lbl.lowbit_matmul(x_os, L["o"], shapes["o"], qt, out=hid)
                allreduce(hid)
            lbl.mlp_forward(x_h, L["gate"], L["up"], 1, HIDDEN, INTER // tp, lbl.SILU, qt, out=mlp_out)
            if fused:
                peer.matmul_all_reduce(mlp_out, L["down"], shapes["down"], qt, out=hid)
            else:
                lbl.lowbit_matmul(mlp_out, L["down"], shapes["down"], qt, out=hid)
                allreduce(hid)
        lbl.lowbit_matmul(x_h, lm_head, (VOCAB, HIDDEN), qt, out=logits)

    def log(msg):
        if rank == 0:
            print(msg, file=sys.stderr, flush=True)

    stream = torch.cuda.Stream()
    graph = None
    with torch.cuda.stream(stream):
        for _ in range(2):
            step()
        torch.cuda.synchronize()
        log("eager steps done")
        # torch's ProcessGroupNCCL collectives hung under CUDA-graph capture on the gpurun box (round 1, unresolved):
        # with collectives the step is launched eagerly (launch-bound: the number is a lower bound)
        if not args.no_graph and (world == 1 or args.no_allreduce or peer is not None):
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=stream, capture_error_mode="thread_local"):
                step()
            log("captured")
    torch.cuda.synchronize()

    def replay():
        with torch.cuda.stream(stream):                # the events below are recorded on this stream
            if graph is not None:
                graph.replay()
            else:
                step()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(3, args.warmup)):
        replay()
    barrier()
    log("warm")
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        replay()
    e1.record(stream)
    barrier()
    ms = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t[0])
    if rank == 0:
        ms_step = ms / args.steps
        print(json.dumps({
            "metric": "sym_int4 decode tok/s (Llama-2-70B LowBitLinear stream, bs1, tensor parallel)",
            "value": 1e3 / ms_step, "unit": "tok/s", "n_gpus": world, "steps": args.steps, "ms_per_step": ms_step,
            "scaling": "strong", "layers": args.layers, "allreduce": (args.allreduce if world > 1 and not args.no_allreduce else None), "cuda_graph": graph is not None,
            "collectives_per_step": 0 if (world == 1 or args.no_allreduce) else 2 * args.layers,
            "packed_bytes_per_rank": bytes_per_rank,
            "per_rank_GBps": bytes_per_rank / (ms_step * 1e-3) / 1e9,
            "config": {"workload": "Llama-2-70B sym_int4 bs1 decode, LowBitLinear stream, column/row sharded, all-reduce "
                                   "of [1, 8192] fp16 after o_proj and down_proj", "parallelism": f"tp{world}"}}))
    if peer is not None:
        assert not peer.timed_out()
        peer.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
