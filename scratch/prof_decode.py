"""development: where does the recorded decode step's time go?  Ablation (pieces replaced by no-ops) + optional
cudaProfilerStart/Stop bracket for an ncu launch list:
    python scratch/prof_decode.py [--layers 32] [--ncu]"""
import argparse, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench
import ipex_llm_b200.ggml.model.llama.llama_cpp as ggml
from ipex_llm_b200.transformers import decode_graph as dg

ap = argparse.ArgumentParser()
ap.add_argument("--layers", type=int, default=32)
ap.add_argument("--ncu", action="store_true")
ap.add_argument("--prompt", type=int, default=512)
args = ap.parse_args()
dev = torch.device("cuda", 0)
model = bench.build_llama(dev, args.layers)
ids = torch.randint(0, 32000, (1, args.prompt), device=dev)


def run(tag, n=64):
    model.__dict__.pop("_b200_decode_graphs", None)
    with torch.no_grad():
        model.generate(ids, max_new_tokens=8, min_new_tokens=8, do_sample=False, pad_token_id=0)
        model.generate(ids, max_new_tokens=n + 1, min_new_tokens=n + 1, do_sample=False, pad_token_id=0)
        model.generate(ids, max_new_tokens=n + 1, min_new_tokens=n + 1, do_sample=False, pad_token_id=0)
    print(f"{tag:28s} {model.rest_cost_mean_device * 1e3:8.4f} ms/token  ({model.rest_cost_mean_device * 1e6 / args.layers:6.2f} us/layer)", flush=True)


if args.ncu:
    with torch.no_grad():
        model.generate(ids, max_new_tokens=4, min_new_tokens=4, do_sample=False, pad_token_id=0)
        torch.cuda.synchronize()
        torch.cuda.cudart().cudaProfilerStart()
        model.generate(ids, max_new_tokens=3, min_new_tokens=3, do_sample=False, pad_token_id=0)
        torch.cuda.synchronize()
        torch.cuda.cudart().cudaProfilerStop()
    sys.exit(0)

run("full step")
noop = lambda *a, **k: None
saved = {n: getattr(ggml, n) for n in ("lowbit_rms_norm", "lowbit_rope_kv_append", "lowbit_sdp_fp8_devlen", "lowbit_greedy_advance")}
for name in saved:
    setattr(ggml, name, noop if name != "lowbit_greedy_advance" else saved[name])
    if name == "lowbit_greedy_advance":
        continue
    run("without " + name)
    setattr(ggml, name, saved[name])
for name in ("lowbit_rms_norm", "lowbit_rope_kv_append", "lowbit_sdp_fp8_devlen"):
    setattr(ggml, name, noop)
run("linears + embed + greedy only")
for name in saved:
    setattr(ggml, name, saved[name])
os.environ["IPEX_LLM_B200_DECODE_PDL"] = "0"
run("full step, no PDL")
