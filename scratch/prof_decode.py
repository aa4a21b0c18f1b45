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
ap.add_argument("--only-full", action="store_true", dest="only_full")
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
if args.only_full:
    sys.exit(0)
noop = lambda *a, **k: None
names = ("lowbit_rms_norm", "lowbit_decode_attention")
saved = {n: getattr(ggml, n) for n in names}
for group in (("lowbit_rms_norm",), ("lowbit_decode_attention",), names):
    for n in group:
        setattr(ggml, n, noop)
    run("without " + " + ".join(g.replace("lowbit_", "") for g in group))
    for n in group:
        setattr(ggml, n, saved[n])
os.environ["IPEX_LLM_B200_DECODE_PDL"] = "0"
run("full step, no PDL")
