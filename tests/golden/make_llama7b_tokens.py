"""Golden greedy tokens of BASELINE.json configs[0]: Llama-2-7B architecture, random-init (tests/golden/llama_init.py),
sym_int4, batch 1, 512-token prompt, 128 new tokens.

CPU side of the check, run once in the build container:

    python tests/golden/make_llama7b_tokens.py            # ~10 minutes, ~30 GB of RAM

Every linear weight goes through the CPU oracle's quantise -> dequantise (oracle/lowbit_oracle.c: the bytes
ggml_quantize_tensor produces and the values ggml_dequantize gives back), the fp32 model is run by transformers'
stock forwards and generate loop -- i.e. the reference's "mode B" semantics, x . dequant(W)^T in fp32
(/root/reference/python/llm/src/ipex_llm/transformers/low_bit_linear.py:581-583, 741-742) -- and the 128 tokens are
stored in tests/golden/llama7b_sym_int4_tokens.json together with the margin of every greedy choice (top-1 minus
top-2 logit), so that a later mismatch can be told apart from a near tie.  tests/test_llama7b_tokens.py reproduces the
tokens on the GPU through ipex_llm.transformers.AutoModelForCausalLM.from_model(...).generate().
"""
import json
import os
import sys
import time

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, HERE)

import llama_init  # noqa: E402
import oracle      # noqa: E402

QTYPE = "sym_int4"


def through_oracle(name, w):
    a = w.numpy()
    return torch.from_numpy(oracle.dequantize(oracle.quantize(a, QTYPE), a.shape, QTYPE))


def main():
    layers = int(os.environ.get("GOLDEN_LAYERS", llama_init.LLAMA2_7B["num_hidden_layers"]))
    out_path = os.path.join(HERE, "llama7b_sym_int4_tokens.json" if layers == 32 else f"llama7b_sym_int4_tokens_{layers}l.json")
    t0 = time.time()
    model = llama_init.build({"num_hidden_layers": layers}, transform=through_oracle)
    print(f"model built in {time.time() - t0:.0f} s", flush=True)
    ids = llama_init.prompt()
    t0 = time.time()
    with torch.no_grad():
        out = model.generate(ids, max_new_tokens=llama_init.NEW_TOKENS, min_new_tokens=llama_init.NEW_TOKENS, do_sample=False,
                             output_scores=True, return_dict_in_generate=True, pad_token_id=0)
    print(f"generated in {time.time() - t0:.0f} s", flush=True)
    tokens = out.sequences[0, ids.shape[1]:].tolist()
    margins = []
    for s in out.scores:
        top = torch.topk(s[0].float(), 2).values
        margins.append(float(top[0] - top[1]))
    json.dump({"config": dict(llama_init.LLAMA2_7B, num_hidden_layers=layers), "qtype": QTYPE, "prompt_seed": 1,
               "prompt_len": llama_init.PROMPT_LEN, "new_tokens": tokens, "margins": margins,
               "how": "oracle-dequantised fp32 model, transformers stock forwards, greedy, CPU"}, open(out_path, "w"))
    print(out_path, "min margin", min(margins), "tokens", tokens[:16])


if __name__ == "__main__":
    main()
