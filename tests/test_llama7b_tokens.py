"""Token identity on the named configuration (BASELINE.json configs[0]): Llama-2-7B architecture, random-init, sym_int4,
batch 1, greedy, 512-token prompt, 128 new tokens -- through the public API
``ipex_llm.transformers.AutoModelForCausalLM.from_model(...).generate()`` on the GPU, against the tokens that the
oracle-dequantised fp32 model produced on the CPU (tests/golden/make_llama7b_tokens.py, stored with the margin of every
greedy choice).  The GPU model computes in fp32 like the reference's CPU LowBitLinear path does
(low_bit_linear.py:498-499, 745): the GEMV keeps fp32 activations to 2^-24, so the two runs differ by summation order
only, far below the stored margins.

The half-precision model (FP8 KV cache, recorded decode step -- what bench.py times) cannot be token-identical to an
fp32 run over 128 steps; it is checked on the same prompt for agreement of its first tokens and of its logits."""
import json
import os
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
GOLDEN = os.path.join(HERE, "golden", "llama7b_sym_int4_tokens.json")


@pytest.mark.timeout(1500)
def test_llama2_7b_512_in_128_out_tokens(native_lib, cuda_device):
    import llama_init
    from ipex_llm.transformers import AutoModelForCausalLM
    gold = json.load(open(GOLDEN))
    assert gold["prompt_len"] == llama_init.PROMPT_LEN and len(gold["new_tokens"]) == llama_init.NEW_TOKENS
    model = AutoModelForCausalLM.from_model(llama_init.build(), low_bit="sym_int4")       # quantised by the CUDA kernel
    for m in model.modules():
        if hasattr(m, "conver_to_half"):
            m.conver_to_half = False            # fp32 in, fp32 through the matmul: the CPU reference's contract
    model = model.to(cuda_device)
    ids = llama_init.prompt().to(cuda_device)
    with torch.no_grad():
        out = model.generate(ids, max_new_tokens=llama_init.NEW_TOKENS, min_new_tokens=llama_init.NEW_TOKENS, do_sample=False,
                             pad_token_id=0)
    got = out[0, ids.shape[1]:].tolist()
    want = gold["new_tokens"]
    first_bad = next((i for i, (a, b) in enumerate(zip(got, want)) if a != b), None)
    assert first_bad is None, (f"token {first_bad} differs: got {got[first_bad]}, want {want[first_bad]}, "
                               f"stored margin {gold['margins'][first_bad]:.3g}")

    # the configuration bench.py times: fp16 activations, FP8 KV cache, recorded decode step
    model = model.half()
    with torch.no_grad():
        fast = model.generate(ids, max_new_tokens=llama_init.NEW_TOKENS, min_new_tokens=llama_init.NEW_TOKENS, do_sample=False)
    assert getattr(model, "last_decode_launches_per_token", 0) > 0
    got16 = fast[0, ids.shape[1]:].tolist()
    # fp16 rounding noise reaches the smallest margins after a while; the opening tokens must agree
    agree = next((i for i, (a, b) in enumerate(zip(got16, want)) if a != b), len(want))
    assert agree >= 4, (agree, got16[:8], want[:8])
