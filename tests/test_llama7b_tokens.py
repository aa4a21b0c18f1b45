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
def test_llama2_7b_512_in_128_out_tokens(native_lib, cuda_device, monkeypatch):
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

    # Half precision.  This random-init 32-layer model amplifies any perturbation (a 2^-11 rounding of the activations moves
    # its final logits by ~15 %, the 2-bit mantissa of an e5m2 KV cache by ~50 %), so a half-precision run cannot reproduce
    # the fp32 tokens and its logits are compared where the perturbation is still small: after the first decoder layer
    # (tensor-core GEMM at M = 512, rotary, attention, fused-free MLP).  The configuration bench.py times adds the e5m2 KV
    # cache and the recorded decode step: it must agree with the same model's eager forwards.
    first = {}
    hook = model.model.layers[0].register_forward_hook(lambda m, a, out: first.__setitem__("h", (out[0] if isinstance(out, tuple) else out).float()))
    with torch.no_grad():
        logits32 = model(ids).logits[0, -1].float()
    h32 = first["h"]
    model = model.half()
    short = 32
    monkeypatch.setenv("IPEX_LLM_QUANTIZE_KV_CACHE", "0")
    with torch.no_grad():
        logits16_fp_cache = model(ids).logits[0, -1].float()
    monkeypatch.delenv("IPEX_LLM_QUANTIZE_KV_CACHE")
    hook.remove()
    rel_h = float((first["h"] - h32).norm() / h32.norm())
    assert rel_h < 3e-2, rel_h
    rel_fp = float((logits16_fp_cache - logits32).norm() / logits32.norm())
    with torch.no_grad():
        logits16 = model(ids).logits[0, -1].float()
        fast = model.generate(ids, max_new_tokens=short, min_new_tokens=short, do_sample=False, pad_token_id=0)
        assert getattr(model, "last_decode_launches_per_token", 0) > 0              # the recorded step ran
        slow = model._hf_generate(ids, max_new_tokens=short, min_new_tokens=short, do_sample=False, pad_token_id=0)
    rel = float((logits16 - logits32).norm() / logits32.norm())
    print(f"vs fp32: first layer output {rel_h:.2e}; prefill logits fp16 + fp cache {rel_fp:.4f}, fp16 + e5m2 cache {rel:.4f} (relative L2)")
    assert rel < 1.0, rel
    if not torch.equal(fast, slow):
        i = int((fast[0] != slow[0]).nonzero()[0, 0])
        with torch.no_grad():
            lg = model(fast[:, :i]).logits[0, -1].float()
        gap = abs(float(lg[fast[0, i]] - lg[slow[0, i]]))
        assert gap <= 4 * 2.0 ** -10 * float(lg.abs().max()), (i, gap)
