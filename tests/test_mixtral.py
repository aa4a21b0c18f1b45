"""Mixtral through the public API (BASELINE.json configs[4], at toy size): transformers >= 5 keeps the experts of a layer
as two fused 3-D parameters; ``optimize_model`` / ``AutoModelForCausalLM.from_model`` turn them into per-expert
``LowBitLinear`` modules, the MoE block runs the one-token fused kernel (``moe_forward_vec``) at decode and per-expert
matmuls at prefill, and ``generate`` replays the recorded decode step.  Checked against the fp32 Hugging Face model whose
expert tensors and linears were replaced by their oracle-dequantised values."""
import os

import numpy as np
import pytest
import torch

import oracle

pytestmark = pytest.mark.gpu
transformers = pytest.importorskip("transformers")


def tiny_mixtral(seed=0):
    cfg = transformers.MixtralConfig(hidden_size=256, intermediate_size=512, num_hidden_layers=2, num_attention_heads=4,
                                     num_key_value_heads=2, vocab_size=320, max_position_embeddings=256, num_local_experts=4,
                                     num_experts_per_tok=2, sliding_window=None, head_dim=64, tie_word_embeddings=False)
    torch.manual_seed(seed)
    return transformers.MixtralForCausalLM(cfg).eval()


def through_oracle(w, qtype):
    a = w.detach().float().numpy()
    return torch.from_numpy(oracle.dequantize(oracle.quantize(np.ascontiguousarray(a), qtype), a.shape, qtype))


def oracle_dequantised_copy(model, qtype):
    ref = tiny_mixtral()
    ref.load_state_dict(model.state_dict())
    for mod in ref.modules():
        if isinstance(mod, torch.nn.Linear) and mod.in_features % 64 == 0:
            mod.weight.data = through_oracle(mod.weight.data, qtype)
        elif hasattr(mod, "gate_up_proj") and isinstance(mod.gate_up_proj, torch.nn.Parameter):
            for e in range(mod.gate_up_proj.shape[0]):
                mod.gate_up_proj.data[e] = through_oracle(mod.gate_up_proj.data[e], qtype)       # rows are independent
                mod.down_proj.data[e] = through_oracle(mod.down_proj.data[e], qtype)
    return ref.float().eval()


@pytest.mark.parametrize("qtype", ["sym_int4", "fp8_e4m3"])
def test_mixtral_experts_are_quantised_and_match_the_oracle_model(native_lib, cuda_device, qtype, tmp_path):
    from ipex_llm.transformers import AutoModelForCausalLM
    from ipex_llm_b200.transformers.models.mixtral import LowBitExperts
    from ipex_llm_b200.transformers.low_bit_linear import LowBitLinear
    base = tiny_mixtral()
    ref = oracle_dequantised_copy(base, qtype) if qtype == "sym_int4" else None
    model = AutoModelForCausalLM.from_model(base, low_bit=qtype)
    experts = model.model.layers[0].mlp.experts
    assert isinstance(experts, LowBitExperts) and len(experts.gate_proj) == 4
    assert sum(isinstance(m, LowBitLinear) for m in model.modules()) == 2 * (2 + 3 * 4) + 1      # qkv, o, 4 x (gate, up, down), lm_head
    if qtype == "sym_int4":
        w = tiny_mixtral().model.layers[0].mlp.experts
        got = experts.up_proj[2].weight.data.view(torch.uint8).numpy()
        np.testing.assert_array_equal(got, oracle.quantize(w.gate_up_proj.data[2, 512:].numpy(), qtype))
        got = experts.down_proj[3].weight.data.view(torch.uint8).numpy()
        np.testing.assert_array_equal(got, oracle.quantize(np.ascontiguousarray(w.down_proj.data[3].numpy()), qtype))
    model = model.to(cuda_device)
    ids = torch.randint(0, 320, (2, 9), generator=torch.Generator().manual_seed(1))
    with torch.no_grad():
        got = model(ids.to(cuda_device)).logits.float().cpu()
        one = model(ids[:1, :1].to(cuda_device)).logits.float().cpu()            # one token: the fused expert kernel
    if ref is not None:
        with torch.no_grad():
            want = ref(ids).logits
            want_one = ref(ids[:1, :1]).logits
        assert float((got - want).abs().mean() / want.abs().mean()) < 2e-3
        assert float((one - want_one).abs().mean() / want_one.abs().mean()) < 2e-3
    # save_low_bit / load_low_bit keep the per-expert packed weights
    d = str(tmp_path / "mixtral_lowbit")
    model.save_low_bit(d)
    again = AutoModelForCausalLM.load_low_bit(d).to(cuda_device)
    assert isinstance(again.model.layers[1].mlp.experts, LowBitExperts)
    with torch.no_grad():
        torch.testing.assert_close(again(ids.to(cuda_device)).logits.float().cpu(), got, rtol=0, atol=0)


def test_mixtral_recorded_decode_step(native_lib, cuda_device):
    from ipex_llm.transformers import AutoModelForCausalLM
    from test_model_e2e import assert_same_tokens_up_to_a_tie
    model = AutoModelForCausalLM.from_model(tiny_mixtral(), low_bit="sym_int4").to(cuda_device).half()
    assert model.generate.__func__.__name__ == "generate_greedy_graph"
    ids = torch.randint(0, 320, (1, 30), generator=torch.Generator().manual_seed(4)).to(cuda_device)
    with torch.no_grad():
        fast = model.generate(ids, max_new_tokens=20, do_sample=False, pad_token_id=0)
        assert getattr(model, "last_decode_launches_per_token", 0) > 0
        slow = model._hf_generate(ids, max_new_tokens=20, do_sample=False, pad_token_id=0)
    assert fast.shape == (1, 50)
    assert_same_tokens_up_to_a_tie(model, fast, slow, torch.float16)
