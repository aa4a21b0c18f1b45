"""End-to-end on a tiny random-init Llama: the loader surface of the reference
(``optimize_model`` / ``AutoModelForCausalLM.from_model`` -> ``.to('cuda')`` -> forward / greedy
``generate`` -> ``save_low_bit`` / ``load_low_bit``), checked against a plain fp32 Hugging Face model
whose linear weights were dequantised by the CPU oracle (the style of the reference's own tests:
test_transformers_api.py:121-169 -- round trip and logits difference)."""
import os

import numpy as np
import pytest
import torch

import oracle

pytestmark = pytest.mark.gpu
transformers = pytest.importorskip("transformers")


def tiny_llama(seed=0):
    cfg = transformers.LlamaConfig(hidden_size=256, intermediate_size=512, num_hidden_layers=2, num_attention_heads=4,
                                   num_key_value_heads=2, vocab_size=320, max_position_embeddings=128,
                                   tie_word_embeddings=False)
    torch.manual_seed(seed)
    m = transformers.LlamaForCausalLM(cfg).eval()
    return m


def oracle_dequantised_copy(model, qtype):
    """fp32 HF model whose nn.Linear weights went through oracle quantise -> dequantise"""
    ref = tiny_llama()
    ref.load_state_dict(model.state_dict())
    for mod in ref.modules():
        if isinstance(mod, torch.nn.Linear) and mod.in_features % 64 == 0:
            w = mod.weight.data.float().numpy()
            mod.weight.data = torch.from_numpy(oracle.dequantize(oracle.quantize(w, qtype), w.shape, qtype))
    return ref.float().eval()


@pytest.mark.parametrize("qtype", ["sym_int4", "asym_int4", "nf4", "fp8_e5m2"])
def test_optimize_model_forward_matches_oracle_model(native_lib, cuda_device, qtype, tmp_path):
    from ipex_llm_b200 import optimize_model, load_low_bit, low_memory_init
    from ipex_llm_b200.transformers.low_bit_linear import LowBitLinear
    base = tiny_llama()
    ref = oracle_dequantised_copy(base, qtype)
    model = optimize_model(base, low_bit=qtype)
    lins = [m for m in model.modules() if isinstance(m, LowBitLinear)]
    assert len(lins) == 2 * 7 + 1 and all(l.weight.quantized for l in lins)           # q,k,v,o,gate,up,down x2 + lm_head
    # packed bytes on the CPU are the canonical layout the oracle produces
    w0 = tiny_llama().model.layers[0].self_attn.q_proj.weight.data.numpy()
    got = model.model.layers[0].self_attn.q_proj.weight.data.view(torch.uint8).numpy()
    np.testing.assert_array_equal(got, oracle.quantize(w0, qtype))

    model = model.to(cuda_device)
    ids = torch.randint(0, 320, (2, 12), generator=torch.Generator().manual_seed(1))
    with torch.no_grad():
        # lm_head was converted with optimize_lm_head: prompts longer than 10 tokens only produce the
        # last position's logits (reference low_bit_linear.py:282-289)
        want = ref(ids).logits[:, -1:, :]
        got = model(ids.to(cuda_device)).logits.float().cpu()
    assert got.shape == want.shape == (2, 1, 320)
    rel = (got - want).abs().mean() / want.abs().mean()
    assert rel < 2e-3, float(rel)                      # reference's own bar is 0.05 (test_transformers_api.py:151-169)

    # greedy decode: token-identical with the oracle-dequantised fp32 model
    with torch.no_grad():
        out_ref = ref.generate(ids[:1], max_new_tokens=16, do_sample=False)
        out_gpu = model.generate(ids[:1].to(cuda_device), max_new_tokens=16, do_sample=False).cpu()
    assert torch.equal(out_ref, out_gpu), (out_ref, out_gpu)

    # save_low_bit / load_low_bit: same bytes, same logits
    save_dir = str(tmp_path / "lowbit")
    model.save_low_bit(save_dir)
    assert os.path.exists(os.path.join(save_dir, "pytorch_model.bin")) and os.path.exists(os.path.join(save_dir, "bigdl_config.json"))
    with low_memory_init():
        shell = tiny_llama()
    loaded = load_low_bit(shell, save_dir)
    np.testing.assert_array_equal(loaded.model.layers[0].self_attn.q_proj.weight.data.view(torch.uint8).numpy(),
                                  oracle.quantize(w0, qtype))
    loaded = loaded.to(cuda_device)
    with torch.no_grad():
        again = loaded(ids.to(cuda_device)).logits.float().cpu()
    torch.testing.assert_close(again, got, rtol=0, atol=0)


def test_auto_model_surface_and_round_trip(native_lib, cuda_device, tmp_path):
    from ipex_llm.transformers import AutoModelForCausalLM         # the reference's import path (alias package)
    from ipex_llm.transformers.low_bit_linear import LowBitLinear
    model = AutoModelForCausalLM.from_model(tiny_llama(), low_bit="sym_int4", modules_to_not_convert=["lm_head"])
    assert model.config.to_dict()["bigdl_transformers_low_bit"] == "sym_int4"
    assert not isinstance(model.lm_head, LowBitLinear) and isinstance(model.model.layers[0].mlp.down_proj, LowBitLinear)
    model = model.to(cuda_device).half()               # dtype cast must leave packed weights alone
    ids = torch.randint(0, 320, (1, 9), generator=torch.Generator().manual_seed(2)).to(cuda_device)
    with torch.no_grad():
        a = model(ids).logits.float().cpu()
    d = str(tmp_path / "hf_lowbit")
    model.save_low_bit(d)
    for f in ("pytorch_model.bin", "config.json", "load_keys.json"):
        assert os.path.exists(os.path.join(d, f)), f
    again = AutoModelForCausalLM.load_low_bit(d, modules_to_not_convert=["lm_head"]).to(cuda_device).half()
    with torch.no_grad():
        b = again(ids).logits.float().cpu()
    torch.testing.assert_close(a, b, rtol=0, atol=0)
    # prefill with a long prompt exercises the tensor-core GEMM (M > 32) inside the same module
    long_ids = torch.randint(0, 320, (1, 100), generator=torch.Generator().manual_seed(3)).to(cuda_device)
    with torch.no_grad():
        lg = model(long_ids).logits
    assert torch.isfinite(lg).all()


def test_optimize_model_low_bit_embedding(native_lib, cuda_device):
    """embedding_qtype: the token embedding becomes a LowBitEmbedding (reference convert.py:730-733); logits match
    the model whose embedding table and linears were replaced by their oracle-dequantised values."""
    from ipex_llm_b200 import optimize_model
    from ipex_llm_b200.transformers.embedding import LowBitEmbedding
    base = tiny_llama()
    ref = oracle_dequantised_copy(base, "sym_int4")
    table = base.model.embed_tokens.weight.data.float().numpy()
    ref.model.embed_tokens.weight.data = torch.from_numpy(
        oracle.dequantize(oracle.quantize(table, "fp8_e4m3"), table.shape, "fp8_e4m3"))
    model = optimize_model(base, low_bit="sym_int4", embedding_qtype="fp8_e4m3").to(cuda_device)
    assert isinstance(model.model.embed_tokens, LowBitEmbedding)
    ids = torch.randint(0, 320, (1, 8), generator=torch.Generator().manual_seed(3))
    with torch.no_grad():
        want = ref(ids).logits
        got = model(ids.to(cuda_device)).logits.float().cpu()
    rel = (got - want).abs().mean() / want.abs().mean()
    assert rel < 2e-3, float(rel)
