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
    assert len(lins) == 2 * 5 + 1 and all(l.weight.quantized for l in lins)           # qkv,o,gate,up,down x2 + lm_head
    # packed bytes on the CPU are the canonical layout the oracle produces; q, k, v were merged into qkv_proj before
    # quantisation (reference convert.py:1000-1003), which is the row concatenation of their packed bytes
    a0 = tiny_llama().model.layers[0].self_attn
    w0 = torch.cat([a0.q_proj.weight.data, a0.k_proj.weight.data, a0.v_proj.weight.data]).numpy()
    got = model.model.layers[0].self_attn.qkv_proj.weight.data.view(torch.uint8).numpy()
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
    np.testing.assert_array_equal(loaded.model.layers[0].self_attn.qkv_proj.weight.data.view(torch.uint8).numpy(),
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


def test_optimize_model_true_vs_false(native_lib, cuda_device, tmp_path):
    """the reference's own check (test/inference/test_transformers_api.py:151-169): logits of the optimised forwards
    against the stock forwards on the same low-bit weights; and a model saved without the optimisations loads back
    with its own key layout"""
    from ipex_llm.transformers import AutoModelForCausalLM
    from ipex_llm.transformers.low_bit_linear import LowBitLinear
    plain = AutoModelForCausalLM.from_model(tiny_llama(), low_bit="sym_int4", optimize_model=False)
    fast = AutoModelForCausalLM.from_model(tiny_llama(), low_bit="sym_int4", optimize_model=True)
    assert hasattr(plain.model.layers[0].self_attn, "q_proj") and hasattr(fast.model.layers[0].self_attn, "qkv_proj")
    assert sum(isinstance(m, LowBitLinear) for m in plain.modules()) == 2 * 7 + 1
    assert plain.generate.__func__.__name__ == "generate"
    plain, fast = plain.to(cuda_device), fast.to(cuda_device)
    ids = torch.randint(0, 320, (2, 9), generator=torch.Generator().manual_seed(5)).to(cuda_device)
    with torch.no_grad():
        a, b = plain(ids).logits.float(), fast(ids).logits.float()
    assert float((a - b).abs().mean() / a.abs().mean()) < 1e-4            # the reference's bar is 0.05
    with torch.no_grad():
        ta = plain.generate(ids[:1], max_new_tokens=12, do_sample=False)
        tb = fast.generate(ids[:1], max_new_tokens=12, do_sample=False)
    assert torch.equal(ta, tb)
    d = str(tmp_path / "plain")
    plain.save_low_bit(d)
    again = AutoModelForCausalLM.load_low_bit(d).to(cuda_device)
    assert hasattr(again.model.layers[0].self_attn, "q_proj")
    with torch.no_grad():
        torch.testing.assert_close(again(ids).logits.float(), a, rtol=0, atol=0)


def tiny_mistral(seed=0):
    cfg = transformers.MistralConfig(hidden_size=256, intermediate_size=512, num_hidden_layers=2, num_attention_heads=4,
                                     num_key_value_heads=2, vocab_size=320, max_position_embeddings=256, sliding_window=None,
                                     head_dim=64, tie_word_embeddings=False)
    torch.manual_seed(seed)
    return transformers.MistralForCausalLM(cfg).eval()


def assert_same_tokens_up_to_a_tie(model, a, b, dtype):
    """The recorded step and the eager forwards run the same kernels, except that the decode attention splits the keys
    for the cache CAPACITY in one case and for the current LENGTH in the other: partial sums meet in a different order,
    an output may move by one unit in the last place, and a greedy choice between two nearly equal logits may flip.
    Sequences must therefore agree up to the first position where the model itself is undecided: there the two
    candidates' logits (eager forward on the common prefix) differ by no more than a few units in the last place."""
    if torch.equal(a, b):
        return
    rows = (a != b).any(dim=1).nonzero()[:, 0].tolist()
    for r in rows:
        i = int((a[r] != b[r]).nonzero()[0, 0])
        with torch.no_grad():
            logits = model(a[r:r + 1, :i]).logits[0, -1].float()
        eps = 2.0 ** -10 if dtype == torch.float16 else 2.0 ** -7
        gap = abs(float(logits[a[r, i]] - logits[b[r, i]]))
        assert gap <= 4 * eps * float(logits.abs().max()), (r, i, gap, a[r, i - 2:i + 3].tolist(), b[r, i - 2:i + 3].tolist())
        top2 = torch.topk(logits, 2).indices.tolist()
        assert int(a[r, i]) in top2 and int(b[r, i]) in top2


@pytest.mark.parametrize("family,qtype,dtype", [("llama", "sym_int4", torch.float16), ("mistral", "fp8_e4m3", torch.float16),
                                                ("mistral", "fp6", torch.bfloat16), ("llama", "asym_int4", torch.bfloat16)])
def test_recorded_decode_step_matches_the_eager_forwards(native_lib, cuda_device, family, qtype, dtype):
    """half-precision model, FP8 KV cache: ``generate`` replays the recorded decode step (CUDA graph); the same
    request through transformers' generate loop runs the patched forwards one kernel at a time.  Same kernels, same
    roundings -> same tokens; and the logits stay close to the oracle-dequantised fp32 model's."""
    from ipex_llm.transformers import AutoModelForCausalLM
    from ipex_llm_b200.transformers import kv
    base = tiny_llama() if family == "llama" else tiny_mistral()
    model = AutoModelForCausalLM.from_model(base, low_bit=qtype).to(cuda_device).to(dtype)
    assert model.generate.__func__.__name__ == "generate_greedy_graph"
    ids = torch.randint(0, 320, (1, 40), generator=torch.Generator().manual_seed(7)).to(cuda_device)
    with torch.no_grad():
        fast = model.generate(ids, max_new_tokens=24, do_sample=False)
        assert getattr(model, "last_decode_launches_per_token", 0) > 0              # the recorded path ran
        slow = model._hf_generate(ids, max_new_tokens=24, do_sample=False)
        out = model(ids, use_cache=True)
    assert isinstance(out.past_key_values, kv.DynamicFp8Cache) and out.past_key_values.key_cache[0].dtype == torch.uint8
    assert fast.shape == (1, 64) and torch.equal(fast[:, :40], ids)
    assert_same_tokens_up_to_a_tie(model, fast, slow, dtype)
    # a second call reuses the recorded step and its cache arena and gives the same answer; batch 2 records another
    with torch.no_grad():
        again = model.generate(ids, max_new_tokens=24, do_sample=False)
        ids2 = torch.cat([ids, ids.flip(1)], dim=0)
        fast2 = model.generate(ids2, max_new_tokens=8, do_sample=False)
        slow2 = model._hf_generate(ids2, max_new_tokens=8, do_sample=False)
    assert torch.equal(again, fast)
    assert_same_tokens_up_to_a_tie(model, fast2, slow2, dtype)
    if family == "llama" and qtype == "sym_int4":                   # a batch beyond the fused gate/up kernel's 8 rows
        ids12 = torch.randint(0, 320, (12, 20), generator=torch.Generator().manual_seed(11)).to(cuda_device)
        with torch.no_grad():
            fast12 = model.generate(ids12, max_new_tokens=6, do_sample=False)
            slow12 = model._hf_generate(ids12, max_new_tokens=6, do_sample=False)
        assert_same_tokens_up_to_a_tie(model, fast12, slow12, dtype)
        # 24 sequences: the linears whose shapes allow it take the skinny tensor-core kernel, in the recorded step and in the eager forward
        ids24 = torch.randint(0, 320, (24, 20), generator=torch.Generator().manual_seed(12)).to(cuda_device)
        with torch.no_grad():
            fast24 = model.generate(ids24, max_new_tokens=6, do_sample=False)
            slow24 = model._hf_generate(ids24, max_new_tokens=6, do_sample=False)
        assert_same_tokens_up_to_a_tie(model, fast24, slow24, dtype)
    assert_same_tokens_up_to_a_tie(model, fast2[:1], fast[:, :48], dtype)
    assert model.rest_cost_mean > 0 and model.first_cost > 0


def test_generate_stops_at_eos_and_falls_back(native_lib, cuda_device):
    from ipex_llm.transformers import AutoModelForCausalLM
    model = AutoModelForCausalLM.from_model(tiny_llama(), low_bit="sym_int4").to(cuda_device).half()
    ids = torch.randint(0, 320, (1, 20), generator=torch.Generator().manual_seed(9)).to(cuda_device)
    with torch.no_grad():
        free = model.generate(ids, max_new_tokens=16, do_sample=False)
        eos = int(free[0, 20 + 5])                                  # the 6th new token becomes the end-of-sequence id
        first = [int(t) for t in free[0, 20:]].index(eos)
        stopped = model.generate(ids, max_new_tokens=16, do_sample=False, eos_token_id=eos, pad_token_id=0)
        want = model._hf_generate(ids, max_new_tokens=16, do_sample=False, eos_token_id=eos, pad_token_id=0)
    assert stopped.shape[1] == 20 + first + 1 and int(stopped[0, -1]) == eos
    assert torch.equal(stopped, want)
    # requests the recorded step does not serve go to transformers' generate (sampling here) and still work
    torch.manual_seed(0)
    with torch.no_grad():
        sampled = model.generate(ids, max_new_tokens=8, do_sample=True, top_k=5)
    assert sampled.shape == (1, 28)
    # fp32 model: floating-point cache, eager path
    model32 = AutoModelForCausalLM.from_model(tiny_llama(), low_bit="sym_int4").to(cuda_device)
    with torch.no_grad():
        t32 = model32.generate(ids, max_new_tokens=8, do_sample=False)
    assert t32.shape == (1, 28)


def test_from_pretrained_load_in_4bit_and_low_bit(native_lib, cuda_device, tmp_path):
    """the reference's headline call: AutoModelForCausalLM.from_pretrained(path, load_in_4bit=True) (model.py:111-388;
    load_in_4bit == load_in_low_bit="sym_int4", model.py:327), from a checkpoint directory written by transformers"""
    from ipex_llm.transformers import AutoModelForCausalLM
    from ipex_llm.transformers.low_bit_linear import LowBitLinear
    from ipex_llm.ggml.quantize import ggml_tensor_qtype
    d = str(tmp_path / "hf")
    tiny_llama().save_pretrained(d)
    model = AutoModelForCausalLM.from_pretrained(d, load_in_4bit=True)
    assert model.config.to_dict()["bigdl_transformers_low_bit"] == "sym_int4"
    qkv = model.model.layers[0].self_attn.qkv_proj
    assert isinstance(qkv, LowBitLinear) and qkv.qtype == ggml_tensor_qtype["sym_int4"] and qkv.weight.data.device.type == "cpu"
    a0 = tiny_llama().model.layers[0].self_attn
    w0 = torch.cat([a0.q_proj.weight.data, a0.k_proj.weight.data, a0.v_proj.weight.data]).numpy()
    np.testing.assert_array_equal(qkv.weight.data.numpy(), oracle.quantize(w0, "sym_int4"))
    ids = torch.randint(0, 320, (1, 12), generator=torch.Generator().manual_seed(6))
    ref = oracle_dequantised_copy(tiny_llama(), "sym_int4")
    model = model.to(cuda_device)
    with torch.no_grad():
        got = model.generate(ids.to(cuda_device), max_new_tokens=10, do_sample=False, pad_token_id=0).cpu()
        want = ref.generate(ids, max_new_tokens=10, do_sample=False, pad_token_id=0)
    assert torch.equal(got, want)
    # another format, the stock forwards, a module left alone
    m2 = AutoModelForCausalLM.from_pretrained(d, load_in_low_bit="fp8_e4m3", optimize_model=False, modules_to_not_convert=["lm_head"])
    assert m2.model.layers[1].mlp.up_proj.qtype == ggml_tensor_qtype["fp8_e4m3"] and not isinstance(m2.lm_head, LowBitLinear)
    assert hasattr(m2.model.layers[0].self_attn, "q_proj")
    with pytest.raises(RuntimeError):
        AutoModelForCausalLM.from_pretrained(d, load_in_low_bit="q4_k")           # a format outside this build: loud, not silent


def test_mistral_sliding_window_mask_is_honoured(native_lib, cuda_device):
    """Mistral-7B-v0.1 style config (sliding_window set): transformers hands the attention a sliding-window mask; the
    optimised forwards keep every position in the cache (like the reference) and apply the mask"""
    from ipex_llm.transformers import AutoModelForCausalLM
    cfg = transformers.MistralConfig(hidden_size=256, intermediate_size=512, num_hidden_layers=2, num_attention_heads=2,
                                     num_key_value_heads=2, vocab_size=320, max_position_embeddings=256, sliding_window=16,
                                     head_dim=128, tie_word_embeddings=False)
    torch.manual_seed(3)
    base = transformers.MistralForCausalLM(cfg).eval()
    ref = transformers.MistralForCausalLM(cfg).eval()
    ref.load_state_dict(base.state_dict())
    for mod in ref.modules():
        if isinstance(mod, torch.nn.Linear) and mod.in_features % 64 == 0:
            w = mod.weight.data.float().numpy()
            mod.weight.data = torch.from_numpy(oracle.dequantize(oracle.quantize(w, "sym_int4"), w.shape, "sym_int4"))
    model = AutoModelForCausalLM.from_model(base, low_bit="sym_int4").to(cuda_device)
    ids = torch.randint(0, 320, (1, 40), generator=torch.Generator().manual_seed(8))     # longer than the window
    with torch.no_grad():
        want = ref(ids).logits[:, -1:, :]
        got = model(ids.to(cuda_device)).logits.float().cpu()
    assert float((got - want).abs().mean() / want.abs().mean()) < 2e-3
    # half precision, FP8 cache, prompt + new tokens beyond the window: transformers' loop on the patched forwards
    model = model.half()
    with torch.no_grad():
        out = model.generate(ids.to(cuda_device), max_new_tokens=6, do_sample=False, pad_token_id=0)
    assert out.shape == (1, 46)
