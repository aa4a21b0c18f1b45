"""Host logic of the optimised llama / mistral forwards, on the CPU with UNQUANTISED linears: ``_optimize_pre``
(q/k/v -> qkv_proj) and ``_optimize_post`` (attention / MLP / RMSNorm / model forwards, KV-cache classes) must leave
the model's function unchanged.  On CPU tensors the patched forwards take their torch branches (the reference's own
off-XPU branches, models/llama.py:166-177, models/common.py:201-205, 287-306), so this checks the plumbing -- head
splitting, cache growth, masks, transformers' generate loop on our cache classes -- without a GPU.  The CUDA branches
are checked in tests/test_model_e2e.py."""
import copy

import pytest
import torch

transformers = pytest.importorskip("transformers")


def tiny(kind, seed=0):
    common = dict(hidden_size=128, intermediate_size=256, num_hidden_layers=2, num_attention_heads=4,
                  num_key_value_heads=2, vocab_size=200, max_position_embeddings=256, tie_word_embeddings=False)
    torch.manual_seed(seed)
    if kind == "llama":
        return transformers.LlamaForCausalLM(transformers.LlamaConfig(**common)).eval()
    return transformers.MistralForCausalLM(transformers.MistralConfig(sliding_window=None, **common)).eval()


@pytest.mark.parametrize("kind", ["llama", "mistral"])
def test_optimised_forwards_keep_the_function(kind):
    from ipex_llm_b200.transformers.convert import _optimize_pre, _optimize_post
    from ipex_llm_b200.transformers.kv import DynamicNormalCache
    ref = tiny(kind)
    opt = _optimize_post(_optimize_pre(copy.deepcopy(ref)))
    attn = opt.model.layers[0].self_attn
    assert hasattr(attn, "qkv_proj") and not hasattr(attn, "q_proj")
    assert attn.qkv_proj.weight.shape == (4 * 32 + 2 * 2 * 32, 128)
    assert torch.equal(attn.qkv_proj.weight[:128], ref.model.layers[0].self_attn.q_proj.weight)
    ids = torch.randint(0, 200, (2, 17), generator=torch.Generator().manual_seed(1))
    with torch.no_grad():
        want = ref(ids).logits
        out = opt(ids, use_cache=True)
    torch.testing.assert_close(out.logits, want, rtol=1e-4, atol=1e-5)
    assert isinstance(out.past_key_values, DynamicNormalCache) and out.past_key_values.get_seq_length() == 17
    # transformers' own generate loop on our cache class, against the untouched model
    with torch.no_grad():
        a = ref.generate(ids, max_new_tokens=12, do_sample=False)
        b = opt.generate(ids, max_new_tokens=12, do_sample=False)
    assert torch.equal(a, b)
    # padded batch: the mask path
    mask = torch.ones_like(ids)
    mask[1, :5] = 0
    with torch.no_grad():
        a = ref.generate(ids, attention_mask=mask, max_new_tokens=6, do_sample=False, pad_token_id=0)
        b = opt.generate(ids, attention_mask=mask, max_new_tokens=6, do_sample=False, pad_token_id=0)
    assert torch.equal(a, b)


def test_normal_cache_grows_in_blocks():
    from ipex_llm_b200.transformers.kv import DynamicNormalCache, KV_ALLOC_BLOCK_LENGTH
    cache = DynamicNormalCache()
    gen = torch.Generator().manual_seed(0)
    parts = [torch.randn(1, 2, n, 16, generator=gen) for n in (10, 1, 1, KV_ALLOC_BLOCK_LENGTH, 1)]
    k, v = cache.update(parts[0], parts[0] * 2, 0)
    ptr = k.data_ptr()
    for p in parts[1:3]:
        k, v = cache.update(p, p * 2, 0)
    assert k.data_ptr() == ptr and k.shape[2] == 12                  # appended in place
    for p in parts[3:]:
        k, v = cache.update(p, p * 2, 0)
    assert k.data_ptr() != ptr                                      # the block ran out: re-allocated and copied
    full = torch.cat(parts, dim=2)
    assert torch.equal(k, full) and torch.equal(v, full * 2)
    assert cache.get_seq_length() == full.shape[2] and len(cache) == 1
    assert torch.equal(cache.key_cache[0], full)


def test_autotp_layers_are_recognised():
    """DeepSpeed AutoTP's sharded linears (reference convert.py:217-228), duck-typed by class name"""
    from ipex_llm_b200.transformers.convert import is_linear_module

    class LinearLayer(torch.nn.Module):
        def __init__(self, weight, bias=None):
            super().__init__()
            self.weight, self.bias = torch.nn.Parameter(weight), bias

    class LinearAllreduce(LinearLayer):
        def __init__(self, weight, bias=None, mp_group=None):
            super().__init__(weight, bias)
            self.mp_group = mp_group

    group = object()
    ok, (i, o, g) = is_linear_module(LinearLayer(torch.zeros(24, 128)))
    assert ok and (i, o, g) == (128, 24, None)
    ok, (i, o, g) = is_linear_module(LinearAllreduce(torch.zeros(128, 64), mp_group=group))
    assert ok and (i, o) == (64, 128) and g is group
    assert is_linear_module(torch.nn.LayerNorm(8)) == (False, None)
