"""Deterministic random-init Llama-2-7B for the token-identity check (shared by make_llama7b_tokens.py, which runs the
CPU side once, and tests/test_llama7b_tokens.py, which runs the GPU side).  Every parameter is drawn on the CPU from
its own generator seeded by the parameter's name, N(0, 0.02^2) (transformers' initializer_range), norms = 1: the
values do not depend on construction order, device or thread count."""
import zlib

import torch

LLAMA2_7B = dict(hidden_size=4096, intermediate_size=11008, num_hidden_layers=32, num_attention_heads=32,
                 num_key_value_heads=32, vocab_size=32000, max_position_embeddings=4096, rms_norm_eps=1e-5,
                 tie_word_embeddings=False)
PROMPT_LEN, NEW_TOKENS = 512, 128          # BASELINE.json configs[0]: 512-in / 128-out, batch 1, greedy


def prompt(vocab=32000, n=PROMPT_LEN):
    return torch.randint(0, vocab, (1, n), generator=torch.Generator().manual_seed(1))


def param_values(name, shape):
    if name.endswith("norm.weight"):
        return torch.ones(shape, dtype=torch.float32)
    g = torch.Generator().manual_seed(zlib.crc32(name.encode()))
    return torch.empty(shape, dtype=torch.float32).normal_(0.0, 0.02, generator=g)


def build(config_overrides=None, transform=None):
    """fp32 LlamaForCausalLM on the CPU.  ``transform(name, tensor) -> tensor`` is applied to every 2-D linear weight
    (the CPU side passes the oracle's quantise -> dequantise)."""
    import transformers
    from ipex_llm_b200.optimize import low_memory_init
    cfg = transformers.LlamaConfig(**dict(LLAMA2_7B, **(config_overrides or {})))
    with low_memory_init():
        model = transformers.LlamaForCausalLM(cfg)
    for mod_name, mod in model.named_modules():
        for pname, p in list(mod._parameters.items()):
            if p is None:
                continue
            name = f"{mod_name}.{pname}" if mod_name else pname
            t = param_values(name, tuple(p.shape))
            if transform is not None and isinstance(mod, torch.nn.Linear) and pname == "weight":
                t = transform(name, t)
            mod._parameters[pname] = torch.nn.Parameter(t, requires_grad=False)
    return model.eval()
