"""ipex_llm_b200 -- B200-native implementation of ipex-llm's low-bit linear hot path.
Mirrors the reference's import surface: ``optimize_model`` / ``load_low_bit`` / ``low_memory_init`` here,
``AutoModelForCausalLM`` under ``.transformers``, the modules under ``.transformers.low_bit_linear``."""


def __getattr__(name):
    if name in ("optimize_model", "load_low_bit", "low_memory_init"):
        from ipex_llm_b200 import optimize
        return getattr(optimize, name)
    raise AttributeError(name)
