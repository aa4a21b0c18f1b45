"""``ipex_llm.transformers`` surface: the loader classes are imported lazily (they pull in
Hugging Face transformers); the low-bit modules are importable directly."""


def __getattr__(name):
    if name in ("AutoModelForCausalLM", "AutoModel"):
        from ipex_llm_b200.transformers import model
        return getattr(model, name)
    raise AttributeError(name)
