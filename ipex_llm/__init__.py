"""Drop-in alias: ``import ipex_llm...`` resolves to the same module objects as ``ipex_llm_b200...``
(``ipex_llm.transformers.AutoModelForCausalLM``, ``ipex_llm.optimize_model``,
``ipex_llm.transformers.low_bit_linear.LowBitLinear``, ``ipex_llm.ggml.quantize.ggml_tensor_qtype`` ...)."""
import importlib
import importlib.abc
import importlib.util
import sys

import ipex_llm_b200 as _impl


class _AliasLoader(importlib.abc.Loader):
    def __init__(self, real):
        self.real = real

    def create_module(self, spec):
        return importlib.import_module(self.real)

    def exec_module(self, module):
        pass


class _AliasFinder(importlib.abc.MetaPathFinder):
    def find_spec(self, fullname, path, target=None):
        if not fullname.startswith("ipex_llm."):
            return None
        real = "ipex_llm_b200." + fullname[len("ipex_llm."):]
        try:
            if importlib.util.find_spec(real) is None:
                return None
        except ModuleNotFoundError:
            return None
        return importlib.util.spec_from_loader(fullname, _AliasLoader(real))


sys.meta_path.insert(0, _AliasFinder())


def __getattr__(name):
    return getattr(_impl, name)
