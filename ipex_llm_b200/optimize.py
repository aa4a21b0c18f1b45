"""``optimize_model`` / ``load_low_bit`` / ``low_memory_init`` -- the generic-module entry points of the
reference's ipex_llm/optimize.py (OP below: ``optimize_model`` OP:199-267, ``_save_low_bit`` OP:41-57,
``load_low_bit`` OP:125-193, ``low_memory_init`` OP:107-122)."""
import json
import os
import types
from contextlib import contextmanager

import torch

from ipex_llm_b200.ggml.quantize import ggml_tensor_qtype
from ipex_llm_b200.utils.common import invalidInputError

PYTORCH_MODEL_NAME = "pytorch_model.bin"
CONFIG_NAME = "bigdl_config.json"


def _save_low_bit(self, save_dir, *args, **kwargs):
    """OP:41-57: state dict with the packed CANONICAL weights + bigdl_config.json"""
    invalidInputError(self._bigdl_config.get("bigdl_transformers_low_bit", False),
                      "Detected this model is not a low-bit model, please use optimize_model first.")
    os.makedirs(save_dir, exist_ok=True)
    devices = {p.device for p in self.parameters()}
    origin = next(iter(devices)) if len(devices) == 1 else None
    self.to("cpu")                                   # cuda -> cpu restores the canonical layout
    torch.save(self.state_dict(), os.path.join(save_dir, PYTORCH_MODEL_NAME))
    with open(os.path.join(save_dir, CONFIG_NAME), "w") as f:
        json.dump(self._bigdl_config, f)
    if origin is not None and origin.type != "cpu":
        self.to(origin)


def assign_low_bit_state_dict(model, state):
    """Install a saved state dict into a model whose linears were converted with
    ``convert_shape_only`` / on meta: packed tensors become quantised ``FP4Params``."""
    from ipex_llm_b200.transformers.low_bit_linear import LowBitLinear, FP4Params
    missing = []
    for mod_name, mod in model.named_modules():
        for pname, p in list(mod._parameters.items()):
            if p is None:
                continue
            key = f"{mod_name}.{pname}" if mod_name else pname
            if key not in state:
                missing.append(key)
                continue
            t = state[key]
            if isinstance(mod, LowBitLinear) and pname == "weight":
                mod._parameters[pname] = FP4Params(t, requires_grad=False, quantized=True,
                                                   _shape=torch.Size(mod.weight_shape), qtype=mod.qtype)
            else:
                mod._parameters[pname] = torch.nn.Parameter(t, requires_grad=False)
        for bname, b in list(mod._buffers.items()):
            key = f"{mod_name}.{bname}" if mod_name else bname
            if b is not None and key in state:
                mod._buffers[bname] = state[key]
    invalidInputError(not missing, f"checkpoint is missing keys: {missing[:8]}")
    return model


@contextmanager
def low_memory_init():
    """OP:107-122: build a model without keeping its weights.  Like accelerate's ``init_empty_weights``
    (which the reference uses), PARAMETERS land on the meta device as they are registered while buffers
    (rotary tables, masks -- often non-persistent, so absent from checkpoints) are computed for real."""
    old_register = torch.nn.Module.register_parameter

    def register_on_meta(module, name, param):
        old_register(module, name, param)
        if param is not None and module._parameters[name].device.type != "meta":
            p = module._parameters[name]
            module._parameters[name] = torch.nn.Parameter(p.detach().to("meta"), requires_grad=p.requires_grad)

    torch.nn.Module.register_parameter = register_on_meta
    try:
        yield
    finally:
        torch.nn.Module.register_parameter = old_register


def load_low_bit(model, model_path):
    """OP:125-193: ``model`` is the architecture (ideally built under ``low_memory_init()``)"""
    from ipex_llm_b200.transformers.convert import ggml_convert_low_bit
    invalidInputError(isinstance(model, torch.nn.Module), "model should be a instance of `torch.nn.Module`")
    invalidInputError(os.path.isdir(model_path), "model_path should be a valid directory path.")
    with open(os.path.join(model_path, CONFIG_NAME)) as f:
        cfg = json.load(f)
    low_bit = cfg.get("bigdl_transformers_low_bit", False)
    invalidInputError(low_bit, "Detect this model is not a low-bit model, please use optimize_model first.")
    qtype = ggml_tensor_qtype[low_bit]
    # the saved keys follow the forwards the model was saved with (merged qkv_proj or separate q/k/v)
    model = ggml_convert_low_bit(model, qtype=qtype, optimize_model=bool(cfg.get("bigdl_optimize_llm", True)),
                                 convert_shape_only=True, device="meta",
                                 mixed_precision=cfg.get("bigdl_lm_head_low_bit") is not None)
    state = torch.load(os.path.join(model_path, PYTORCH_MODEL_NAME), map_location="cpu", weights_only=True)
    assign_low_bit_state_dict(model, state)
    model._bigdl_config = {"bigdl_transformers_low_bit": low_bit, "bigdl_optimize_llm": bool(cfg.get("bigdl_optimize_llm", True))}
    if cfg.get("bigdl_lm_head_low_bit") is not None:
        model._bigdl_config["bigdl_lm_head_low_bit"] = cfg["bigdl_lm_head_low_bit"]
    model.save_low_bit = types.MethodType(_save_low_bit, model)
    return model


def _embedding_qtype(name):
    """``embedding_qtype=`` of from_pretrained / optimize_model (reference model.py:148, 436-439): the reference
    offers its k-quants there; here any of the in-scope block formats."""
    if name is None:
        return None
    invalidInputError(name in ggml_tensor_qtype, f"Unknown embedding_qtype: {name}")
    return ggml_tensor_qtype[name]


def optimize_model(model, low_bit="sym_int4", optimize_llm=True, modules_to_not_convert=None,
                   cpu_embedding=False, **kwargs):
    """OP:199-267.  Requires the model on cpu or meta (OP:236-239); returns it with its eligible linears
    replaced by ``LowBitLinear`` and a ``save_low_bit`` method attached."""
    from ipex_llm_b200.transformers.convert import ggml_convert_low_bit
    invalidInputError(low_bit is None or low_bit in ggml_tensor_qtype,
                      f"Unknown load_in_low_bit value: {low_bit}, expected: sym_int4, asym_int4, nf4, "
                      f"fp8, fp8_e4m3, fp8_e5m2 or fp6.")
    invalidInputError(isinstance(model, torch.nn.Module),
                      f"model should be an instance of `torch.nn.Module`, but got {type(model)} at last.")
    if hasattr(model, "device") and isinstance(getattr(model, "device"), torch.device):
        invalidInputError(model.device.type in ("cpu", "meta"),
                          f"Expect model on device `cpu` or `meta`, but got device type {model.device.type}")
    if low_bit is not None:
        model = ggml_convert_low_bit(model, qtype=ggml_tensor_qtype[low_bit], optimize_model=optimize_llm,
                                     modules_to_not_convert=modules_to_not_convert,
                                     torch_dtype=kwargs.get("torch_dtype", "auto"),
                                     embedding_qtype=_embedding_qtype(kwargs.get("embedding_qtype", None)),
                                     mixed_precision=bool(kwargs.get("mixed_precision", False)))
    model._bigdl_config = {"bigdl_transformers_low_bit": low_bit, "bigdl_optimize_llm": bool(optimize_llm)}
    if low_bit is not None and kwargs.get("mixed_precision", False):
        from ipex_llm_b200.transformers.convert import MIXED_LM_HEAD_QTYPE
        model._bigdl_config["bigdl_lm_head_low_bit"] = MIXED_LM_HEAD_QTYPE
    model.save_low_bit = types.MethodType(_save_low_bit, model)
    return model
