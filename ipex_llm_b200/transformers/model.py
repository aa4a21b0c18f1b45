"""``AutoModelForCausalLM.from_pretrained(load_in_4bit= / load_in_low_bit=)``, ``save_low_bit`` and
``load_low_bit`` -- the loader surface of the reference's ipex_llm/transformers/model.py (MD below:
``from_pretrained`` MD:111-388, ``load_convert`` MD:436-530, ``save_low_bit`` MD:59-103,
``load_low_bit`` MD:532-788), written against transformers 5.x without accelerate."""
import json
import os
import types

import torch
import transformers

from ipex_llm_b200.ggml.quantize import ggml_tensor_qtype
from ipex_llm_b200.optimize import PYTORCH_MODEL_NAME, assign_low_bit_state_dict
from ipex_llm_b200.utils.common import invalidInputError


def save_low_bit(self, save_dir, *args, **kwargs):
    """MD:59-103: canonical packed weights in pytorch_model.bin, ``bigdl_transformers_low_bit`` in
    config.json, and load_keys.json"""
    invalidInputError(self.config.to_dict().get("bigdl_transformers_low_bit", False),
                      "Detected this model is not a low-bit model, please use from_pretrained's "
                      "load_in_4bit or load_in_low_bit parameter to load a 4-bit model first.")
    os.makedirs(save_dir, exist_ok=True)
    origin_device = self.device
    self.to("cpu")
    state = self.state_dict()
    torch.save(state, os.path.join(save_dir, PYTORCH_MODEL_NAME))
    self.config.save_pretrained(save_dir)
    if self.can_generate() and getattr(self, "generation_config", None) is not None:
        self.generation_config.save_pretrained(save_dir)
    with open(os.path.join(save_dir, "load_keys.json"), "w") as f:
        json.dump({"all_checkpoint_keys": list(state.keys())}, f)
    if origin_device.type != "cpu":
        self.to(origin_device)


class _BaseAutoModelClass:
    HF_MODEL = None

    @classmethod
    def from_pretrained(cls, *args, **kwargs):
        """MD:111-388.  ``load_in_4bit=True`` == ``load_in_low_bit="sym_int4"`` (MD:327)."""
        load_in_4bit = kwargs.pop("load_in_4bit", False)
        load_in_low_bit = kwargs.pop("load_in_low_bit", None)
        optimize_model = kwargs.pop("optimize_model", True)
        modules_to_not_convert = kwargs.pop("modules_to_not_convert", None)
        mixed_precision = kwargs.pop("mixed_precision", False)
        for k in ("cpu_embedding", "speculative", "pipeline_parallel_stages", "embedding_qtype", "imatrix",
                  "quantization_group_size", "disk_embedding"):
            kwargs.pop(k, None)
        if load_in_4bit or load_in_low_bit:
            q_k = load_in_low_bit if load_in_low_bit else "sym_int4"
            return cls.load_convert(q_k, optimize_model, modules_to_not_convert, *args, mixed_precision=mixed_precision, **kwargs)
        return cls.HF_MODEL.from_pretrained(*args, **kwargs)

    @classmethod
    def load_convert(cls, q_k, optimize_model, modules_to_not_convert, *args, mixed_precision=False, **kwargs):
        """MD:436-530"""
        from ipex_llm_b200.transformers.convert import ggml_convert_low_bit
        invalidInputError(q_k in ggml_tensor_qtype, f"Unknown load_in_low_bit value: {q_k}")
        model = cls.HF_MODEL.from_pretrained(*args, **kwargs)
        model = model.to("cpu")
        model = ggml_convert_low_bit(model, ggml_tensor_qtype[q_k], optimize_model,
                                     modules_to_not_convert=modules_to_not_convert,
                                     torch_dtype=kwargs.get("torch_dtype", "auto"), mixed_precision=mixed_precision)
        return cls._finish(model, q_k, mixed_precision, optimize_model)

    @classmethod
    def from_model(cls, model, low_bit="sym_int4", optimize_model=True, modules_to_not_convert=None, mixed_precision=False):
        """convert an already-built Hugging Face model (random-init models in tests and benchmarks).  A model that
        already sits on a GPU is quantised there, linear by linear (FP4Params.to quantises on the weight's device)."""
        from ipex_llm_b200.transformers.convert import ggml_convert_low_bit
        first = next(model.parameters(), None)
        if first is None or first.device.type != "cuda":
            model = model.to("cpu")
        model = ggml_convert_low_bit(model, ggml_tensor_qtype[low_bit], optimize_model,
                                     modules_to_not_convert=modules_to_not_convert, mixed_precision=mixed_precision)
        return cls._finish(model, low_bit, mixed_precision, optimize_model)

    @staticmethod
    def _finish(model, q_k, mixed_precision=False, optimize_model=True):
        # bigdl_optimize_model: whether q/k/v were merged (the saved keys differ); read back by load_low_bit
        model.config.update({"bigdl_transformers_low_bit": q_k, "bigdl_optimize_model": bool(optimize_model)})
        if mixed_precision:                              # recorded so that load_low_bit sizes the lm_head correctly
            from ipex_llm_b200.transformers.convert import MIXED_LM_HEAD_QTYPE
            model.config.update({"bigdl_lm_head_low_bit": MIXED_LM_HEAD_QTYPE})
        model.save_low_bit = types.MethodType(save_low_bit, model)
        return model

    @classmethod
    def load_low_bit(cls, pretrained_model_name_or_path, *model_args, **kwargs):
        """MD:532-788: build the architecture on meta, size the packed buffers
        (``convert_shape_only``), then install the saved canonical bytes."""
        from ipex_llm_b200.transformers.convert import ggml_convert_low_bit
        config = transformers.AutoConfig.from_pretrained(pretrained_model_name_or_path)
        low_bit = config.to_dict().get("bigdl_transformers_low_bit", False)
        invalidInputError(low_bit, "Detect this model is not a low-bit model, Please use from_pretrained "
                                   "with load_in_4bit or load_in_low_bit to get a low-bit model , and "
                                   " serialize the model using save_low_bit first.")
        invalidInputError(low_bit in ggml_tensor_qtype, f"Unknown bigdl_transformers_low_bit value: {low_bit}")
        from ipex_llm_b200.optimize import low_memory_init
        with low_memory_init():
            model = cls.HF_MODEL.from_config(config)
        mixed_precision = bool(kwargs.pop("mixed_precision", False)) or config.to_dict().get("bigdl_lm_head_low_bit") is not None
        optimize_model = bool(kwargs.pop("optimize_model", config.to_dict().get("bigdl_optimize_model", True)))
        model = ggml_convert_low_bit(model, ggml_tensor_qtype[low_bit], optimize_model,
                                     convert_shape_only=True, device="meta",
                                     modules_to_not_convert=kwargs.pop("modules_to_not_convert", None),
                                     mixed_precision=mixed_precision)
        state = torch.load(os.path.join(pretrained_model_name_or_path, PYTORCH_MODEL_NAME), map_location="cpu",
                           weights_only=True)
        assign_low_bit_state_dict(model, state)
        if hasattr(model, "tie_weights"):
            model.tie_weights()
        model.eval()
        return cls._finish(model, low_bit, optimize_model=optimize_model)


class AutoModelForCausalLM(_BaseAutoModelClass):
    HF_MODEL = transformers.AutoModelForCausalLM


class AutoModel(_BaseAutoModelClass):
    HF_MODEL = transformers.AutoModel
