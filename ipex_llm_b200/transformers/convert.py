"""Model conversion: swap eligible ``nn.Linear`` modules for ``LowBitLinear`` -- the part of the
reference's ipex_llm/transformers/convert.py (CV below) that CONSTRUCTS the hot path
(``is_linear_module`` CV:138-234, ``is_lm_head`` CV:126-130, ``use_scale_search`` CV:459-469,
``_replace_with_low_bit_linear`` CV:472-752, ``ggml_convert_low_bit`` CV:1092-1193).

``_optimize_pre`` (CV:890-1089) and ``_optimize_post`` (CV:1275-2140) are built for the two model families the
hot path's configs name, llama (CV:1000-1003, 1327-1340) and mistral (CV:1007-1009, 1724-1735): q/k/v merged
into one ``qkv_proj`` before quantisation, and the attention / MLP / RMSNorm / model forwards replaced by the ones
in ``transformers/models`` that call the fused operators.  ``optimize_model=False`` leaves the stock Hugging Face
forwards on top of ``LowBitLinear``.  The reference's 40 other model families are out of scope.
"""
import types
import logging
import os
import warnings

import torch
from torch import nn

from ipex_llm_b200.ggml.quantize import ggml_tensor_qtype, B200_SUPPORTED_QTYPE_IDS
from ipex_llm_b200.utils.common import invalidInputError

logger = logging.getLogger("ipex_llm_b200")


def is_linear_module(module):
    """CV:138-234: (is_linear, (in, out, mp_group)).  Besides ``nn.Linear`` this recognises DeepSpeed AutoTP's
    already-sharded layers (CV:217-228) by class name -- ``LinearLayer`` (column-parallel: no communication) and
    ``LinearAllreduce`` (row-parallel: carries ``mp_group``, the output is summed over it) -- whose ``weight`` is the
    local ``[out, in]`` shard; DeepSpeed itself is not imported."""
    from ipex_llm_b200.transformers.low_bit_linear import LowBitLinear
    if isinstance(module, nn.Linear) and not isinstance(module, LowBitLinear):
        return True, (module.in_features, module.out_features, None)
    cls = module.__class__.__name__
    if cls in ("LinearLayer", "LinearAllreduce") and getattr(module, "weight", None) is not None and module.weight.dim() == 2:
        out_features, in_features = module.weight.shape
        mp_group = getattr(module, "mp_group", None) if cls == "LinearAllreduce" else None
        return True, (in_features, out_features, mp_group)
    return False, None


def _is_fused_experts(module):
    from ipex_llm_b200.transformers.models.mixtral import is_fused_experts
    return is_fused_experts(module)


def is_lm_head(name, model_config, out_features):
    """CV:126-130"""
    return name == "lm_head" or getattr(model_config, "vocab_size", None) == out_features


def use_scale_search(model_config, qtype):
    """CV:459-469: scale search for fp6 and fp8_e4m3, with the reference's model exceptions"""
    model_type = getattr(model_config, "model_type", None)
    if qtype == ggml_tensor_qtype["fp6"] and model_type not in ["qwen2"]:
        return True
    if qtype == ggml_tensor_qtype["fp8_e4m3"] and model_type not in ["qwen2", "baichuan"]:
        if model_type == "llama" and getattr(model_config, "vocab_size", None) == 128256 and \
                "instruct" in str(getattr(model_config, "_name_or_path", "")).lower():
            return False
        return True
    return False


MIXED_LM_HEAD_QTYPE = "fp6"     # lm_head format under mixed_precision (the reference: q6_k)


def _replace_with_low_bit_linear(model, qtype, modules_to_not_convert=None, convert_shape_only=False,
                                 prefix_name="", model_config=None, enable_scale_search=False, embedding_qtype=None,
                                 mixed_precision=False):
    """CV:472-752.  Linears whose in_features is not a multiple of 64 stay as they are (CV:615-617);
    lm_head (by name or vocab-sized output) gets ``optimize_lm_head`` unless IPEX_LLM_LAST_LM_HEAD=0
    (CV:504-513).  ``mixed_precision`` (CV:622-626) keeps a 4-bit model's lm_head at a finer format: the reference
    picks q6_k, a k-quant (out of scope here); this build uses its own 6-bit block format, fp6 (MIXED_LM_HEAD_QTYPE).
    Quantisation happens inside ``FP4Params.to(device)`` of the module's own device."""
    from ipex_llm_b200.transformers.low_bit_linear import LowBitLinear, FP4Params
    modules_to_not_convert = modules_to_not_convert or []
    has_been_replaced = False
    for name, module in model.named_children():
        is_linear, linear_args = is_linear_module(module)
        full_module_name = prefix_name + "." + name if prefix_name != "" else name
        if any(key in full_module_name for key in modules_to_not_convert):
            continue
        if is_linear:
            in_features, out_features, mp_group = linear_args
            if in_features % 64 != 0:
                continue
            optimize_lm_head = is_lm_head(name, model_config, out_features) and \
                os.environ.get("IPEX_LLM_LAST_LM_HEAD", None) != "0"
            cur_qtype, cur_scale_search = qtype, enable_scale_search
            if mixed_precision and is_lm_head(name, model_config, out_features) and \
                    qtype in (ggml_tensor_qtype["sym_int4"], ggml_tensor_qtype["asym_int4"]):
                cur_qtype = ggml_tensor_qtype[MIXED_LM_HEAD_QTYPE]
                cur_scale_search = use_scale_search(model_config, cur_qtype)
            with torch.device("meta"):
                new_linear = LowBitLinear(in_features, out_features, cur_qtype, module.bias is not None,
                                          mp_group=mp_group, optimize_lm_head=optimize_lm_head,
                                          enable_scale_search=cur_scale_search)
            device = module.weight.data.device
            new_linear._parameters["weight"] = FP4Params(data=module.weight.data, requires_grad=False, quantized=False,
                                                         _shape=None, convert_shape_only=convert_shape_only,
                                                         qtype=cur_qtype, in_features=in_features,
                                                         enable_scale_search=cur_scale_search).to(device)
            if module.bias is not None:
                new_linear._parameters["bias"] = nn.Parameter(module.bias.data, requires_grad=False).to(device)
            if not module.training:
                new_linear.eval()
            model._modules[name] = new_linear
            model._modules[name].requires_grad_(False)
            has_been_replaced = True
            module.weight = None
        elif _is_fused_experts(module) and module.gate_up_proj.shape[2] % 64 == 0 and module.down_proj.shape[2] % 64 == 0:
            # transformers >= 5 mixture-of-experts layers: every expert of the layer in two 3-D parameters (SURVEY 7,
            # "hard parts"); they become per-expert LowBitLinear modules, quantised like any other linear
            from ipex_llm_b200.transformers.models.mixtral import LowBitExperts
            model._modules[name] = LowBitExperts.from_fused(module, qtype, convert_shape_only, enable_scale_search)
            has_been_replaced = True
            continue
        elif embedding_qtype is not None and type(module) == nn.Embedding and module.embedding_dim % 64 == 0:
            # CV:730-733: the embedding table in a low-bit block format, dequantised row by row on the GPU
            from ipex_llm_b200.transformers.embedding import LowBitEmbedding
            model._modules[name] = LowBitEmbedding.from_embedding(module, convert_shape_only, embedding_qtype)
        if len(list(module.children())) > 0:
            _, flag = _replace_with_low_bit_linear(module, qtype, modules_to_not_convert, convert_shape_only,
                                                   prefix_name=full_module_name, model_config=model_config,
                                                   enable_scale_search=enable_scale_search,
                                                   embedding_qtype=embedding_qtype, mixed_precision=mixed_precision)
            has_been_replaced = flag or has_been_replaced
    return model, has_been_replaced


def convert_bigdl_other_module(model, dtype):
    """CV:1196-1229: every floating-point parameter / buffer outside the low-bit linears -> dtype"""
    from ipex_llm_b200.transformers.low_bit_linear import LowBitLinear
    for module in model.modules():
        if list(module.children()) or isinstance(module, LowBitLinear):
            continue
        module.to(dtype)


def convert_forward(m, target_m, new_forward):
    """CV:1232-1242 (convert_forward): rebind ``forward`` on every module of class ``target_m``"""
    for _, sub_m in m.named_children():
        if sub_m.__class__ == target_m:
            bound_method = new_forward.__get__(sub_m, sub_m.__class__)
            setattr(sub_m, "forward", bound_method)
        convert_forward(sub_m, target_m, new_forward)


def _optimize_pre(model, qtype=None):
    """CV:890-1089, llama (CV:1000-1003) and mistral (CV:1007-1009) branches: merge q/k/v into ``qkv_proj``
    BEFORE the linears are quantised (one GEMV instead of three per layer)."""
    model_type = getattr(getattr(model, "config", None), "model_type", None)
    if model_type == "llama":
        from ipex_llm_b200.transformers.models.llama import merge_qkv
        model.apply(merge_qkv)
    elif model_type == "mistral":
        from ipex_llm_b200.transformers.models.mistral import merge_qkv
        model.apply(merge_qkv)
    elif model_type == "mixtral":
        from ipex_llm_b200.transformers.models.mixtral import merge_qkv
        model.apply(merge_qkv)
    return model


def _optimize_post(model):
    """CV:1275-2140, llama (CV:1327-1340) and mistral (CV:1724-1735) branches: RMSNorm -> rms_norm_forward, MLP ->
    mlp_silu_forward (fused gate/up at decode sizes), Model -> *_model_forward (KV-cache class), Attention ->
    *_attention_forward (qkv GEMV, in-place rotary, FP8 cache, fused decode attention).  Also installs the recorded
    decode step behind ``generate`` (transformers/decode_graph.py)."""
    import importlib
    model_type = getattr(getattr(model, "config", None), "model_type", None)
    from ipex_llm_b200.transformers.models.common import rms_norm_forward, mlp_silu_forward
    if model_type == "llama":
        module = importlib.import_module(model.__class__.__module__)
        from ipex_llm_b200.transformers.models.llama import llama_model_forward, llama_attention_forward
        convert_forward(model, module.LlamaRMSNorm, rms_norm_forward)
        convert_forward(model, module.LlamaMLP, mlp_silu_forward)
        convert_forward(model, module.LlamaModel, llama_model_forward)
        convert_forward(model, module.LlamaAttention, llama_attention_forward)
    elif model_type == "mistral":
        module = importlib.import_module(model.__class__.__module__)
        from ipex_llm_b200.transformers.models.mistral import mistral_model_forward, mistral_attention_forward
        convert_forward(model, module.MistralRMSNorm, rms_norm_forward)
        convert_forward(model, module.MistralMLP, mlp_silu_forward)
        convert_forward(model, module.MistralModel, mistral_model_forward)
        convert_forward(model, module.MistralAttention, mistral_attention_forward)
    elif model_type == "mixtral":
        module = importlib.import_module(model.__class__.__module__)
        from ipex_llm_b200.transformers.models.mixtral import mixtral_model_forward, mixtral_attention_forward
        convert_forward(model, module.MixtralRMSNorm, rms_norm_forward)
        convert_forward(model, module.MixtralModel, mixtral_model_forward)
        convert_forward(model, module.MixtralAttention, mixtral_attention_forward)
    else:
        return model
    if hasattr(model, "generate"):
        from ipex_llm_b200.transformers.decode_graph import install_fast_generate
        install_fast_generate(model)
    return model


def ggml_convert_low_bit(model, qtype, optimize_model=True, convert_shape_only=False, device="cpu",
                         modules_to_not_convert=None, torch_dtype="auto", embedding_qtype=None, mixed_precision=False,
                         **unused):
    """CV:1092-1193.  ``woq_int4`` is the same format as ``sym_int4`` (CV:1101-1109)."""
    if qtype == ggml_tensor_qtype["woq_int4"]:
        qtype = ggml_tensor_qtype["sym_int4"]
    invalidInputError(qtype in B200_SUPPORTED_QTYPE_IDS,
                      f"qtype {qtype} is not implemented by the B200 low-bit kernels")
    model_config = getattr(model, "config", None)
    if optimize_model:
        model = _optimize_pre(model, qtype)                 # CV:1133
    enable_scale_search = use_scale_search(model_config, qtype)
    model, has_been_replaced = _replace_with_low_bit_linear(model, qtype, modules_to_not_convert or [],
                                                            convert_shape_only, model_config=model_config,
                                                            enable_scale_search=enable_scale_search,
                                                            embedding_qtype=embedding_qtype, mixed_precision=mixed_precision)
    if not has_been_replaced:
        warnings.warn("No linear modules were found in your model.")
    elif device == "cpu" and torch_dtype != "auto":
        convert_bigdl_other_module(model, torch_dtype)
    if optimize_model:
        model = _optimize_post(model)                       # CV:1174
    return model


def chain_decode_prefetch(model):
    """Name each LowBitLinear's successor in module order so that the decode GEMV can prefetch the next
    weights into L2 (``LowBitLinear.next_linear``; a pure performance hint, see low_bit_linear.py)."""
    from ipex_llm_b200.transformers.low_bit_linear import LowBitLinear
    lins = [m for m in model.modules() if isinstance(m, LowBitLinear)]
    for a, b in zip(lins, lins[1:] + lins[:1]):
        a.next_linear = b
    return model


def convert_gptq(module, awq=False, llm_awq=False, act_order=False):
    """GPTQ / AWQ 4-bit linear -> asym_int4 block records, without touching the weights' values
    (reference CV:382-456: ``qs[j] = w[j] | w[j + QK/2] << 4``, record = ``d, m = -(zero * scale), qs``).

    ``module`` carries ``qweight`` (int32, eight 4-bit codes per word), ``qzeros`` (int32), ``scales``
    (fp16 [groups, out]), ``bits`` == 4, ``group_size`` (a multiple of 64) and, for GPTQ with
    ``act_order``, ``g_idx``.  GPTQ packs codes along the input dimension and stores zero - 1; AWQ
    packs along the output dimension (``wf`` gives its nibble order) and stores the zero itself.
    Returns (packed uint8 canonical [out, in] asym_int4, g_idx_map or None).
    """
    from ipex_llm_b200.transformers.low_bit_linear import get_block_size
    qk = get_block_size("asym_int4")
    invalidInputError(module.bits == 4, "only 4-bit GPTQ/AWQ checkpoints map onto asym_int4")
    invalidInputError(module.group_size % qk == 0, f"group_size must be a multiple of {qk}")
    shifts = module.wf.reshape(-1).to(torch.int32)                     # bit offset of each of the 8 codes in a word
    scales = module.scales.to(torch.float16)                           # [groups, out]
    groups, out_features = scales.shape

    def unpack_last(t):                                                # [..., n] int32 -> [..., n, 8] codes
        return (t.unsqueeze(-1) >> shifts) & 0xF

    zeros = unpack_last(module.qzeros).reshape(module.qzeros.shape[0], -1)[:, :out_features]
    if not awq:
        zeros = zeros + 1
    if awq:
        codes = unpack_last(module.qweight).reshape(module.qweight.shape[0], -1)      # [in, out]
        if llm_awq:
            codes = codes.t()
    else:
        codes = unpack_last(module.qweight.t()).reshape(out_features, -1).t()       # words run along `in`
    g_idx_map = None
    if act_order and not awq:
        invalidInputError(module.g_idx.shape[0] == codes.shape[0], "g_idx and weight shape mismatch")
        _, g_idx_map = torch.sort(module.g_idx)
        codes = codes[g_idx_map, :]
    in_features = codes.shape[0]
    if llm_awq:
        real = in_features // module.group_size
        zeros, scales = zeros[:, :real].t() if zeros.shape[0] != real else zeros[:real], scales[:real]
    w = codes.t().reshape(out_features, in_features // qk, qk).to(torch.uint8)        # [out, blocks, 64]
    qs = w[:, :, :qk // 2] | (w[:, :, qk // 2:] << 4)
    per_block = module.group_size // qk
    d = scales.t().repeat_interleave(per_block, dim=1).contiguous()                    # [out, blocks]
    z = zeros.t().to(torch.float16).repeat_interleave(per_block, dim=1)
    m = -(z * d)
    rec = torch.cat([d.unsqueeze(-1).view(torch.uint8), m.contiguous().unsqueeze(-1).view(torch.uint8), qs], dim=-1)
    return rec.reshape(-1), g_idx_map


def convert_vllm_awq_or_gptq(module, gptq=False, act_order=False):
    """vLLM's AWQ / GPTQ linear -> asym_int4 block records (reference CV:301-379).  Same repacking as
    ``convert_gptq``; vLLM keeps the bit width and group size on ``module.quant_method.quant_config`` and uses
    AutoAWQ's nibble order (0, 4, 1, 5, 2, 6, 3, 7) for AWQ, the natural order for GPTQ."""
    import types
    cfg = module.quant_method.quant_config
    order = [0, 1, 2, 3, 4, 5, 6, 7] if gptq else [0, 4, 1, 5, 2, 6, 3, 7]
    shim = types.SimpleNamespace(
        qweight=module.qweight, qzeros=module.qzeros, scales=module.scales,
        bits=int(cfg.weight_bits) if gptq else 4, group_size=int(cfg.group_size),
        wf=(torch.tensor(order, dtype=torch.int32) * 4).unsqueeze(0), g_idx=getattr(module, "g_idx", None))
    return convert_gptq(shim, awq=not gptq, llm_awq=False, act_order=act_order)


def merge_quantized_qkv(q, k, v):
    """packed CANONICAL q, k, v (same qtype, same in_features) -> one packed qkv weight: rows are
    block-contiguous in the canonical layout, so this is a byte concatenation on the CPU side
    (reference models/common.py:379-413, which also hops through the CPU layout)."""
    parts = [t.view(torch.uint8) if t.dtype != torch.uint8 else t for t in (q, k, v)]
    out = torch.cat([p.reshape(-1) for p in parts])
    return out.view(q.dtype) if q.dtype != torch.uint8 else out
