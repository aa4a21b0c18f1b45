"""Mistral forwards -- the reference's transformers/models/mistral.py: ``mistral_model_forward`` :55-107, ``merge_qkv``
:110-111, ``mistral_attention_forward`` :114-171.  The bodies are the Llama ones (models/llama.py) bound to the Mistral
classes.  Like the reference, the cache keeps every position (no sliding-window eviction); the sliding-window mask
that transformers builds is honoured through ``attention_mask``."""
import torch

from ipex_llm_b200.transformers.models.common import merge_qkv_base
from ipex_llm_b200.transformers.models.llama import decoder_model_forward, decoder_attention_forward


def merge_qkv(module: torch.nn.Module):
    from transformers.models.mistral.modeling_mistral import MistralAttention
    merge_qkv_base(module, MistralAttention)


def _mistral_bindings():
    from transformers.models.mistral import modeling_mistral as m
    return decoder_model_forward(m.MistralModel), decoder_attention_forward(m.apply_rotary_pos_emb)


mistral_model_forward, mistral_attention_forward = _mistral_bindings()
