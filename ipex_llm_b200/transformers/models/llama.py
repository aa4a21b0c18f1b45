"""Llama forwards that call the low-bit linears and the fused operators around them -- the B200 readings of the
reference's transformers/models/llama.py (LL below): ``llama_model_forward`` LL:56-115 (picks the KV-cache class),
``merge_qkv`` LL:118-119, ``llama_attention_forward`` LL:129-205.  Written against transformers 5.x, whose attention
modules receive ``position_embeddings`` (the reference's ``transformers >= 4.46`` branch, LL:160-165) and return
``(attn_output, attn_weights)``.  Mistral's forwards (models/mistral.py) are the same code."""
from typing import Optional

import torch

from ipex_llm_b200.transformers.kv import DynamicFp8Cache, DynamicNormalCache
from ipex_llm_b200.transformers.models.common import merge_qkv_base, scaled_dot_product_attention
from ipex_llm_b200.transformers.models.common import rotary_half_with_cache_inplaced
from ipex_llm_b200.transformers.models.utils import use_quantize_kv_cache


def pick_kv_cache(self, inputs, past_key_values, use_cache):
    """LL:66-90 without the SnapKV branch (KV compression is out of scope): FP8 cache or block-grown fp cache"""
    if not use_cache:
        return past_key_values
    use_quantize_kv = use_quantize_kv_cache(self.layers[0].self_attn.o_proj, inputs, self.config.num_attention_heads,
                                            self.config.num_key_value_heads)
    if use_quantize_kv and not isinstance(past_key_values, DynamicFp8Cache):
        past_key_values = DynamicFp8Cache.from_legacy_cache(past_key_values)
    elif not use_quantize_kv and not isinstance(past_key_values, DynamicNormalCache):
        past_key_values = DynamicNormalCache.from_legacy_cache(past_key_values)
    return past_key_values


def decoder_model_forward(base_class):
    """LL:56-115 / mistral.py:55-107 for a given transformers model class"""

    def model_forward(self, input_ids: torch.LongTensor = None, attention_mask: Optional[torch.Tensor] = None,
                      position_ids: Optional[torch.LongTensor] = None, past_key_values=None,
                      inputs_embeds: Optional[torch.FloatTensor] = None, use_cache: Optional[bool] = None, **kwargs):
        inputs = inputs_embeds if inputs_embeds is not None else self.norm.weight      # device and compute dtype of the model
        use_cache = use_cache if use_cache is not None else self.config.use_cache
        if input_ids is not None and input_ids.device.type == "cuda":
            use_cache = True                                                # LL:70
        past_key_values = pick_kv_cache(self, inputs, past_key_values, use_cache)
        return base_class.forward(self, input_ids=input_ids, attention_mask=attention_mask, position_ids=position_ids,
                                  past_key_values=past_key_values, inputs_embeds=inputs_embeds, use_cache=use_cache, **kwargs)

    return model_forward


def merge_qkv(module: torch.nn.Module):
    from transformers.models.llama.modeling_llama import LlamaAttention
    merge_qkv_base(module, LlamaAttention)


def decoder_attention_forward(apply_rotary_pos_emb):
    """LL:129-205 / mistral.py:114-171.  ``apply_rotary_pos_emb`` = the model family's torch rotary (CPU tensors)."""

    def attention_forward(self, hidden_states: torch.Tensor, position_embeddings=None, attention_mask: Optional[torch.Tensor] = None,
                          past_key_values=None, **kwargs):
        bsz, q_len, _ = hidden_states.size()
        n_heads, n_kv_heads, head_dim = self.config.num_attention_heads, self.config.num_key_value_heads, self.head_dim

        qkv_flat = self.qkv_proj(hidden_states)
        qkv = qkv_flat.view(bsz, q_len, n_heads + 2 * n_kv_heads, head_dim).transpose(1, 2)
        query_states, key_states, value_states = qkv.split([n_heads, n_kv_heads, n_kv_heads], dim=1)

        cos, sin = position_embeddings
        cached = None
        if (q_len > 1 and qkv_flat.device.type == "cuda" and cos.shape[-1] == head_dim and cos.shape[0] == 1 and cos.shape[1] == q_len
                and qkv_flat.dtype in (torch.float16, torch.bfloat16) and cos.dtype == qkv_flat.dtype and qkv_flat.is_contiguous()
                and cos.stride(-1) == 1 and sin.stride(-1) == 1 and hasattr(past_key_values, "rope_append")):
            # prompt into an empty FP8 cache: rotary + the cache write of all its tokens in one launch (the bytes of the two
            # separate operators below: tests/test_decode_ops.py)
            cached = past_key_values.rope_append(self.layer_idx, qkv_flat, n_heads, n_kv_heads, head_dim, cos[0], sin[0])
        if cached is not None:
            key_states, value_states = cached
        else:
            if query_states.device.type == "cuda" and cos.shape[-1] == head_dim:
                rotary_half_with_cache_inplaced(query_states, key_states, cos, sin)     # in place, on the views of qkv
            else:
                query_states, key_states = apply_rotary_pos_emb(query_states, key_states, cos, sin)
            if past_key_values is not None:
                key_states, value_states = past_key_values.update(key_states, value_states, self.layer_idx)

        attn_output = scaled_dot_product_attention(query_states, key_states, value_states, attention_mask,
                                                   q_len == key_states.size(2), self.scaling)
        attn_output = attn_output.transpose(1, 2).contiguous().reshape(bsz, q_len, -1)
        return self.o_proj(attn_output), None

    return attention_forward


def _llama_bindings():
    from transformers.models.llama import modeling_llama as m
    return decoder_model_forward(m.LlamaModel), decoder_attention_forward(m.apply_rotary_pos_emb)


llama_model_forward, llama_attention_forward = _llama_bindings()
