"""Mixtral: low-bit experts and the one-token MoE forward.

transformers >= 5 keeps a layer's experts as two fused 3-D parameters (``gate_up_proj [E, 2I, H]``, ``down_proj
[E, H, I]``) that ``is_linear_module`` never sees.  ``LowBitExperts`` replaces such a module: one ``LowBitLinear`` per
expert and projection (the reference's layout -- ``expert.gate_proj / up_proj / down_proj`` -- which is what its fused
decode path takes the weight addresses from), quantised exactly like any other linear.  Its forward follows the
reference's MoE block forwards (/root/reference/python/llm/src/ipex_llm/transformers/models/qwen3_moe.py:70-132,
deepseek.py:280-330; the fork has no Mixtral forward of its own): one token -> ``moe_forward_vec`` over device tables of
the experts' weight addresses behind the ``use_fuse_moe`` gate; more tokens -> per-expert ``LowBitLinear`` calls over
the tokens routed to each expert.  Attention, RMSNorm and the model forward are the Mistral ones.
"""
import torch
from torch import nn

from ipex_llm_b200.transformers.models.common import merge_qkv_base
from ipex_llm_b200.transformers.models.llama import decoder_model_forward, decoder_attention_forward
from ipex_llm_b200.transformers.models.utils import use_fuse_moe


def merge_qkv(module: torch.nn.Module):
    merge_qkv_base(module, "MixtralAttention")


def is_fused_experts(module) -> bool:
    """a module holding every expert of a layer as [E, 2I, H] / [E, H, I] parameters (transformers >= 5 MixtralExperts
    and its look-alikes)"""
    g, d = getattr(module, "gate_up_proj", None), getattr(module, "down_proj", None)
    return (isinstance(g, nn.Parameter) and isinstance(d, nn.Parameter) and g.dim() == 3 and d.dim() == 3
            and g.shape[0] == d.shape[0] and g.shape[1] == 2 * d.shape[2] and g.shape[2] == d.shape[1])


class LowBitExperts(nn.Module):
    """the experts of one MoE layer as per-expert LowBitLinear modules; call signature of MixtralExperts.forward"""

    def __init__(self, num_experts, hidden_dim, intermediate_dim, qtype, act_fn, enable_scale_search=False):
        super().__init__()
        from ipex_llm_b200.transformers.low_bit_linear import LowBitLinear
        self.num_experts, self.hidden_dim, self.intermediate_dim, self.qtype = num_experts, hidden_dim, intermediate_dim, qtype
        self.act_fn = act_fn

        def lins(i, o):
            return nn.ModuleList([LowBitLinear(i, o, qtype, bias=False, enable_scale_search=enable_scale_search)
                                  for _ in range(num_experts)])

        self.gate_proj = lins(hidden_dim, intermediate_dim)
        self.up_proj = lins(hidden_dim, intermediate_dim)
        self.down_proj = lins(intermediate_dim, hidden_dim)
        self._tables = None

    @classmethod
    def from_fused(cls, module, qtype, convert_shape_only=False, enable_scale_search=False):
        """quantise a fused-experts module; gate = the first I rows of gate_up_proj, up = the last I (the chunk(2) of
        its forward)"""
        from ipex_llm_b200.transformers.low_bit_linear import FP4Params
        E, two_i, H = module.gate_up_proj.shape
        inter = two_i // 2
        device = module.gate_up_proj.device
        with torch.device("meta"):
            new = cls(E, H, inter, qtype, getattr(module, "act_fn", nn.SiLU()), enable_scale_search)
        for e in range(E):
            parts = {"gate_proj": module.gate_up_proj.data[e, :inter], "up_proj": module.gate_up_proj.data[e, inter:],
                     "down_proj": module.down_proj.data[e]}
            for name, w in parts.items():
                lin = getattr(new, name)[e]
                lin._parameters["weight"] = FP4Params(data=w.contiguous(), requires_grad=False, quantized=False, _shape=None,
                                                      convert_shape_only=convert_shape_only, qtype=qtype,
                                                      in_features=w.shape[1], enable_scale_search=enable_scale_search).to(device)
        new.requires_grad_(False)
        if not module.training:
            new.eval()
        return new

    def _address_tables(self, device):
        """u64 device arrays of the experts' weight addresses (qwen3_moe.py:88-99), rebuilt when the weights moved"""
        key = (device, self.gate_proj[0].weight.data.data_ptr(), self.down_proj[-1].weight.data.data_ptr())
        if self._tables is None or self._tables[0] != key:
            def table(mods):
                return torch.tensor([m.weight.data.data_ptr() for m in mods], dtype=torch.int64, device=device)
            self._tables = (key, table(self.gate_proj), table(self.up_proj), table(self.down_proj))
        return self._tables[1:]

    def fused_one_token(self, x, top_k_index, top_k_weights, out=None, accumulate=False):
        from ipex_llm_b200.transformers.low_bit_linear import moe_forward_vec
        gates, ups, downs = self._address_tables(x.device)
        return moe_forward_vec(x, top_k_index, top_k_weights, gates, ups, downs, self.hidden_dim, self.intermediate_dim,
                               self.qtype, out=out, accumulate=accumulate)

    def forward(self, hidden_states, top_k_index, top_k_weights):
        if hidden_states.shape[0] == 1 and use_fuse_moe(hidden_states, self.qtype) and \
                self.gate_proj[0].weight.data.device.type == "cuda" and not self.training:
            return self.fused_one_token(hidden_states, top_k_index, top_k_weights)
        final = torch.zeros_like(hidden_states)
        expert_mask = torch.nn.functional.one_hot(top_k_index, num_classes=self.num_experts).permute(2, 1, 0)
        for e in torch.greater(expert_mask.sum(dim=(-1, -2)), 0).nonzero()[:, 0].tolist():
            top_k_pos, token_idx = torch.where(expert_mask[e])
            cur = hidden_states[token_idx]
            cur = self.act_fn(self.gate_proj[e](cur)) * self.up_proj[e](cur)
            cur = self.down_proj[e](cur) * top_k_weights[token_idx, top_k_pos, None].to(cur.dtype)
            final.index_add_(0, token_idx, cur.to(final.dtype))
        return final


def _mixtral_bindings():
    from transformers.models.mixtral import modeling_mixtral as m
    return decoder_model_forward(m.MixtralModel), decoder_attention_forward(m.apply_rotary_pos_emb)


mixtral_model_forward, mixtral_attention_forward = _mixtral_bindings()
