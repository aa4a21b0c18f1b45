"""One decode step of a llama / mistral-family model recorded in a CUDA graph, and the greedy ``generate`` that
replays it.

Why it exists: at B200 speeds a decoded token of Llama-2-7B is about a millisecond of GPU work in ~300 kernels; issuing
them from Python (the way the reference's XPU path runs its patched forwards,
/root/reference/python/llm/src/ipex_llm/transformers/models/llama.py:129-205) costs ten times that.  The step below is
those same forwards -- qkv GEMV, rotary + FP8 cache write, fused decode attention, o_proj, fused gate/up MLP, down_proj,
RMSNorm, lm_head, greedy pick -- written against static buffers, with every per-token quantity (cache length, position,
token) in device memory, so that one graph launch per token replays it.  The residual stream is accumulated by the
GEMV epilogues (``accumulate``); nothing runs on the host between tokens.

``install_fast_generate(model)`` (called by ``_optimize_post``) puts ``generate_greedy_graph`` in front of
``model.generate``: plain greedy requests (no sampling, beams, processors, padding) on a CUDA model whose KV cache is
the FP8 one take the recorded path, everything else goes to transformers' ``generate`` unchanged (which still runs the
patched forwards).  Both paths produce the same tokens: the recorded step uses the same kernels with the same roundings
as the eager forwards.  Timing attributes follow the reference's BenchmarkWrapper
(utils/benchmark_util_4_47.py:3300-3393): ``first_cost`` (prefill + first token, s), ``rest_cost_mean`` (mean wall time
of every later token, s).
"""
import copy
import os
import time
from collections import OrderedDict

import torch

import ipex_llm_b200.ggml.model.llama.llama_cpp as ggml
from ipex_llm_b200.transformers import kv as kvmod
from ipex_llm_b200.transformers.low_bit_linear import LowBitLinear, lowbit_matmul, mlp_forward
from ipex_llm_b200.transformers.models.utils import SILU, mlp_fusion_check, use_quantize_kv_cache

MAX_BATCH = 32                # rows the decode GEMV takes in one pass (the fused gate/up kernel serves <= 8 of them)
CAPACITY_STEP = 256           # cache capacities are rounded up to this many positions (one graph per capacity)
MAX_GRAPHS = 4                # recorded steps kept per model (each owns its KV arena)


def _stream(device):
    return torch.cuda.current_stream(device).cuda_stream


def local_heads(attn):
    """(query heads, kv heads) of this rank's slice of the merged qkv projection (all of them without sharding)"""
    cfg = attn.config
    total = attn.qkv_proj.out_features // attn.head_dim
    n_kv = total * cfg.num_key_value_heads // (cfg.num_attention_heads + 2 * cfg.num_key_value_heads)
    return total - 2 * n_kv, n_kv


class DecodeGraph:
    """the recorded step for (batch, cache capacity) and the buffers it reads and writes"""

    def __init__(self, model, batch, capacity):
        core, cfg = model.model, model.config
        self.model, self.core, self.B, self.capacity = model, core, batch, capacity
        dev, dtype = core.norm.weight.device, core.norm.weight.dtype
        self.device, self.dtype = dev, dtype
        attn0 = core.layers[0].self_attn
        self.Hq, self.Hkv = local_heads(attn0)
        self.D, self.H = attn0.head_dim, cfg.hidden_size
        self.eps = cfg.rms_norm_eps
        self.n_layers = len(core.layers)
        tot = (self.Hq + 2 * self.Hkv) * self.D
        B = batch
        z = lambda *shape, dt=dtype: torch.zeros(*shape, dtype=dt, device=dev)           # noqa: E731
        # {tokens in the cache, tokens generated, arrival counter, suppress_below, n_suppress, -, -, -}
        self.state = z(8, dt=torch.int32)
        self.suppress_ids = z(8, dt=torch.int64)    # ids left out of the pick while tokens generated < suppress_below
        self.next_ids = z(B, dt=torch.int64)
        self.tokens = z(B, capacity, dt=torch.int64)
        self.h, self.xn = z(B, self.H), z(B, self.H)
        self.qkv = z(B, 1, tot)
        self.attn = z(B, self.Hq, 1, self.D)
        mlp0 = core.layers[0].mlp
        self.moe = hasattr(mlp0, "experts") and hasattr(mlp0, "gate")          # mixture-of-experts layers (Mixtral)
        inter = mlp0.experts.intermediate_dim if self.moe else mlp0.gate_proj.out_features
        self.act = z(B, inter)
        self.logits = z(B, model.lm_head.out_features)
        self.k_cache = [z(B, self.Hkv, capacity, self.D, dt=torch.uint8) for _ in range(self.n_layers)]
        self.v_cache = [z(B, self.Hkv, capacity, self.D, dt=torch.uint8) for _ in range(self.n_layers)]
        with torch.no_grad():
            pos = torch.arange(capacity, device=dev)[None, :]
            cos, sin = core.rotary_emb(self.h[:1, None, :], position_ids=pos)           # [1, capacity, D], model dtype
        self.cos, self.sin = cos[0].contiguous(), sin[0].contiguous()
        self.sdp_ws = torch.empty(max(256, ggml.lowbit_sdp_fp8_workspace_size(B, self.Hq, self.Hkv, 1, capacity, self.D)),
                                  dtype=torch.uint8, device=dev)
        self.fuse_mlp = cfg.hidden_act == "silu"
        self.launches_per_step = 0
        # host-in-the-loop variant (streaming, end-of-sequence checks, host-side token choice): the step's input token
        # comes from pinned host memory and its output token goes back there, as copy nodes of the same graph
        self.tok_in_host = torch.zeros(B, dtype=torch.int64).pin_memory()
        self.tok_out_host = torch.zeros(B, dtype=torch.int64).pin_memory()
        self.graph = self._record(self._step)
        self.graph_host = None

    # ---- pieces of the step
    def _linear(self, lin, x, out, accumulate=False):
        """out (+)= lin(x): the GEMV with its fused epilogue when the module allows it, the module's own forward
        (tensor-parallel groups, unquantised or act-order linears) otherwise"""
        if isinstance(lin, LowBitLinear) and lin.mp_group is None and not lin.act_order and lin.weight.data.device.type == "cuda" \
                and (lin.bias is None or lin.bias.dtype == out.dtype):
            lowbit_matmul(x, lin.weight.data, lin.weight_shape, lin.qtype, out=out, bias=lin.bias, accumulate=accumulate)
            self.launches_per_step += ggml.last_launch_count()
        elif isinstance(lin, LowBitLinear) and hasattr(lin.mp_group, "matmul_all_reduce") and not lin.act_order \
                and lin.mp_group.fused_supported(x.shape[0], lin.out_len, lin.qtype):
            # row-parallel linear under tensor parallelism: matmul, all-reduce, bias and residual in one kernel
            lin.mp_group.matmul_all_reduce(x, lin.weight.data, lin.weight_shape, lin.qtype, bias=lin.bias, out=out, accumulate=accumulate)
            self.launches_per_step += ggml.last_launch_count()
        else:
            y = lin(x)
            if accumulate:
                out.add_(y)
            else:
                out.copy_(y)

    def _moe(self, block):
        """router (torch: an 8-row matmul, softmax, top-k) + the routed experts' MLPs of the token in two launches
        (moe_forward_vec reads the routed ids on the device), accumulated into the residual stream"""
        from ipex_llm_b200.transformers.models.mixtral import LowBitExperts
        _, weights, index = block.gate(self.xn)
        if self.B == 1 and isinstance(block.experts, LowBitExperts):
            block.experts.fused_one_token(self.xn, index, weights, out=self.h, accumulate=True)
            self.launches_per_step += ggml.last_launch_count()
        else:
            self.h.add_(block.experts(self.xn, index, weights))

    def _rms(self, norm, x, out):
        ggml.lowbit_rms_norm(x, norm.weight, self.eps, out, _stream(self.device))
        self.launches_per_step += 1

    def _step(self):
        core, st, B = self.core, _stream(self.device), self.B
        self.launches_per_step = 0
        emb = core.embed_tokens
        if type(emb) is torch.nn.Embedding and emb.weight.dtype == self.dtype:
            torch.index_select(emb.weight, 0, self.next_ids, out=self.h)
        else:
            self.h.copy_(emb(self.next_ids))
        qkv2d, attn2d = self.qkv.view(B, -1), self.attn.view(B, -1)
        for i, layer in enumerate(core.layers):
            attn, mlp = layer.self_attn, layer.mlp
            self._rms(layer.input_layernorm, self.h, self.xn)
            self._linear(attn.qkv_proj, self.xn, qkv2d)
            # cache_complete: the rows below the cache length and the length itself come from earlier steps, and every
            # step starts with an ordinarily launched kernel (the embedding lookup above) that nothing overtakes
            ggml.lowbit_decode_attention(self.qkv, self.Hq, self.Hkv, self.D, self.cos, self.sin, self.state,
                                         self.k_cache[i], self.v_cache[i], attn.scaling, self.attn, self.sdp_ws, st,
                                         cache_complete=True)
            self.launches_per_step += ggml.last_launch_count()
            self._linear(attn.o_proj, attn2d, self.h, accumulate=True)
            self._rms(layer.post_attention_layernorm, self.h, self.xn)
            if self.moe:
                self._moe(mlp)
                continue
            gate, up = mlp.gate_proj, mlp.up_proj
            if self.fuse_mlp and isinstance(gate, LowBitLinear) and isinstance(up, LowBitLinear) and gate.bias is None \
                    and up.bias is None and gate.mp_group is None and gate.qtype == up.qtype \
                    and gate.weight.data.device.type == "cuda":
                # one launch for <= 8 sequences; lowbit_mlp_cuda runs two matmuls and one activation kernel beyond
                mlp_forward(self.xn, gate.weight.data, up.weight.data, B, self.H, gate.out_len, SILU, gate.qtype, out=self.act)
                self.launches_per_step += ggml.last_launch_count()
            else:
                self.act.copy_(mlp.act_fn(gate(self.xn)) * up(self.xn))
            self._linear(mlp.down_proj, self.act, self.h, accumulate=True)
        self._rms(core.norm, self.h, self.xn)
        self._linear(self.model.lm_head, self.xn, self.logits)
        ggml.lowbit_greedy_advance(self.logits, self.next_ids, self.tokens, self.state, st, self.suppress_ids)
        self.launches_per_step += 1

    def _step_host_io(self):
        self.next_ids.copy_(self.tok_in_host, non_blocking=True)           # H2D, 8 bytes per sequence
        self._step()
        self.tok_out_host.copy_(self.next_ids, non_blocking=True)          # D2H, 8 bytes per sequence

    def _record(self, step):
        use_pdl = os.environ.get("IPEX_LLM_B200_DECODE_PDL", "1") != "0"
        old = ggml.lowbit_get_programmatic_dependent_launch()
        ggml.lowbit_set_programmatic_dependent_launch(use_pdl)
        try:
            stream = torch.cuda.Stream(self.device)
            stream.wait_stream(torch.cuda.current_stream(self.device))
            with torch.cuda.device(self.device), torch.cuda.stream(stream), torch.no_grad():
                saved = self.state.clone(), self.next_ids.clone()
                for _ in range(2):                          # sizes every workspace; the cache rows it writes lie at or
                    step()                                  # beyond the current length and are overwritten later
                stream.synchronize()
                graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(graph, stream=stream):
                    step()
                self.state.copy_(saved[0])
                self.next_ids.copy_(saved[1])
            torch.cuda.current_stream(self.device).wait_stream(stream)
            torch.cuda.synchronize(self.device)
            return graph
        finally:
            ggml.lowbit_set_programmatic_dependent_launch(old)

    # ---- one generate() call
    def bind_cache(self):
        """a DynamicFp8Cache whose layers write into this step's arenas (for the prefill through the model's forward)"""
        cache = kvmod.DynamicFp8Cache()
        cache.layers = [kvmod.Fp8KVLayer(storage=(self.k_cache[i], self.v_cache[i])) for i in range(self.n_layers)]
        return cache

    def start(self, first_logits, prompt_len, min_new=0, eos=()):
        """after the prefill: pick the first token from its logits [B, V] and point the step's state at the cache.
        ``eos`` ids cannot be picked among the first ``min_new`` tokens (transformers' min_new_tokens rule)."""
        eos = list(eos)[:8]
        self.suppress_ids.copy_(torch.tensor(eos + [-1] * (8 - len(eos)), dtype=torch.int64))
        self.state.copy_(torch.tensor([prompt_len - 1, 0, 0, min_new if eos else 0, len(eos), 0, 0, 0], dtype=torch.int32))
        logits = first_logits.contiguous()
        ggml.lowbit_greedy_advance(logits, self.next_ids, self.tokens, self.state, _stream(self.device), self.suppress_ids)

    def replay(self, n=1):
        for _ in range(n):
            self.graph.replay()

    def replay_host_io(self):
        """one step fed from ``tok_in_host``; returns the chosen tokens on the host (synchronises the stream)"""
        if self.graph_host is None:
            self.graph_host = self._record(self._step_host_io)
        self.graph_host.replay()
        torch.cuda.current_stream(self.device).synchronize()
        return self.tok_out_host.clone()


def _capacity_for(total_len):
    return max(2 * CAPACITY_STEP, -(-total_len // CAPACITY_STEP) * CAPACITY_STEP)


def get_decode_graph(model, batch, total_len):
    graphs = model.__dict__.setdefault("_b200_decode_graphs", OrderedDict())
    core = model.model
    key = (batch, _capacity_for(total_len), core.norm.weight.device, core.norm.weight.dtype)
    g = graphs.get(key)
    if g is None:
        while len(graphs) >= MAX_GRAPHS:
            graphs.popitem(last=False)
        g = graphs[key] = DecodeGraph(model, batch, key[1])
    else:
        graphs.move_to_end(key)
    return g


# ------------------------------------------------------------------ generate
_PLAIN_KEYS = {"input_ids", "attention_mask", "max_new_tokens", "min_new_tokens", "max_length", "do_sample", "num_beams",
               "eos_token_id", "pad_token_id", "use_cache", "temperature", "top_p", "top_k", "streamer"}


def _plan(model, inputs, generation_config, kwargs):
    """the request as (input_ids, max_new, min_new, eos ids, pad id) if the recorded path serves it, else None"""
    if os.environ.get("IPEX_LLM_B200_DECODE_GRAPH", "1") == "0" or not torch.cuda.is_available():
        return None
    if any(k not in _PLAIN_KEYS for k in kwargs):
        return None
    input_ids = kwargs.get("input_ids", inputs)
    core = getattr(model, "model", None)
    if core is None or not hasattr(core, "layers") or not hasattr(core.layers[0].self_attn, "qkv_proj"):
        return None
    w = core.norm.weight
    if not (torch.is_tensor(input_ids) and input_ids.dim() == 2 and input_ids.dtype == torch.int64
            and input_ids.device == w.device and w.device.type == "cuda" and 1 <= input_ids.shape[0] <= MAX_BATCH
            and input_ids.shape[1] >= 1):
        return None
    mask = kwargs.get("attention_mask")
    if mask is not None and not bool(mask.to(torch.bool).all()):
        return None                                           # padded batches: positions differ per sequence
    gc = copy.deepcopy(generation_config if generation_config is not None else model.generation_config)
    gc.update(**{k: v for k, v in kwargs.items() if k not in ("input_ids", "attention_mask", "streamer")})
    plain = (not gc.do_sample and (gc.num_beams or 1) == 1 and (getattr(gc, "num_beam_groups", 1) or 1) == 1
             and (gc.repetition_penalty or 1.0) == 1.0 and not gc.no_repeat_ngram_size and not gc.bad_words_ids
             and not getattr(gc, "force_words_ids", None) and (gc.num_return_sequences or 1) == 1
             and not gc.output_scores and not getattr(gc, "output_logits", False) and not gc.return_dict_in_generate
             and not gc.output_attentions and not gc.output_hidden_states and not getattr(gc, "penalty_alpha", None)
             and not getattr(gc, "suppress_tokens", None) and not getattr(gc, "begin_suppress_tokens", None)
             and not getattr(gc, "sequence_bias", None) and getattr(gc, "forced_eos_token_id", None) is None
             and not getattr(gc, "stop_strings", None) and not getattr(gc, "max_time", None)
             and gc.use_cache is not False and getattr(gc, "prompt_lookup_num_tokens", None) is None)
    if not plain:
        return None
    prompt_len = input_ids.shape[1]
    max_new = gc.max_new_tokens if gc.max_new_tokens is not None else (gc.max_length or 20) - prompt_len
    if max_new < 1:
        return None
    min_new = min(gc.min_new_tokens or 0, max_new)
    if gc.min_length and gc.min_length > prompt_len:
        min_new = min(max(min_new, gc.min_length - prompt_len), max_new)
    eos = gc.eos_token_id
    eos = [] if eos is None else ([int(eos)] if not isinstance(eos, (list, tuple)) else [int(e) for e in eos])
    cfg = model.config
    attn = core.layers[0].self_attn
    n_q, n_kv = local_heads(attn)
    # the step is written for the FP8 cache (half-precision models) and full-width rotary embedding; a sliding-window
    # model is served while the whole sequence fits the window (the cache never evicts, like the reference's)
    window = getattr(cfg, "sliding_window", None)
    if not (w.dtype in (torch.float16, torch.bfloat16) and use_quantize_kv_cache(core.layers[0].self_attn.o_proj, w)
            and ggml.lowbit_sdp_fp8_supported(n_q, n_kv, 1, attn.head_dim)
            and (not window or prompt_len + max_new <= window)):
        return None
    partial = getattr(cfg, "partial_rotary_factor", None)
    if partial not in (None, 1, 1.0):
        return None
    # the step indexes cos / sin tables computed once per arena: rotary variants whose frequencies depend on the
    # sequence length seen so far ("dynamic", "longrope") keep to the eager forwards
    if getattr(getattr(core, "rotary_emb", None), "rope_type", "default") not in ("default", "linear", "llama3", "yarn"):
        return None
    pad = gc.pad_token_id if gc.pad_token_id is not None else (eos[0] if eos else 0)
    return input_ids, max_new, min_new, eos, pad


@torch.no_grad()
def generate_greedy_graph(self, inputs=None, generation_config=None, **kwargs):
    """``generate`` of an optimised llama / mistral model.  Plain greedy requests run the recorded decode step; any
    other request is transformers' ``generate`` on the patched forwards."""
    plan = _plan(self, inputs, generation_config, kwargs)
    if plan is None:
        return self._hf_generate(inputs, generation_config=generation_config, **kwargs)
    input_ids, max_new, min_new, eos, pad = plan
    streamer = kwargs.get("streamer")
    B, prompt_len = input_ids.shape
    dev = input_ids.device
    with torch.cuda.device(dev):
        g = get_decode_graph(self, B, prompt_len + max_new)
        if streamer is not None:
            streamer.put(input_ids.cpu())
        torch.cuda.synchronize(dev)
        t0 = time.perf_counter()
        # prefill through the model's own (patched) forward, the FP8 cache writing into the step's arenas
        out = self(input_ids=input_ids, past_key_values=g.bind_cache(), use_cache=True)
        g.start(out.logits[:, -1, :], prompt_len, min_new, eos)
        del out
        torch.cuda.synchronize(dev)
        t1 = time.perf_counter()
        n_done = 1
        check_eos = bool(eos) and min_new < max_new
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        if check_eos or streamer is not None:
            # the host sees every token as it is produced (and feeds it back: it could have changed it)
            eos_t = torch.tensor(eos if eos else [-1])
            finished = torch.zeros(B, dtype=torch.bool)
            tok = g.next_ids.cpu()
            if streamer is not None:
                streamer.put(tok)
            if check_eos and min_new <= 1:
                finished |= torch.isin(tok, eos_t)
            ev0.record()
            while n_done < max_new and not bool(finished.all()):
                g.tok_in_host.copy_(tok)
                tok = g.replay_host_io()
                n_done += 1
                if streamer is not None:
                    streamer.put(tok)
                if check_eos and n_done >= min_new:
                    finished |= torch.isin(tok, eos_t)
            ev1.record()
            if streamer is not None:
                streamer.end()
        else:
            ev0.record()
            g.replay(max_new - 1)
            ev1.record()
            n_done = max_new
        torch.cuda.synchronize(dev)
        t2 = time.perf_counter()
        new_tokens = g.tokens[:, :n_done].clone()
    if check_eos and B > 1:
        # sequences that ended earlier are padded behind their end-of-sequence token, as transformers does
        is_eos = torch.isin(new_tokens, eos_t.to(dev))
        is_eos[:, :max(min_new - 1, 0)] = False
        after = (is_eos.cumsum(1) - is_eos.to(torch.int64)) > 0
        new_tokens = new_tokens.masked_fill(after, pad)
    self.first_cost = t1 - t0
    self.rest_cost_mean = (t2 - t1) / (n_done - 1) if n_done > 1 else 0.0
    self.rest_cost_mean_device = ev0.elapsed_time(ev1) * 1e-3 / (n_done - 1) if n_done > 1 else 0.0    # CUDA events
    self.last_decode_launches_per_token = g.launches_per_step
    return torch.cat([input_ids, new_tokens], dim=1)


def install_fast_generate(model):
    if getattr(model, "_hf_generate", None) is None:
        model._hf_generate = model.generate                      # the bound transformers method
        model.generate = generate_greedy_graph.__get__(model, model.__class__)
    return model
