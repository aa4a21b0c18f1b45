"""Host-side conversion logic that needs no GPU: GPTQ -> asym_int4 repacking against the bytes the
reference's own convert_gptq produced (tests/golden/ref_python_q4.npz, made by make_golden.py), qkv
merging, scale-search policy, meta-device conversion."""
import os
import types

import numpy as np
import pytest
import torch

import oracle

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def test_convert_gptq_matches_reference_bytes(native_lib):
    from ipex_llm_b200.transformers.convert import convert_gptq
    g = np.load(os.path.join(GOLD, "ref_python_q4.npz"))
    ints, zeros, scales = g["gptq_ints"].astype(np.int64), g["gptq_zeros"].astype(np.int64), g["gptq_scales"]
    in_f, out_f = ints.shape
    group = in_f // zeros.shape[0]
    shifts = np.arange(0, 32, 4, dtype=np.int64)
    qweight = np.zeros((in_f // 8, out_f), dtype=np.int64)
    for i in range(8):
        qweight |= ints[i::8, :] << shifts[i]
    qzeros = np.zeros((zeros.shape[0], out_f // 8), dtype=np.int64)
    for i in range(8):
        qzeros |= zeros[:, i::8] << shifts[i]
    mod = types.SimpleNamespace(scales=torch.from_numpy(scales),
                                qzeros=torch.from_numpy(qzeros.astype(np.uint32).view(np.int32)),
                                qweight=torch.from_numpy(qweight.astype(np.uint32).view(np.int32)),
                                wf=torch.tensor(shifts, dtype=torch.int32).unsqueeze(0), bits=4, group_size=group,
                                in_features=in_f, g_idx=None)
    packed, gmap = convert_gptq(mod)
    assert gmap is None
    np.testing.assert_array_equal(packed.numpy(), g["gptq_bytes"])
    # and those records mean GPTQ's weights
    d = oracle.dequantize(packed.numpy(), (out_f, in_f), "asym_int4")
    np.testing.assert_allclose(d, g["gptq_weight"], rtol=0, atol=2e-3 * np.abs(g["gptq_weight"]).max())


def test_merge_quantized_qkv_is_row_concat(native_lib):
    from ipex_llm_b200.transformers.convert import merge_quantized_qkv
    rng = np.random.default_rng(0)
    q, k, v = [(rng.standard_normal((n, 128)) * 0.02).astype(np.float32) for n in (64, 16, 16)]
    merged = merge_quantized_qkv(*[torch.from_numpy(oracle.quantize(t, "sym_int4")) for t in (q, k, v)])
    np.testing.assert_array_equal(merged.numpy(), oracle.quantize(np.concatenate([q, k, v]), "sym_int4"))


def test_scale_search_policy_and_meta_conversion(native_lib):
    from ipex_llm_b200.transformers.convert import use_scale_search, ggml_convert_low_bit
    from ipex_llm_b200.transformers.low_bit_linear import LowBitLinear
    cfg = types.SimpleNamespace(model_type="llama", vocab_size=32000, _name_or_path="x")
    assert use_scale_search(cfg, 29) and use_scale_search(cfg, 15) and not use_scale_search(cfg, 2)
    assert not use_scale_search(types.SimpleNamespace(model_type="qwen2"), 29)
    assert not use_scale_search(types.SimpleNamespace(model_type="llama", vocab_size=128256, _name_or_path="Llama-3-Instruct"), 15)

    class Net(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.a = torch.nn.Linear(128, 64)
            self.odd = torch.nn.Linear(100, 8)          # in_features % 64 != 0: left alone (convert.py:615)
            self.lm_head = torch.nn.Linear(128, 320, bias=False)
    with torch.device("meta"):
        net = Net()
    net = ggml_convert_low_bit(net, 34, convert_shape_only=True, device="meta", modules_to_not_convert=[])   # woq_int4 alias
    assert isinstance(net.a, LowBitLinear) and net.a.qtype == 2 and net.a.weight.data.shape == (64 * 2 * 34,)
    assert isinstance(net.odd, torch.nn.Linear) and not isinstance(net.odd, LowBitLinear)
    assert net.lm_head.optimize_lm_head and not net.a.optimize_lm_head
    with pytest.raises(RuntimeError):
        ggml_convert_low_bit(Net(), 8)                  # sym_int8: not implemented, loud


def test_convert_vllm_awq_or_gptq(native_lib):
    """vLLM flavour (attributes on quant_method.quant_config): GPTQ gives the reference bytes of the golden
    fixture; AWQ (words along the output dimension, AutoAWQ nibble order, zero stored as is) dequantises to
    (code - zero) * scale."""
    from ipex_llm_b200.transformers.convert import convert_vllm_awq_or_gptq
    g = np.load(os.path.join(GOLD, "ref_python_q4.npz"))
    ints, zeros, scales = g["gptq_ints"].astype(np.int64), g["gptq_zeros"].astype(np.int64), g["gptq_scales"]
    in_f, out_f = ints.shape
    group = in_f // zeros.shape[0]
    cfg = types.SimpleNamespace(weight_bits=4, group_size=group)
    qm = types.SimpleNamespace(quant_config=cfg)

    def pack_last(a, order):                      # [r, c] codes -> [r, c / 8] int32 words, code i of a word at nibble order[i]
        w = np.zeros((a.shape[0], a.shape[1] // 8), dtype=np.int64)
        for i in range(8):
            w |= a[:, i::8] << (4 * order[i])
        return torch.from_numpy(w.astype(np.uint32).view(np.int32))

    nat = list(range(8))
    qweight = pack_last(ints.T.copy(), nat).t().contiguous()                   # GPTQ: words along `in`
    mod = types.SimpleNamespace(scales=torch.from_numpy(scales), qzeros=pack_last(zeros, nat), qweight=qweight, quant_method=qm)
    packed, gmap = convert_vllm_awq_or_gptq(mod, gptq=True)
    assert gmap is None
    np.testing.assert_array_equal(packed.numpy(), g["gptq_bytes"])

    rng = np.random.default_rng(7)
    codes = rng.integers(0, 16, size=(in_f, out_f)).astype(np.int64)
    z = rng.integers(0, 16, size=zeros.shape).astype(np.int64)

    def pack_awq(a):                               # AutoAWQ: nibble p of a word holds column awq-order^-1
        w = np.zeros((a.shape[0], a.shape[1] // 8), dtype=np.int64)
        for p_, col in enumerate([0, 2, 4, 6, 1, 3, 5, 7]):
            w |= a[:, col::8] << (4 * p_)
        return torch.from_numpy(w.astype(np.uint32).view(np.int32))

    mod = types.SimpleNamespace(scales=torch.from_numpy(scales), qzeros=pack_awq(z), qweight=pack_awq(codes), quant_method=qm)
    packed, _ = convert_vllm_awq_or_gptq(mod, gptq=False)
    want = (codes.T - np.repeat(z, group, axis=0).T) * np.repeat(scales.astype(np.float32), group, axis=0).T
    got = oracle.dequantize(packed.numpy(), (out_f, in_f), "asym_int4")
    np.testing.assert_allclose(got, want, rtol=0, atol=2e-3 * np.abs(want).max())


def test_mixed_precision_lm_head(native_lib, tmp_path):
    """convert.py:622-626: with mixed_precision a 4-bit model keeps its lm_head at a finer format (the reference:
    q6_k; this build: fp6, 50 bytes per 64 weights); the choice is recorded so that load_low_bit sizes the lm_head
    buffer correctly, and other target formats are left alone"""
    import json, os
    from ipex_llm_b200.transformers.convert import ggml_convert_low_bit, MIXED_LM_HEAD_QTYPE
    from ipex_llm_b200 import optimize
    from ipex_llm_b200.ggml.quantize import ggml_tensor_qtype as Q

    class Net(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.a = torch.nn.Linear(128, 64, bias=False)
            self.lm_head = torch.nn.Linear(128, 320, bias=False)

    def build(low_bit, mixed):
        with torch.device("meta"):
            net = Net()
        return ggml_convert_low_bit(net, Q[low_bit], convert_shape_only=True, device="meta", mixed_precision=mixed)

    net = build("sym_int4", True)
    assert MIXED_LM_HEAD_QTYPE == "fp6"
    assert net.a.qtype == Q["sym_int4"] and net.a.weight.data.shape == (64 * 2 * 34,)
    assert net.lm_head.qtype == Q["fp6"] and net.lm_head.weight.data.shape == (320 * 2 * 50,)
    assert net.lm_head.weight.enable_scale_search          # fp6 is quantised with the scale search (convert.py:459-469)
    assert build("asym_int4", True).lm_head.qtype == Q["fp6"]
    assert build("sym_int4", False).lm_head.qtype == Q["sym_int4"]
    assert build("fp8_e4m3", True).lm_head.qtype == Q["fp8_e4m3"]      # only 4-bit targets are raised

    # save / load: the state dict of a converted model and the recorded lm_head format
    saved = {"a.weight": torch.arange(64 * 2 * 34, dtype=torch.int64).to(torch.uint8),
             "lm_head.weight": torch.arange(320 * 2 * 50, dtype=torch.int64).to(torch.uint8)}
    torch.save(saved, os.path.join(tmp_path, optimize.PYTORCH_MODEL_NAME))
    with open(os.path.join(tmp_path, optimize.CONFIG_NAME), "w") as f:
        json.dump({"bigdl_transformers_low_bit": "sym_int4", "bigdl_lm_head_low_bit": "fp6"}, f)
    with optimize.low_memory_init():
        fresh = Net()
    loaded = optimize.load_low_bit(fresh, str(tmp_path))
    assert loaded.lm_head.qtype == Q["fp6"] and loaded.lm_head.weight.quantized
    assert torch.equal(loaded.lm_head.weight.data, saved["lm_head.weight"]) and torch.equal(loaded.a.weight.data, saved["a.weight"])
    assert loaded._bigdl_config == {"bigdl_transformers_low_bit": "sym_int4", "bigdl_lm_head_low_bit": "fp6",
                                    "bigdl_optimize_llm": True}
