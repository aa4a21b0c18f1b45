"""Generate golden vectors for the low-bit formats by RUNNING the reference's own in-repo
Python statements of them (run once in the build container; /root/reference does not exist
on the GPU box, so the outputs are committed as tests/golden/*.npz).

Sources executed (never copied into this repo -- they are loaded from /root/reference):
  * utils/convert_chatglm.py:88-133  quantize_q8_0 / quantize_q4_0 / quantize_q4_1  (QK = 64)
  * transformers/gguf/gguf.py:273-296 convert_q4_0_tensor / convert_q4_1_tensor (dequant)
  * transformers/convert.py:382-456  convert_gptq (GPTQ/AWQ -> asym_int4 block records)
The closed native library that actually runs the hot path is absent, so these are the only
executable statements of the block formats the reference holds.  A third, independent
source is the installed `gguf` package (numpy restatement of upstream ggml; QK = 32).

Usage:  python tests/golden/make_golden.py
"""
import ast
import os
import sys
import types

import numpy as np
import torch

REF = "/root/reference/python/llm/src/ipex_llm"
OUT = os.path.dirname(os.path.abspath(__file__))


def _extract_functions(path, names, extra_globals):
    """exec only the named top-level (or class-level) function defs of a reference file."""
    src = open(path).read()
    tree = ast.parse(src)
    ns = dict(extra_globals)
    found = []
    for node in ast.walk(tree):
        if isinstance(node, ast.FunctionDef) and node.name in names:
            mod = ast.Module(body=[node], type_ignores=[])
            exec(compile(mod, path, "exec"), ns)
            found.append(node.name)
    missing = set(names) - set(found)
    if missing:
        raise RuntimeError(f"{path}: missing {missing}")
    return ns


def _invalid_input_error(cond, msg, fix=None):
    if not cond:
        raise RuntimeError(msg)


def chatglm_quantizers():
    g = {"torch": torch, "invalidInputError": _invalid_input_error,
         "GGML_QK8_0": 64, "GGML_QK4_0": 64, "GGML_QK4_1": 64}
    # the constants are re-read from the file so that a change upstream is noticed
    src = open(f"{REF}/utils/convert_chatglm.py").read()
    for line in src.splitlines():
        for c in ("GGML_QK8_0", "GGML_QK4_0", "GGML_QK4_1"):
            if line.startswith(c + " ="):
                g[c] = int(line.split("=")[1])
    ns = _extract_functions(f"{REF}/utils/convert_chatglm.py",
                            ["quantize_q4_0", "quantize_q4_1"], g)
    return ns["quantize_q4_0"], ns["quantize_q4_1"], g["GGML_QK4_0"]


def gguf_dequantizers(qk_bytes):
    ns = _extract_functions(f"{REF}/transformers/gguf/gguf.py",
                            ["convert_q4_0_tensor", "convert_q4_1_tensor"], {"torch": torch})
    holder = types.SimpleNamespace(block_size=qk_bytes)
    return (lambda t, dims: ns["convert_q4_0_tensor"](holder, t, 0, 0, dims),
            lambda t, dims: ns["convert_q4_1_tensor"](holder, t, 0, 0, dims))


def gptq_converter(qk):
    # convert_gptq imports get_block_size from the (unimportable) low_bit_linear module;
    # provide the one value it asks for.
    fake = types.ModuleType("ipex_llm.transformers.low_bit_linear")
    fake.get_block_size = lambda name: qk
    for m in ("ipex_llm", "ipex_llm.transformers"):
        sys.modules.setdefault(m, types.ModuleType(m))
    sys.modules["ipex_llm.transformers.low_bit_linear"] = fake
    ns = _extract_functions(f"{REF}/transformers/convert.py", ["convert_gptq"],
                            {"torch": torch, "invalidInputError": _invalid_input_error})
    return ns["convert_gptq"]


def main():
    torch.manual_seed(1234)
    rng = np.random.default_rng(1234)
    q4_0, q4_1, qk = chatglm_quantizers()
    assert qk == 64

    # ---- (1) reference Python quantisers, QK=64 ---------------------------------
    rows, k = 48, 512
    w = (rng.standard_normal((rows, k)) * 0.02).astype(np.float32)
    w[0, :64] = 0.0                       # all-zero block
    w[1, :64] = 0.013                     # constant block
    w[2, 5] = 0.5                         # single large positive
    w[3, 70] = -0.75                      # single large negative
    wt = torch.from_numpy(w)
    g40 = q4_0(wt.clone()).numpy().view(np.uint8).reshape(-1)
    g41 = q4_1(wt.clone()).numpy().view(np.uint8).reshape(-1)

    # ---- (2) reference GGUF-loader dequantisation applied to those bytes --------
    dq0, dq1 = gguf_dequantizers({2: 2 + qk // 2, 3: 4 + qk // 2})
    d40 = dq0(torch.from_numpy(g40.copy()), (rows, k)).float().numpy()
    d41 = dq1(torch.from_numpy(g41.copy()), (rows, k)).float().numpy()

    # ---- (3) convert_gptq: synthetic GPTQ module -> asym_int4 records ------------
    conv = gptq_converter(qk)
    in_f, out_f, group, bits = 256, 32, 128, 4
    ints = rng.integers(0, 16, size=(in_f, out_f), dtype=np.int64)          # [in, out]
    zeros = rng.integers(0, 15, size=(in_f // group, out_f), dtype=np.int64)  # stored zero-1
    scales = (rng.random((in_f // group, out_f)) * 0.01 + 0.002).astype(np.float16)
    shifts = np.arange(0, 32, bits, dtype=np.int64)
    qweight = np.zeros((in_f // 8, out_f), dtype=np.int64)
    for i in range(8):
        qweight |= ints[i::8, :] << shifts[i]
    qzeros = np.zeros((in_f // group, out_f // 8), dtype=np.int64)
    for i in range(8):
        qzeros |= zeros[:, i::8] << shifts[i]
    mod = types.SimpleNamespace(
        scales=torch.from_numpy(scales),
        qzeros=torch.from_numpy(qzeros.astype(np.uint32).view(np.int32)),
        qweight=torch.from_numpy(qweight.astype(np.uint32).view(np.int32)),
        wf=torch.tensor(shifts, dtype=torch.int32).unsqueeze(0),
        bits=bits, group_size=group, in_features=in_f, g_idx=None)
    packed, _ = conv(mod, awq=False, llm_awq=False, act_order=False)
    gptq_bytes = packed.numpy().view(np.uint8).reshape(-1)
    # the weights those records mean, by GPTQ's definition: (q - (zero+1)) * scale
    zfull = np.repeat(zeros + 1, group, axis=0)
    sfull = np.repeat(scales.astype(np.float32), group, axis=0)
    gptq_w = ((ints - zfull) * sfull).T.astype(np.float32)                   # [out, in]

    np.savez_compressed(os.path.join(OUT, "ref_python_q4.npz"),
                        w=w, q4_0_bytes=g40, q4_1_bytes=g41, q4_0_dequant=d40,
                        q4_1_dequant=d41, qk=np.int64(qk),
                        gptq_ints=ints.astype(np.uint8), gptq_zeros=zeros.astype(np.uint8),
                        gptq_scales=scales, gptq_bytes=gptq_bytes, gptq_weight=gptq_w)

    # ---- (4) installed gguf package, QK=32 (upstream ggml arithmetic) ------------
    import gguf.quants as gq
    w32 = (rng.standard_normal((64, 256)) * 0.05).astype(np.float32)
    w32[0, :32] = 0.0
    np.savez_compressed(os.path.join(OUT, "gguf_q4_qk32.npz"), w=w32,
                        q4_0_bytes=gq.Q4_0.quantize_blocks(w32.reshape(-1, 32)).reshape(-1),
                        q4_1_bytes=gq.Q4_1.quantize_blocks(w32.reshape(-1, 32)).reshape(-1))
    print("wrote", sorted(f for f in os.listdir(OUT) if f.endswith(".npz")))


if __name__ == "__main__":
    main()
