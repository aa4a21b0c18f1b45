"""qtype table of the reference (ipex_llm/ggml/quantize.py:28-64), numbers unchanged so that
saved checkpoints (`bigdl_transformers_low_bit` in config.json) stay readable.

Only the qtypes of the low-bit linear hot path are implemented by libb200lowbit.so:
sym_int4, asym_int4, nf4, fp8_e4m3, fp8_e5m2 (alias fp8), fp6, and woq_int4 as an alias of
sym_int4 (reference convert.py:1101-1109).  The other names are kept so that a caller's
lookup does not KeyError, and are rejected with a clear error when used.
"""

ggml_tensor_qtype = {"sym_int4": 2,
                     "asym_int4": 3,
                     "sym_int5": 6,
                     "asym_int5": 7,
                     "sym_int8": 8,
                     "nf4": 10,
                     "nf3": 11,
                     "fp16": 12,
                     "fp8_e4m3": 15,
                     "fp4": 16,
                     "mixed_fp4": 17,
                     "mixed_fp8": 18,
                     "fp8_e5m2": 19,
                     "fp8": 19,
                     "bf16": 20,
                     "gguf_iq2_xxs": 21,
                     "gguf_iq2_xs": 22,
                     "q2_k": 23,
                     "gguf_iq1_s": 24,
                     "gguf_iq1_m": 25,
                     "q6_k": 26,
                     "q4_k": 27,
                     "q5_k": 28,
                     "fp6": 29,
                     "fp6_k": 30,
                     "sym_int4_rtn": 31,
                     "sym_int8_rtn": 32,
                     "asym_int4_rtn": 33,
                     "woq_int4": 34,
                     "torch_fp8_e5m2": 35,
                     "torch_fp8": 35,
                     "torch_fp8_e4m3": 36}

gguf_mixed_qtype = {"gguf_q4k_s": 101,
                    "gguf_q4k_m": 102}

# what the B200 kernels implement
B200_SUPPORTED_QTYPES = {"sym_int4", "asym_int4", "nf4", "fp8_e4m3", "fp8_e5m2", "fp8", "fp6",
                         "woq_int4"}
B200_SUPPORTED_QTYPE_IDS = {ggml_tensor_qtype[n] for n in B200_SUPPORTED_QTYPES}
