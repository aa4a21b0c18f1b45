"""CPU oracle for the low-bit linear hot path -- TEST INFRASTRUCTURE ONLY.

ctypes wrapper over ``oracle/liblowbit_oracle.so`` (built from ``lowbit_oracle.c`` by
``oracle/Makefile``).  Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s
``cpu_baseline`` / ``--impl reference`` legs may import this package; the product
package ``ipex_llm_b200`` never does.

Parity status: byte-level parity with the reference's closed binary is UNPINNED (the
binary and any golden vectors are absent from /root/reference).  sym_int4/asym_int4 are
pinned against gguf.quants and against vectors generated from the reference's own
in-repo Python statements (tests/golden/make_golden.py); nf4/fp8/fp6 are build-defined.
See lowbit_oracle.c header and DESIGN.md.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liblowbit_oracle.so")

QTYPES = {"sym_int4": 2, "asym_int4": 3, "nf4": 10, "fp8_e4m3": 15, "fp8_e5m2": 19,
          "fp6": 29, "woq_int4": 34}


def build(force=False):
    """Compile the oracle with the committed Makefile (gcc only, a second or two)."""
    src = os.path.join(_HERE, "lowbit_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = ctypes.CDLL(_SO)
        L.oracle_qk_size.argtypes = [ctypes.c_int]
        L.oracle_qk_size.restype = ctypes.c_int
        L.oracle_type_size.argtypes = [ctypes.c_int]
        L.oracle_type_size.restype = ctypes.c_size_t
        L.oracle_set_qk32.argtypes = [ctypes.c_int]
        L.oracle_quantize_tensor.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int,
                                             ctypes.c_size_t, ctypes.c_int, ctypes.c_void_p,
                                             ctypes.c_int]
        L.oracle_quantize_tensor.restype = ctypes.c_size_t
        L.oracle_dequantize.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t,
                                        ctypes.c_int]
        for name in ("oracle_mul_mat_q_fp32", "oracle_mul_mat_q_f64",
                     "oracle_mul_mat_q_fp32_modeA"):
            getattr(L, name).argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int,
                                         ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
            getattr(L, name).restype = None
        L.oracle_f32_to_f16.argtypes = [ctypes.c_float]
        L.oracle_f32_to_f16.restype = ctypes.c_uint16
        L.oracle_f16_to_f32.argtypes = [ctypes.c_uint16]
        L.oracle_f16_to_f32.restype = ctypes.c_float
        L.oracle_mini_encode.argtypes = [ctypes.c_int, ctypes.c_float]
        L.oracle_mini_encode.restype = ctypes.c_uint32
        L.oracle_mini_decode.argtypes = [ctypes.c_int, ctypes.c_uint32]
        L.oracle_mini_decode.restype = ctypes.c_float
        L.oracle_modeA_uses_vnni.restype = ctypes.c_int
        _lib = L
    return _lib


def _qt(qtype):
    return QTYPES[qtype] if isinstance(qtype, str) else int(qtype)


def set_qk32(on):
    lib().oracle_set_qk32(1 if on else 0)


def qk_size(qtype):
    return lib().oracle_qk_size(_qt(qtype))


def type_size(qtype):
    return lib().oracle_type_size(_qt(qtype))


def quantize(w, qtype, scale_search=False, return_hist=False):
    """fp32 ndarray [..., k] -> canonical packed uint8 ndarray (1-D)."""
    w = np.ascontiguousarray(w, dtype=np.float32)
    qt = _qt(qtype)
    qk, bs = qk_size(qt), type_size(qt)
    k = w.shape[-1]
    if k % qk:
        raise ValueError(f"last dim must be a multiple of {qk}")
    n = w.size
    dst = np.empty((n // qk) * bs, dtype=np.uint8)
    hist = np.zeros(16, dtype=np.int64)
    wrote = lib().oracle_quantize_tensor(w.ctypes.data, dst.ctypes.data, qt, n, k,
                                         hist.ctypes.data, int(bool(scale_search)))
    if wrote != dst.size:
        raise RuntimeError(f"oracle_quantize_tensor wrote {wrote}, expected {dst.size}")
    return (dst, hist) if return_hist else dst


def dequantize(packed, shape, qtype):
    packed = np.ascontiguousarray(packed, dtype=np.uint8)
    n = int(np.prod(shape))
    out = np.empty(n, dtype=np.float32)
    lib().oracle_dequantize(packed.ctypes.data, out.ctypes.data, n, _qt(qtype))
    return out.reshape(shape)


def _mm(fn, packed, shape, qtype, x, out_dtype):
    packed = np.ascontiguousarray(packed, dtype=np.uint8)
    x = np.ascontiguousarray(x, dtype=np.float32)
    N, K = shape
    M = x.shape[0]
    assert x.shape[1] == K
    ne0 = (ctypes.c_int64 * 2)(K, N)
    ne1 = (ctypes.c_int64 * 2)(K, M)
    y = np.empty((M, N), dtype=out_dtype)
    fn(ne0, packed.ctypes.data, _qt(qtype), ne1, x.ctypes.data, y.ctypes.data)
    return y


def matmul(packed, shape, qtype, x):
    """Mode B: x[M,K] @ dequant(W[N,K]).T, fp32 serial-k accumulate."""
    return _mm(lib().oracle_mul_mat_q_fp32, packed, shape, qtype, x, np.float32)


def matmul_f64(packed, shape, qtype, x):
    return _mm(lib().oracle_mul_mat_q_f64, packed, shape, qtype, x, np.float64)


def matmul_mode_a(packed, shape, qtype, x):
    """Mode A: int8-quantised activations (ggml-faithful CPU arithmetic); CPU baseline."""
    return _mm(lib().oracle_mul_mat_q_fp32_modeA, packed, shape, qtype, x, np.float32)


def uses_vnni():
    return bool(lib().oracle_modeA_uses_vnni())
