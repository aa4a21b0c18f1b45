"""ctypes binding of libb200lowbit.so -- the B200 twin of the "GGML API" block of the
reference's ipex_llm/ggml/model/llama/llama_cpp.py:945-1193 (same Python function names and
argument order, so low_bit_linear.py reads like the reference's).

Library discovery follows llama_cpp.py:71-109: ``<package>/libs/libb200lowbit.so`` unless the
``LLAMA_CPP_LIB`` environment variable points elsewhere; a missing or unloadable library is a
``RuntimeError`` at import (there is no Python or CPU fallback for the kernels).
"""
import ctypes
import os
import pathlib

from ipex_llm_b200.utils.common import invalidInputError

_PKG_ROOT = pathlib.Path(__file__).resolve().parents[3]


def _load_shared_library():
    lib_path = _PKG_ROOT / "libs" / "libb200lowbit.so"
    if "LLAMA_CPP_LIB" in os.environ:
        lib_path = pathlib.Path(os.environ["LLAMA_CPP_LIB"]).resolve()
    if not lib_path.exists():
        invalidInputError(False, f"Shared library '{lib_path}' not found.",
                          "build it with `python -m ipex_llm_b200.build_native` "
                          "(nvcc, sm_100a); there is no CPU fallback.")
    try:
        return ctypes.CDLL(str(lib_path))
    except Exception as e:  # noqa: BLE001 - mirror llama_cpp.py:97-100
        invalidInputError(False, f"Failed to load shared library '{lib_path}': {e}.")


_lib = _load_shared_library()
LIB_PATH = _lib._name


class LowBitNativeError(RuntimeError):
    pass


def _check(rc=None, what=""):
    """Raise if the native call reported an error (return code or the per-thread channel)."""
    code = _lib.b200_lowbit_last_error() if rc is None else rc
    if code:
        msg = _lib.b200_lowbit_last_error_string().decode(errors="replace")
        _lib.b200_lowbit_clear_error()
        raise LowBitNativeError(f"{what}: native error {code}: {msg}")


# ---------------------------------------------------------------- reference-shaped (host pointers)
def ggml_quantize_tensor(src, dst, qtype, n, k, hist, scale_search):
    _lib.b200_lowbit_clear_error()
    r = _lib.ggml_quantize_tensor(src, dst, qtype, n, k, hist, scale_search)
    _check(what="ggml_quantize_tensor")
    return r


_lib.ggml_quantize_tensor.argtypes = [ctypes.POINTER(ctypes.c_float), ctypes.c_void_p, ctypes.c_int,
                                      ctypes.c_size_t, ctypes.c_int, ctypes.POINTER(ctypes.c_int64),
                                      ctypes.c_bool]
_lib.ggml_quantize_tensor.restype = ctypes.c_size_t


def ggml_type_size(qtype):
    return _lib.ggml_type_size(qtype)


_lib.ggml_type_size.argtypes = [ctypes.c_int]
_lib.ggml_type_size.restype = ctypes.c_size_t


def ggml_qk_size(qtype):
    return _lib.ggml_qk_size(qtype)


_lib.ggml_qk_size.argtypes = [ctypes.c_int]
_lib.ggml_qk_size.restype = ctypes.c_int


def ggml_dequantize_q4_0(src, dst, k):
    _lib.b200_lowbit_clear_error()
    _lib.ggml_dequantize_q4_0(src, dst, k)
    _check(what="ggml_dequantize_q4_0")


_lib.ggml_dequantize_q4_0.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t]
_lib.ggml_dequantize_q4_0.restype = None


def ggml_dequantize(src, dst, k, qtype):
    _lib.b200_lowbit_clear_error()
    _lib.ggml_dequantize(src, dst, k, qtype)
    _check(what="ggml_dequantize")


_lib.ggml_dequantize.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int]
_lib.ggml_dequantize.restype = None


def ggml_q_format_convet_cpu2xpu(src, dst, n, qtype, k=None):
    """Reference signature plus the row length ``k`` the tiled B200 layout needs."""
    _lib.b200_lowbit_clear_error()
    if k is not None:
        _lib.b200_lowbit_set_row_length(int(k))
    _lib.ggml_q_format_convet_cpu2xpu(src, dst, n, qtype)
    _check(what="ggml_q_format_convet_cpu2xpu")


def ggml_q_format_convet_xpu2cpu(src, dst, n, qtype, k=None):
    _lib.b200_lowbit_clear_error()
    if k is not None:
        _lib.b200_lowbit_set_row_length(int(k))
    _lib.ggml_q_format_convet_xpu2cpu(src, dst, n, qtype)
    _check(what="ggml_q_format_convet_xpu2cpu")


for _n in ("ggml_q_format_convet_cpu2xpu", "ggml_q_format_convet_xpu2cpu"):
    getattr(_lib, _n).argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int]
    getattr(_lib, _n).restype = None
_lib.b200_lowbit_set_row_length.argtypes = [ctypes.c_int]
_lib.b200_lowbit_set_row_length.restype = None
for _n in ("ggml_q_format_convert_cpu2cuda_2d", "ggml_q_format_convert_cuda2cpu_2d"):
    getattr(_lib, _n).argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_int64,
                                  ctypes.c_int]
    getattr(_lib, _n).restype = ctypes.c_int


def ggml_compute_forward_mul_mat_q_fp32(src_0_ne, src_0_data, src_0_qtype, src_1_ne, src_1_data, result):
    _lib.b200_lowbit_clear_error()
    _lib.ggml_compute_forward_mul_mat_q_fp32(src_0_ne, src_0_data, src_0_qtype, src_1_ne, src_1_data,
                                             result)
    _check(what="ggml_compute_forward_mul_mat_q_fp32")


_lib.ggml_compute_forward_mul_mat_q_fp32.argtypes = [ctypes.POINTER(ctypes.c_int64), ctypes.c_void_p,
                                                     ctypes.c_int, ctypes.POINTER(ctypes.c_int64),
                                                     ctypes.c_void_p, ctypes.c_void_p]
_lib.ggml_compute_forward_mul_mat_q_fp32.restype = None

_lib.b200_lowbit_last_error.restype = ctypes.c_int
_lib.b200_lowbit_last_error_string.restype = ctypes.c_char_p
_lib.b200_lowbit_clear_error.restype = None
_lib.b200_lowbit_last_launch_count.restype = ctypes.c_int
_lib.b200_lowbit_sm_count.restype = ctypes.c_int

# ---------------------------------------------------------------- device-pointer entry points
DT_F32, DT_F16, DT_BF16 = 0, 1, 2

_lib.ggml_quantize_tensor_cuda.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_size_t,
                                           ctypes.c_int, ctypes.c_void_p, ctypes.c_bool, ctypes.c_void_p]
_lib.ggml_quantize_tensor_cuda.restype = ctypes.c_int
_lib.ggml_dequantize_cuda.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int,
                                      ctypes.c_int, ctypes.c_void_p]
_lib.ggml_dequantize_cuda.restype = ctypes.c_int
_lib.lowbit_dequantize_device_layout_cuda.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64,
                                                      ctypes.c_int64, ctypes.c_int, ctypes.c_int,
                                                      ctypes.c_void_p]
_lib.lowbit_dequantize_device_layout_cuda.restype = ctypes.c_int
_lib.lowbit_relayout_cuda.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_int64,
                                      ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
_lib.lowbit_relayout_cuda.restype = ctypes.c_int
_lib.lowbit_matmul_workspace_size.argtypes = [ctypes.c_int64, ctypes.c_int64, ctypes.c_int64, ctypes.c_int]
_lib.lowbit_matmul_workspace_size.restype = ctypes.c_size_t
_lib.lowbit_matmul_cuda.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_int64,
                                    ctypes.c_void_p, ctypes.c_int, ctypes.c_int64,
                                    ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int,
                                    ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p]
_lib.lowbit_matmul_cuda.restype = ctypes.c_int


def ggml_quantize_tensor_cuda(d_src, d_dst, qtype, n, k, d_hist, scale_search, stream):
    _check(_lib.ggml_quantize_tensor_cuda(d_src, d_dst, qtype, n, k, d_hist, scale_search, stream),
           "ggml_quantize_tensor_cuda")


def ggml_dequantize_cuda(d_src, d_dst, k, qtype, out_dtype, stream):
    _check(_lib.ggml_dequantize_cuda(d_src, d_dst, k, qtype, out_dtype, stream), "ggml_dequantize_cuda")


def lowbit_dequantize_device_layout_cuda(d_src, d_dst, rows, k, qtype, out_dtype, stream):
    _check(_lib.lowbit_dequantize_device_layout_cuda(d_src, d_dst, rows, k, qtype, out_dtype, stream),
           "lowbit_dequantize_device_layout_cuda")


def lowbit_relayout_cuda(d_src, d_dst, rows, k, qtype, to_device, stream):
    _check(_lib.lowbit_relayout_cuda(d_src, d_dst, rows, k, qtype, int(to_device), stream),
           "lowbit_relayout_cuda")


def lowbit_matmul_workspace_size(n_rows, k, m, x_dtype):
    return _lib.lowbit_matmul_workspace_size(n_rows, k, m, x_dtype)


def lowbit_matmul_cuda(d_w, qtype, n_rows, k, d_x, x_dtype, m, d_y, y_dtype, d_bias, accumulate,
                       workspace, workspace_bytes, stream):
    _check(_lib.lowbit_matmul_cuda(d_w, qtype, n_rows, k, d_x, x_dtype, m, d_y, y_dtype, d_bias,
                                   int(accumulate), workspace, workspace_bytes, stream),
           "lowbit_matmul_cuda")


_lib.lowbit_dequantize_rows_cuda.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_int64, ctypes.c_void_p,
                                             ctypes.c_int, ctypes.c_int64, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
_lib.lowbit_dequantize_rows_cuda.restype = ctypes.c_int


def lowbit_dequantize_rows_cuda(d_w, qtype, rows, k, d_ids, ids_dtype, n_ids, d_out, out_dtype, stream):
    _check(_lib.lowbit_dequantize_rows_cuda(d_w, qtype, rows, k, d_ids, ids_dtype, n_ids, d_out, out_dtype, stream),
           "lowbit_dequantize_rows_cuda")


_I64P = ctypes.POINTER(ctypes.c_int64)
_lib.lowbit_kv_quantize_cuda.argtypes = [ctypes.c_void_p, ctypes.c_int, _I64P, ctypes.c_void_p, _I64P, ctypes.c_int64,
                                         ctypes.c_int64, ctypes.c_int64, ctypes.c_int64, ctypes.c_void_p]
_lib.lowbit_kv_quantize_cuda.restype = ctypes.c_int
_lib.lowbit_kv_dequantize_cuda.argtypes = [ctypes.c_void_p, _I64P, ctypes.c_void_p, ctypes.c_int, _I64P, ctypes.c_int64,
                                           ctypes.c_int64, ctypes.c_int64, ctypes.c_int64, ctypes.c_void_p]
_lib.lowbit_kv_dequantize_cuda.restype = ctypes.c_int
_lib.lowbit_sdp_fp8_supported.argtypes = [ctypes.c_int] * 4
_lib.lowbit_sdp_fp8_supported.restype = ctypes.c_int
_lib.lowbit_sdp_fp8_workspace_size.argtypes = [ctypes.c_int] * 6
_lib.lowbit_sdp_fp8_workspace_size.restype = ctypes.c_size_t
_lib.lowbit_sdp_fp8_cuda.argtypes = [ctypes.c_void_p, ctypes.c_int, _I64P, ctypes.c_void_p, _I64P, ctypes.c_void_p, _I64P,
                                     ctypes.c_void_p, ctypes.c_int, _I64P, ctypes.c_int, ctypes.c_float] + [ctypes.c_int] * 6 + \
                                    [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p]
_lib.lowbit_sdp_fp8_cuda.restype = ctypes.c_int


def _strides3(t):
    return (ctypes.c_int64 * 3)(*t.stride()[:3])


def lowbit_kv_quantize(src, dst, stream):
    """src float [B,H,S,D] -> dst uint8 view of the same shape (both with D contiguous)"""
    B, H, S, D = src.shape
    _check(_lib.lowbit_kv_quantize_cuda(src.data_ptr(), _DT_OF[src.dtype], _strides3(src), dst.data_ptr(), _strides3(dst),
                                        B, H, S, D, stream), "lowbit_kv_quantize_cuda")


def lowbit_kv_dequantize(src, dst, stream):
    B, H, S, D = src.shape
    _check(_lib.lowbit_kv_dequantize_cuda(src.data_ptr(), _strides3(src), dst.data_ptr(), _DT_OF[dst.dtype], _strides3(dst),
                                          B, H, S, D, stream), "lowbit_kv_dequantize_cuda")


def lowbit_sdp_fp8_supported(hq, hkv, s, d):
    return bool(_lib.lowbit_sdp_fp8_supported(hq, hkv, s, d))


def lowbit_sdp_fp8_workspace_size(b, hq, hkv, s, l, d):
    return _lib.lowbit_sdp_fp8_workspace_size(b, hq, hkv, s, l, d)


def lowbit_sdp_fp8(q, k, v, mask, causal, scale, out, workspace, stream):
    B, Hq, S, D = q.shape
    Hkv, L = k.shape[1], k.shape[2]
    _check(_lib.lowbit_sdp_fp8_cuda(q.data_ptr(), _DT_OF[q.dtype], _strides3(q), k.data_ptr(), _strides3(k), v.data_ptr(),
                                    _strides3(v), mask.data_ptr() if mask is not None else None,
                                    _DT_OF[mask.dtype] if mask is not None else 0,
                                    _strides3(mask) if mask is not None else None, int(bool(causal)), float(scale),
                                    B, Hq, Hkv, S, L, D, out.data_ptr(), _DT_OF[out.dtype], workspace.data_ptr(),
                                    workspace.numel() * workspace.element_size(), stream), "lowbit_sdp_fp8_cuda")


class _DtypeIds(dict):
    """str(torch dtype) -> DT_* id; a dtype OBJECT is looked up once through its name and remembered (str() of a dtype costs
    about as much as the rest of a binding's argument marshalling)"""

    def __missing__(self, key):
        val = dict.__getitem__(self, str(key))
        self[key] = val
        return val


_DT_OF = _DtypeIds({"torch.float32": 0, "torch.float16": 1, "torch.bfloat16": 2})

_lib.lowbit_mlp_workspace_size.argtypes = [ctypes.c_int64, ctypes.c_int64, ctypes.c_int64, ctypes.c_int]
_lib.lowbit_mlp_workspace_size.restype = ctypes.c_size_t
_lib.lowbit_mlp_cuda.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_int64,
                                 ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_void_p, ctypes.c_int,
                                 ctypes.c_int, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p]
_lib.lowbit_mlp_cuda.restype = ctypes.c_int


def lowbit_mlp_workspace_size(n_rows, k, m, x_dtype):
    return _lib.lowbit_mlp_workspace_size(n_rows, k, m, x_dtype)


def lowbit_mlp_cuda(d_w_gate, d_w_up, qtype, n_rows, k, d_x, x_dtype, m, d_y, y_dtype, act, workspace,
                    workspace_bytes, stream):
    _check(_lib.lowbit_mlp_cuda(d_w_gate, d_w_up, qtype, n_rows, k, d_x, x_dtype, m, d_y, y_dtype, act,
                                workspace, workspace_bytes, stream), "lowbit_mlp_cuda")


_lib.lowbit_moe_supported.argtypes = [ctypes.c_int, ctypes.c_int64, ctypes.c_int64, ctypes.c_int]
_lib.lowbit_moe_supported.restype = ctypes.c_int
_lib.lowbit_moe_workspace_size.argtypes = [ctypes.c_int64, ctypes.c_int64, ctypes.c_int, ctypes.c_int]
_lib.lowbit_moe_workspace_size.restype = ctypes.c_size_t
_lib.lowbit_moe_cuda.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int,
                                 ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int,
                                 ctypes.c_int64, ctypes.c_int64, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_int,
                                 ctypes.c_int, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p]
_lib.lowbit_moe_cuda.restype = ctypes.c_int


def lowbit_moe_supported(qtype, hidden, inter, top_k):
    return bool(_lib.lowbit_moe_supported(qtype, hidden, inter, top_k))


def lowbit_moe_workspace_size(hidden, inter, top_k, x_dtype):
    return _lib.lowbit_moe_workspace_size(hidden, inter, top_k, x_dtype)


def lowbit_moe_cuda(d_x, x_dtype, d_ids, ids_dtype, d_weight, weight_dtype, top_k, d_gates, d_ups, d_downs, n_experts,
                    hidden, inter, qtype, act, d_y, y_dtype, accumulate, workspace, workspace_bytes, stream):
    _check(_lib.lowbit_moe_cuda(d_x, x_dtype, d_ids, ids_dtype, d_weight, weight_dtype, top_k, d_gates, d_ups, d_downs,
                                n_experts, hidden, inter, qtype, act, d_y, y_dtype, int(accumulate), workspace,
                                workspace_bytes, stream), "lowbit_moe_cuda")


_lib.lowbit_set_programmatic_dependent_launch.argtypes = [ctypes.c_int]
_lib.lowbit_set_programmatic_dependent_launch.restype = None
_lib.lowbit_get_programmatic_dependent_launch.argtypes = []
_lib.lowbit_get_programmatic_dependent_launch.restype = ctypes.c_int


def lowbit_set_programmatic_dependent_launch(on):
    _lib.lowbit_set_programmatic_dependent_launch(int(bool(on)))


def lowbit_get_programmatic_dependent_launch():
    return bool(_lib.lowbit_get_programmatic_dependent_launch())


_lib.lowbit_hint_next_weights.argtypes = [ctypes.c_void_p, ctypes.c_size_t]
_lib.lowbit_hint_next_weights.restype = None


def lowbit_hint_next_weights(d_w, nbytes):
    _lib.lowbit_hint_next_weights(d_w, nbytes)


_lib.lowbit_matmul_f32_noalloc_cuda.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_int64,
                                                ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p, ctypes.c_int,
                                                ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
_lib.lowbit_matmul_f32_noalloc_cuda.restype = ctypes.c_int


def lowbit_matmul_f32_noalloc_cuda(d_w, qtype, n_rows, k, d_x, m, d_y, y_dtype, d_bias, accumulate, stream):
    """fp32-activation matmul that allocates nothing (scratch is a fixed array inside the library,
    m * k <= 262144); for hosts that cannot let the library or themselves call cudaMalloc."""
    _check(_lib.lowbit_matmul_f32_noalloc_cuda(d_w, qtype, n_rows, k, d_x, m, d_y, y_dtype, d_bias,
                                               int(accumulate), stream), "lowbit_matmul_f32_noalloc_cuda")


def last_launch_count():
    return _lib.b200_lowbit_last_launch_count()


def sm_count():
    return _lib.b200_lowbit_sm_count()


def llama_init_backend():
    _lib.llama_init_backend()


_lib.llama_init_backend.restype = None

_llama_initialized = False
if not _llama_initialized:
    llama_init_backend()
    _llama_initialized = True


_lib.lowbit_rotary_inplaced_cuda.argtypes = [ctypes.c_void_p, ctypes.POINTER(ctypes.c_int64), ctypes.c_int, ctypes.c_void_p,
                                             ctypes.POINTER(ctypes.c_int64), ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                             ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_void_p,
                                             ctypes.c_void_p, ctypes.POINTER(ctypes.c_int64), ctypes.c_void_p, ctypes.c_void_p,
                                             ctypes.c_int, ctypes.POINTER(ctypes.c_int64), ctypes.c_int, ctypes.c_void_p]
_lib.lowbit_rotary_inplaced_cuda.restype = ctypes.c_int


def lowbit_rotary_inplaced(q, k, rotary_dim, two, inv_freq, position_ids, cos, sin, cs_interleaved, stream):
    """q [B,Hq,S,D], k [B,Hkv,S,D] or None (views with D contiguous), rotated in place"""
    B, Hq, S, D = q.shape
    pos_strides = cs_strides = None
    if inv_freq is not None:
        pos_strides = (ctypes.c_int64 * 2)(position_ids.stride(0) if position_ids.shape[0] > 1 else 0, position_ids.stride(1))
    else:
        cs_strides = (ctypes.c_int64 * 2)(cos.stride(0) if cos.shape[0] > 1 else 0, cos.stride(1))
    _check(_lib.lowbit_rotary_inplaced_cuda(q.data_ptr(), _strides3(q), Hq, k.data_ptr() if k is not None else None,
                                            _strides3(k) if k is not None else None, k.shape[1] if k is not None else 0,
                                            _DT_OF[q.dtype], B, S, D, rotary_dim, int(bool(two)),
                                            inv_freq.data_ptr() if inv_freq is not None else None,
                                            position_ids.data_ptr() if inv_freq is not None else None, pos_strides,
                                            cos.data_ptr() if cos is not None else None, sin.data_ptr() if sin is not None else None,
                                            _DT_OF[cos.dtype] if cos is not None else 0, cs_strides, int(bool(cs_interleaved)),
                                            stream), "lowbit_rotary_inplaced_cuda")


# ---- peer-memory all-reduce (csrc/comm.cu)
_lib.lowbit_comm_max_bytes.argtypes = []
_lib.lowbit_comm_max_bytes.restype = ctypes.c_size_t
_lib.lowbit_comm_create.argtypes = [ctypes.c_size_t, ctypes.POINTER(ctypes.c_void_p), ctypes.c_void_p]
_lib.lowbit_comm_create.restype = ctypes.c_int
_lib.lowbit_comm_open.argtypes = [ctypes.c_void_p, ctypes.POINTER(ctypes.c_void_p)]
_lib.lowbit_comm_open.restype = ctypes.c_int
_lib.lowbit_comm_close.argtypes = [ctypes.c_void_p]
_lib.lowbit_comm_close.restype = ctypes.c_int
_lib.lowbit_comm_destroy.argtypes = [ctypes.c_void_p]
_lib.lowbit_comm_destroy.restype = ctypes.c_int
_lib.lowbit_comm_status.argtypes = [ctypes.c_void_p, ctypes.POINTER(ctypes.c_int)]
_lib.lowbit_comm_status.restype = ctypes.c_int
_lib.lowbit_allreduce_sum_cuda.argtypes = [ctypes.POINTER(ctypes.c_void_p), ctypes.c_int, ctypes.c_int, ctypes.c_size_t,
                                           ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_int, ctypes.c_void_p]
_lib.lowbit_allreduce_sum_cuda.restype = ctypes.c_int


def lowbit_comm_max_bytes():
    return _lib.lowbit_comm_max_bytes()


def lowbit_comm_create(max_bytes):
    """-> (device pointer, 64-byte IPC handle) of this rank's buffer on the current device"""
    ptr = ctypes.c_void_p()
    handle = ctypes.create_string_buffer(64)
    _check(_lib.lowbit_comm_create(max_bytes, ctypes.byref(ptr), handle), "lowbit_comm_create")
    return ptr.value, handle.raw


def lowbit_comm_open(handle):
    ptr = ctypes.c_void_p()
    _check(_lib.lowbit_comm_open(ctypes.create_string_buffer(handle, 64), ctypes.byref(ptr)), "lowbit_comm_open")
    return ptr.value


def lowbit_comm_close(ptr):
    _check(_lib.lowbit_comm_close(ptr), "lowbit_comm_close")


def lowbit_comm_destroy(ptr):
    _check(_lib.lowbit_comm_destroy(ptr), "lowbit_comm_destroy")


def lowbit_comm_status(ptr):
    v = ctypes.c_int()
    _check(_lib.lowbit_comm_status(ptr, ctypes.byref(v)), "lowbit_comm_status")
    return v.value


def lowbit_allreduce_sum(peer_ptrs, rank, max_bytes, src, dst, stream):
    arr = (ctypes.c_void_p * len(peer_ptrs))(*peer_ptrs)
    _check(_lib.lowbit_allreduce_sum_cuda(arr, rank, len(peer_ptrs), max_bytes, src.data_ptr(), dst.data_ptr(), src.numel(),
                                          _DT_OF[src.dtype], stream), "lowbit_allreduce_sum_cuda")


_lib.lowbit_matmul_allreduce_cuda.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_int64, ctypes.c_void_p, ctypes.c_int,
                                              ctypes.c_int64, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int,
                                              ctypes.POINTER(ctypes.c_void_p), ctypes.c_int, ctypes.c_int, ctypes.c_size_t,
                                              ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p]
_lib.lowbit_matmul_allreduce_cuda.restype = ctypes.c_int


def lowbit_matmul_allreduce(w, qtype, n_rows, k, x, y, bias, accumulate, peer_ptrs, rank, max_bytes, workspace, stream):
    arr = (ctypes.c_void_p * len(peer_ptrs))(*peer_ptrs)
    _check(_lib.lowbit_matmul_allreduce_cuda(w.data_ptr(), qtype, n_rows, k, x.data_ptr(), _DT_OF[x.dtype], x.shape[0],
                                             y.data_ptr(), _DT_OF[y.dtype], bias.data_ptr() if bias is not None else None,
                                             int(bool(accumulate)), arr, rank, len(peer_ptrs), max_bytes,
                                             workspace.data_ptr() if workspace is not None else None,
                                             workspace.numel() if workspace is not None else 0, stream), "lowbit_matmul_allreduce_cuda")


# ---- the small operators of a recorded decode step (csrc/decode_ops.cu, csrc/kvcache.cu)
_I32P = ctypes.c_void_p
_lib.lowbit_rms_norm_cuda.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_void_p, ctypes.c_int, ctypes.c_float,
                                      ctypes.c_int64, ctypes.c_int64, ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_void_p]
_lib.lowbit_rms_norm_cuda.restype = ctypes.c_int
_lib.lowbit_rope_kv_append_cuda.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_int64] + [ctypes.c_int] * 5 + \
                                           [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_int, _I32P, ctypes.c_int,
                                            ctypes.c_void_p, _I64P, ctypes.c_void_p, _I64P, ctypes.c_int, ctypes.c_void_p]
_lib.lowbit_rope_kv_append_cuda.restype = ctypes.c_int
_lib.lowbit_sdp_fp8_devlen_cuda.argtypes = [ctypes.c_void_p, ctypes.c_int, _I64P, ctypes.c_void_p, _I64P, ctypes.c_void_p, _I64P,
                                            ctypes.c_int, ctypes.c_float] + [ctypes.c_int] * 5 + [_I32P, ctypes.c_int, ctypes.c_int,
                                            ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p]
_lib.lowbit_sdp_fp8_devlen_cuda.restype = ctypes.c_int
_lib.lowbit_greedy_advance_cuda.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_int, ctypes.c_int, ctypes.c_void_p,
                                            ctypes.c_void_p, ctypes.c_int64, ctypes.c_int, _I32P, ctypes.c_void_p, ctypes.c_void_p]
_lib.lowbit_greedy_advance_cuda.restype = ctypes.c_int


def lowbit_rms_norm(x_2d, weight, eps, out, stream):
    """x_2d [rows, hidden] (last dim contiguous), weight [hidden] -> out [rows, hidden]"""
    rows, hidden = x_2d.shape
    _check(_lib.lowbit_rms_norm_cuda(x_2d.data_ptr(), _DT_OF[x_2d.dtype], x_2d.stride(0), weight.data_ptr(),
                                     _DT_OF[weight.dtype], float(eps), rows, hidden, out.data_ptr(), _DT_OF[out.dtype],
                                     out.stride(0), stream), "lowbit_rms_norm_cuda")


def lowbit_rope_kv_append(qkv, n_heads, n_kv_heads, head_dim, cos, sin, pos, pos_add, k_cache, v_cache, stream):
    """qkv [B, S, (Hq + 2 Hkv) * D]; cos / sin [positions, >= D/2] tables; pos: int32 device scalar or None;
    k_cache / v_cache uint8 [B, Hkv, capacity, D]"""
    B, S = qkv.shape[0], qkv.shape[1]
    _check(_lib.lowbit_rope_kv_append_cuda(qkv.data_ptr(), _DT_OF[qkv.dtype], qkv.stride(0), qkv.stride(1), B, S, n_heads,
                                           n_kv_heads, head_dim, cos.data_ptr(), sin.data_ptr(), _DT_OF[cos.dtype],
                                           cos.stride(0), cos.shape[0], pos.data_ptr() if pos is not None else None, int(pos_add),
                                           k_cache.data_ptr(), _strides3(k_cache), v_cache.data_ptr(), _strides3(v_cache),
                                           k_cache.shape[2], stream), "lowbit_rope_kv_append_cuda")


def lowbit_sdp_fp8_devlen(q, k, v, causal, scale, kv_len, len_add, out, workspace, stream):
    """q [B, Hq, S, D] view; k / v uint8 [B, Hkv, capacity, D]; kv_len: int32 device scalar (keys = kv_len + len_add)"""
    B, Hq, S, D = q.shape
    Hkv, L_max = k.shape[1], k.shape[2]
    _check(_lib.lowbit_sdp_fp8_devlen_cuda(q.data_ptr(), _DT_OF[q.dtype], _strides3(q), k.data_ptr(), _strides3(k), v.data_ptr(),
                                           _strides3(v), int(bool(causal)), float(scale), B, Hq, Hkv, S, L_max, kv_len.data_ptr(),
                                           int(len_add), D, out.data_ptr(), _DT_OF[out.dtype], workspace.data_ptr(),
                                           workspace.numel() * workspace.element_size(), stream), "lowbit_sdp_fp8_devlen_cuda")


def lowbit_greedy_advance(logits, next_ids, tokens, state, stream, suppress_ids=None):
    """logits [B, V]; next_ids int64 [B]; tokens int64 [B, capacity]; state int32 [8]; suppress_ids int64 [8] or None"""
    B, V = logits.shape
    _check(_lib.lowbit_greedy_advance_cuda(logits.data_ptr(), _DT_OF[logits.dtype], logits.stride(0), B, V, next_ids.data_ptr(),
                                           tokens.data_ptr(), tokens.stride(0), tokens.shape[1], state.data_ptr(),
                                           suppress_ids.data_ptr() if suppress_ids is not None else None, stream),
           "lowbit_greedy_advance_cuda")


_lib.lowbit_decode_attention_workspace_size.argtypes = [ctypes.c_int] * 5
_lib.lowbit_decode_attention_workspace_size.restype = ctypes.c_size_t
_lib.lowbit_decode_attention_cuda.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int64] + [ctypes.c_int] * 4 + \
                                             [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_int, _I32P,
                                              ctypes.c_void_p, _I64P, ctypes.c_void_p, _I64P, ctypes.c_int, ctypes.c_float,
                                              ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p]
_lib.lowbit_decode_attention_cuda.restype = ctypes.c_int


def lowbit_decode_attention_workspace_size(b, hq, hkv, capacity, d):
    return _lib.lowbit_decode_attention_workspace_size(b, hq, hkv, capacity, d)


def lowbit_decode_attention(qkv, n_heads, n_kv_heads, head_dim, cos, sin, cache_len, k_cache, v_cache, scale, out, workspace, stream,
                            cache_complete=False):
    """qkv [B, (Hq + 2 Hkv) * D] (or [B, 1, ...]); cos / sin [positions, >= D/2]; cache_len: int32 device scalar = tokens
    already cached; k_cache / v_cache uint8 [B, Hkv, capacity, D]; out [B, Hq, 1, D] contiguous"""
    B = qkv.shape[0]
    _check(_lib.lowbit_decode_attention_cuda(qkv.data_ptr(), _DT_OF[qkv.dtype], qkv.stride(0), B, n_heads, n_kv_heads, head_dim,
                                             cos.data_ptr(), sin.data_ptr(), _DT_OF[cos.dtype], cos.stride(0), cos.shape[0],
                                             cache_len.data_ptr(), k_cache.data_ptr(), _strides3(k_cache), v_cache.data_ptr(),
                                             _strides3(v_cache), k_cache.shape[2], float(scale), out.data_ptr(),
                                             _DT_OF[out.dtype], int(bool(cache_complete)), workspace.data_ptr(),
                                             workspace.numel() * workspace.element_size(), stream), "lowbit_decode_attention_cuda")


_lib.lowbit_sdp_fp8_prefill_supported.argtypes = [ctypes.c_int] * 4
_lib.lowbit_sdp_fp8_prefill_supported.restype = ctypes.c_int
_lib.lowbit_sdp_fp8_prefill_cuda.argtypes = [ctypes.c_void_p, ctypes.c_int, _I64P, ctypes.c_void_p, _I64P, ctypes.c_void_p, _I64P,
                                             ctypes.c_void_p, ctypes.c_int, _I64P, ctypes.c_int, ctypes.c_float] + [ctypes.c_int] * 6 + \
                                            [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
_lib.lowbit_sdp_fp8_prefill_cuda.restype = ctypes.c_int
_lib.lowbit_sdp_fp8_prefill_strided_cuda.argtypes = [ctypes.c_void_p, ctypes.c_int, _I64P, ctypes.c_void_p, _I64P, ctypes.c_void_p, _I64P,
                                                     ctypes.c_void_p, ctypes.c_int, _I64P, ctypes.c_int, ctypes.c_float] + [ctypes.c_int] * 6 + \
                                                    [ctypes.c_void_p, ctypes.c_int, _I64P, ctypes.c_void_p]
_lib.lowbit_sdp_fp8_prefill_strided_cuda.restype = ctypes.c_int


def lowbit_sdp_fp8_prefill_supported(hq, hkv, d, dtype):
    return str(dtype) in ("torch.float16", "torch.bfloat16") and bool(_lib.lowbit_sdp_fp8_prefill_supported(hq, hkv, d, _DT_OF[str(dtype)]))


def lowbit_sdp_fp8_prefill(q, k, v, mask, causal, scale, out, stream):
    """q [B,Hq,S,D] view, k / v uint8 [B,Hkv,L,D] views, mask additive [B,1|Hq,S,L] or None, out [B,Hq,S,D] with contiguous
    rows (any batch / head / token strides: a transposed view of a [B,S,Hq,D] buffer is written in place)"""
    B, Hq, S, D = q.shape
    Hkv, L = k.shape[1], k.shape[2]
    _check(_lib.lowbit_sdp_fp8_prefill_strided_cuda(q.data_ptr(), _DT_OF[q.dtype], _strides3(q), k.data_ptr(), _strides3(k), v.data_ptr(),
                                                    _strides3(v), mask.data_ptr() if mask is not None else None,
                                                    _DT_OF[mask.dtype] if mask is not None else 0,
                                                    _strides3(mask) if mask is not None else None, int(bool(causal)), float(scale),
                                                    B, Hq, Hkv, S, L, D, out.data_ptr(), _DT_OF[out.dtype], _strides3(out), stream),
           "lowbit_sdp_fp8_prefill_cuda")
