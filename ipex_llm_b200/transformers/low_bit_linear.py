"""B200 host side of the low-bit linear hot path -- mirrors the public surface of the
reference's ipex_llm/transformers/low_bit_linear.py (LBL below) so that callers and tests read
the same: ``ggml_convert_qtype``, ``ggml_q_format_convet_cpu2xpu/xpu2cpu``,
``ggml_int4_convert_fp32``, ``ggml_convert_fp32``, ``ggml_matmul_src1_x_src0_t``,
``reshape_lm_head_input``, ``use_batch_forward``, ``FP4Params`` and ``LowBitLinear``.

Differences that are the point of this build:
  * the native library is libb200lowbit.so (hand-written sm_100a CUDA); the accelerator device
    type is "cuda" where the reference says "xpu";
  * there is no CPU arithmetic anywhere: tensors that live on the CPU are processed by the same
    CUDA kernels through the reference-shaped host-pointer C ABI;
  * the device layout is tiled by 16 rows, so the relayout helpers take the row length;
  * a quantised ``FP4Params`` ignores dtype casts (LBL only guards torch_fp8, :443-444): the
    nf4 buffer is bf16-typed (LBL:139-145) and ``model.to(torch.float16)`` must not touch it.
"""
import ctypes
import functools
import os
from functools import reduce
from operator import mul

import torch
import torch.nn.functional as F
from torch import nn

import ipex_llm_b200.ggml.model.llama.llama_cpp as ggml
from ipex_llm_b200.ggml.quantize import ggml_tensor_qtype, B200_SUPPORTED_QTYPE_IDS
from ipex_llm_b200.utils.common import invalidInputError

TORCH_LINEAR_THRESHOLD = int(os.getenv("BIGDL_LLM_LINEAR_THRESHOLD", "512"))
SYM_INT4 = ggml_tensor_qtype["sym_int4"]
ASYM_INT4 = ggml_tensor_qtype["asym_int4"]
NF4 = ggml_tensor_qtype["nf4"]
FP8E4 = ggml_tensor_qtype["fp8_e4m3"]
FP8E5 = ggml_tensor_qtype["fp8_e5m2"]
FP6 = ggml_tensor_qtype["fp6"]
WOQ_INT4 = ggml_tensor_qtype["woq_int4"]

_TORCH_TO_DT = {torch.float32: ggml.DT_F32, torch.float16: ggml.DT_F16, torch.bfloat16: ggml.DT_BF16}


def _check_qtype(qtype):
    invalidInputError(qtype in B200_SUPPORTED_QTYPE_IDS,
                      f"qtype {qtype} is not implemented by the B200 low-bit kernels "
                      f"(supported: sym_int4, asym_int4, nf4, fp8_e4m3, fp8_e5m2/fp8, fp6, woq_int4)")


def _stream_ptr(device):
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def _raw_stream(device):
    """the current stream's handle as an int (the wrappers below run a few hundred times per prefill: every microsecond of
    host time between two GEMMs is a microsecond the GPU may idle)"""
    return torch._C._cuda_getCurrentRawStream(device.index if device.index is not None else torch.cuda.current_device())


class _on_device:
    """``with torch.cuda.device(dev)`` only when dev is not the current device already"""
    __slots__ = ("ctx",)

    def __init__(self, device):
        idx = device.index
        self.ctx = None if (idx is None or idx == torch.cuda.current_device()) else torch.cuda.device(device)

    def __enter__(self):
        if self.ctx is not None:
            self.ctx.__enter__()

    def __exit__(self, *exc):
        if self.ctx is not None:
            self.ctx.__exit__(*exc)
        return False


def get_block_size(qtype: str):
    """LBL:98-99"""
    return ggml.ggml_qk_size(ggml_tensor_qtype[qtype])


def get_qk_size(qtype: int):
    """LBL:102-103"""
    return ggml.ggml_qk_size(qtype)


def ggml_convert_qtype(tensor: torch.Tensor, qtype: int, device=None, convert_shape_only=False,
                       imatrix: torch.Tensor = None, in_features: int = None,
                       enable_scale_search: bool = False):
    """fp32 tensor [.., k] -> packed canonical blocks (LBL:106-195).

    Buffer sizing and dtype rules are the reference's: ``(n // QK) * type_size`` uint8, held as a
    bf16 tensor of half that length for nf4 (LBL:139-145).  ``device`` is where the result
    lives: "meta" / convert_shape_only only size the buffer (LBL:150); a CUDA ``tensor`` is
    quantised in place on its device, a CPU one through the host-pointer ABI.
    """
    _check_qtype(qtype)
    invalidInputError(imatrix is None, "imatrix quantisation (k-quants / iq) is out of scope of this build")
    QK = ggml.ggml_qk_size(qtype)
    block_size_in_bytes = ggml.ggml_type_size(qtype)
    invalidInputError(tensor.dtype == torch.float, "Input tensor must be float32")
    n = tensor.numel()
    k = tensor.shape[-1]
    invalidInputError(k % QK == 0, f"Last dim of input tensor must be multiple of {QK}")
    dst_size = (n // QK) * block_size_in_bytes
    if device is None:
        device = tensor.device
    device = torch.device(device) if not isinstance(device, torch.device) else device
    if device.type == "cuda" and device.index is None:
        device = torch.device("cuda", torch.cuda.current_device())
    if qtype == NF4:
        dst_tensor = torch.empty(dst_size // 2, dtype=torch.bfloat16, device=device)
    else:
        dst_tensor = torch.empty(dst_size, dtype=torch.uint8, device=device)
    if convert_shape_only or device.type == "meta":
        return dst_tensor
    if n == 0:
        return dst_tensor
    if device.type == "cuda":
        src = tensor.contiguous().to(device)
        with torch.cuda.device(device):
            ggml.ggml_quantize_tensor_cuda(ctypes.c_void_p(src.data_ptr()), ctypes.c_void_p(dst_tensor.data_ptr()),
                                           qtype, n, k, None, enable_scale_search, _stream_ptr(device))
        return dst_tensor
    invalidInputError(device.type == "cpu", f"unsupported device {device}")
    tensor = tensor.contiguous().cpu()
    src = ctypes.cast(tensor.data_ptr(), ctypes.POINTER(ctypes.c_float))
    dst = ctypes.c_void_p(dst_tensor.data_ptr())
    hist = (ctypes.c_int64 * 16)()
    ggml.ggml_quantize_tensor(src, dst, qtype, n, k, hist, enable_scale_search)
    return dst_tensor


def _check_packed(tensor, qtype):
    if qtype == NF4:
        invalidInputError(tensor.dtype == torch.bfloat16, "NF4 Input tensor must be bfloat16")
    else:
        invalidInputError(tensor.dtype == torch.uint8, "Input tensor except NF4 must be uint8")


def _relayout(tensor, shape, qtype, to_device_layout):
    """canonical <-> tiled on whatever device ``tensor`` is (CPU: via the host-pointer ABI)."""
    _check_qtype(qtype)
    _check_packed(tensor, qtype)
    rows = reduce(mul, shape[:-1], 1)
    k = shape[-1]
    tensor = tensor.contiguous()
    dst_tensor = torch.empty_like(tensor)
    if tensor.numel() == 0:
        return dst_tensor
    if tensor.device.type == "cuda":
        with torch.cuda.device(tensor.device):
            ggml.lowbit_relayout_cuda(ctypes.c_void_p(tensor.data_ptr()), ctypes.c_void_p(dst_tensor.data_ptr()),
                                      rows, k, qtype, to_device_layout, _stream_ptr(tensor.device))
    else:
        invalidInputError(tensor.device.type == "cpu", "Input tensor must be on cpu or cuda")
        fn = ggml.ggml_q_format_convet_cpu2xpu if to_device_layout else ggml.ggml_q_format_convet_xpu2cpu
        fn(ctypes.c_void_p(tensor.data_ptr()), ctypes.c_void_p(dst_tensor.data_ptr()), rows * k, qtype, k=k)
    return dst_tensor


def ggml_q_format_convet_cpu2xpu(tensor: torch.Tensor, num_elem: int, qtype: int, shape=None):
    """LBL:198-224.  ``shape`` = (rows, k) of the weight; required because the B200 layout is
    tiled by rows (the reference's XPU layout is not, so it only passes num_elem)."""
    invalidInputError(shape is not None and reduce(mul, shape, 1) == num_elem,
                      "the B200 device layout needs the weight shape: pass shape=(rows, k)")
    return _relayout(tensor, tuple(shape), qtype, True)


def ggml_q_format_convet_xpu2cpu(tensor: torch.Tensor, num_elem: int, qtype: int, shape=None):
    """LBL:227-253."""
    invalidInputError(shape is not None and reduce(mul, shape, 1) == num_elem,
                      "the B200 device layout needs the weight shape: pass shape=(rows, k)")
    return _relayout(tensor, tuple(shape), qtype, False)


ggml_q_format_convert_cpu2cuda = ggml_q_format_convet_cpu2xpu
ggml_q_format_convert_cuda2cpu = ggml_q_format_convet_xpu2cpu


def ggml_convert_fp32(tensor: torch.Tensor, weight_shape: tuple, k: int, qtype: int):
    """packed CANONICAL blocks -> fp32 [weight_shape] (LBL:269-279); k = element count."""
    _check_qtype(qtype)
    if qtype == NF4 and tensor.dtype == torch.bfloat16:
        tensor = tensor.view(torch.uint8)
    invalidInputError(tensor.dtype == torch.uint8, "Input tensor must be uint8")
    tensor = tensor.contiguous()
    dst_tensor = torch.empty(weight_shape, dtype=torch.float, device=tensor.device)
    if k == 0:
        return dst_tensor
    if tensor.device.type == "cuda":
        with torch.cuda.device(tensor.device):
            ggml.ggml_dequantize_cuda(ctypes.c_void_p(tensor.data_ptr()), ctypes.c_void_p(dst_tensor.data_ptr()),
                                      k, qtype, ggml.DT_F32, _stream_ptr(tensor.device))
    else:
        ggml.ggml_dequantize(ctypes.c_void_p(tensor.data_ptr()), ctypes.c_void_p(dst_tensor.data_ptr()), k, qtype)
    return dst_tensor


def ggml_int4_convert_fp32(tensor: torch.Tensor, weight_shape: tuple, k: int):
    """LBL:256-266 (sym_int4 only)."""
    return ggml_convert_fp32(tensor, weight_shape, k, SYM_INT4)


def dequantize_device_layout(tensor: torch.Tensor, weight_shape: tuple, qtype: int, dtype=torch.float32):
    """weight in the tiled device layout (a CUDA FP4Params.data) -> dense [rows, k]."""
    if tensor.dtype == torch.bfloat16:
        tensor = tensor.view(torch.uint8)
    rows, k = reduce(mul, weight_shape[:-1], 1), weight_shape[-1]
    out = torch.empty(weight_shape, dtype=dtype, device=tensor.device)
    with torch.cuda.device(tensor.device):
        ggml.lowbit_dequantize_device_layout_cuda(ctypes.c_void_p(tensor.data_ptr()), ctypes.c_void_p(out.data_ptr()),
                                                  rows, k, qtype, _TORCH_TO_DT[dtype], _stream_ptr(tensor.device))
    return out


def reshape_lm_head_input(x):
    """LBL:282-289: in prefill only the last position needs logits."""
    if x.dim() > 3:
        x = x.reshape([-1, x.shape[-2], x.shape[-1]])
    shape = list(x.size())
    if shape[1] > 10:
        shape[1] = 1
        x = x[:, -1, :].view(shape)
    return x


GEMV_MAX_M = 32   # rows of x the HBM-bound GEMV kernel takes in one pass (gemv_args.h)
SKINNY_MIN_M, SKINNY_MAX_M = 17, 64   # half-precision batches in this window take the skinny tensor-core kernel (capi.cu use_skinny)


def takes_tensor_core_path(M: int, N: int, K: int, x_dtype) -> bool:
    """whether lowbit_matmul runs this shape on the tcgen05 kernels (weights rounded once to the activation dtype:
    the prefill precision contract) rather than on the GEMV (exact up to fp32 summation order)"""
    if x_dtype == torch.float32 or N % 128 or K % 64 or N < 128:
        return False
    return M > GEMV_MAX_M or (SKINNY_MIN_M <= M <= SKINNY_MAX_M and K % 128 == 0)


def use_batch_forward(x: torch.Tensor, qtype: int, output_len: int):
    """LBL:292-327 picks the reference's small-batch kernel; here: does the dequant-GEMV (not the
    tensor-core GEMM) serve this activation?  Same intent, B200 crossover."""
    return x.shape[0] <= GEMV_MAX_M and qtype in B200_SUPPORTED_QTYPE_IDS


# ------------------------------------------------------------------ workspace (per device)
_workspaces = {}


def _get_workspace(device, need):
    ws = _workspaces.get(device)
    if ws is None or ws.numel() < need:
        ws = torch.zeros(int(need * 1.25) + 256, dtype=torch.uint8, device=device)
        _workspaces[device] = ws
    return ws


def lowbit_matmul(x_2d: torch.Tensor, weight: torch.Tensor, weight_shape, qtype: int, out: torch.Tensor = None,
                  bias: torch.Tensor = None, accumulate: bool = False, out_dtype=None, next_weight: torch.Tensor = None):
    """x_2d [M,K] (fp32/fp16/bf16, CUDA) @ dequant(W [N,K], tiled device layout).T -> [M,N].

    The B200 twin of ``xe_linear.forward_new(x_2d, w, qtype, out_len)`` (LBL:704-708); ``out`` /
    ``accumulate`` / ``bias`` expose the fused epilogue of the kernel.
    """
    N, K = weight_shape
    M = x_2d.shape[0]
    x_dt = _TORCH_TO_DT.get(x_2d.dtype)
    if x_dt is None:
        invalidInputError(False, f"unsupported activation dtype {x_2d.dtype}")
    if x_2d.shape[1] != K:
        invalidInputError(False, f"activation last dim {x_2d.shape[1]} != in_features {K}")
    if not x_2d.is_contiguous():
        x_2d = x_2d.contiguous()
    out_dtype = out_dtype or (out.dtype if out is not None else x_2d.dtype)
    if out is None:
        invalidInputError(not accumulate, "accumulate needs an output tensor")
        out = torch.empty((M, N), dtype=out_dtype, device=x_2d.device)
    elif not (out.is_contiguous() and out.shape == (M, N) and out.dtype == out_dtype):
        invalidInputError(False, "out must be a contiguous [M, N] tensor of out_dtype")
    if bias is not None and bias.dtype != out_dtype:
        bias = bias.to(out_dtype)
    if M == 0:
        return out
    w = weight.view(torch.uint8) if weight.dtype != torch.uint8 else weight
    dev = x_2d.device
    ws = _get_workspace(dev, _matmul_workspace_size(N, K, M, x_dt))
    if next_weight is not None:      # L2 prefetch hint: the weights the following matmul on this stream reads
        ggml.lowbit_hint_next_weights(ctypes.c_void_p(next_weight.data_ptr()), next_weight.numel() * next_weight.element_size())
    with _on_device(dev):
        ggml.lowbit_matmul_cuda(w.data_ptr(), qtype, N, K, x_2d.data_ptr(), x_dt, M, out.data_ptr(), _TORCH_TO_DT[out_dtype],
                                bias.data_ptr() if bias is not None else None, accumulate, ws.data_ptr(), ws.numel(), _raw_stream(dev))
    return out


@functools.lru_cache(maxsize=4096)
def _matmul_workspace_size(N, K, M, x_dt):
    return ggml.lowbit_matmul_workspace_size(N, K, M, x_dt)


SILU = 0     # activation ids of the reference (transformers/models/utils.py: SILU / GELU)
GELU = 1


def mlp_forward(x_2d: torch.Tensor, gate_weight: torch.Tensor, up_weight: torch.Tensor, M: int, K: int, N: int,
                act: int, qtype: int, out: torch.Tensor = None):
    """``xe_linear.mlp_forward_xpu(x_2d, gate_w, up_w, M, K, N, act, qtype)`` (models/common.py:146-161):
    act(x . Wgate^T) * (x . Wup^T) -> [M, N] in the activation dtype.  Both weights are [N, K]
    in the tiled device layout; at decode sizes (M <= 8) one kernel streams both matrices."""
    invalidInputError(x_2d.dtype in _TORCH_TO_DT, f"unsupported activation dtype {x_2d.dtype}")
    invalidInputError(tuple(x_2d.shape) == (M, K), f"activation shape {tuple(x_2d.shape)} != ({M}, {K})")
    invalidInputError(act in (SILU, GELU), f"unknown activation id {act}")
    invalidInputError(gate_weight.numel() * gate_weight.element_size() == up_weight.numel() * up_weight.element_size(),
                      "gate and up weights must have the same packed size")
    x_2d = x_2d.contiguous()
    if out is None:
        out = torch.empty((M, N), dtype=x_2d.dtype, device=x_2d.device)
    if M == 0:
        return out
    dev = x_2d.device
    dt = _TORCH_TO_DT[x_2d.dtype]
    ws = _get_workspace(dev, ggml.lowbit_mlp_workspace_size(N, K, M, dt))
    g = gate_weight.view(torch.uint8) if gate_weight.dtype != torch.uint8 else gate_weight
    u = up_weight.view(torch.uint8) if up_weight.dtype != torch.uint8 else up_weight
    with torch.cuda.device(dev):
        ggml.lowbit_mlp_cuda(ctypes.c_void_p(g.data_ptr()), ctypes.c_void_p(u.data_ptr()), qtype, N, K,
                             ctypes.c_void_p(x_2d.data_ptr()), dt, M, ctypes.c_void_p(out.data_ptr()),
                             _TORCH_TO_DT[out.dtype], act, ctypes.c_void_p(ws.data_ptr()), ws.numel(), _stream_ptr(dev))
    return out


def use_fuse_moe(hidden: int, inter: int, top_k: int, qtype: int) -> bool:
    """shape / qtype test of the fused one-token MoE (the reference's use_fuse_moe, models/utils.py)"""
    return ggml.lowbit_moe_supported(qtype, hidden, inter, top_k)


def moe_forward_vec(x: torch.Tensor, topk_ids: torch.Tensor, topk_weight: torch.Tensor, gates: torch.Tensor,
                    ups: torch.Tensor, downs: torch.Tensor, hidden: int, inter: int, qtype: int, act: int = SILU,
                    out: torch.Tensor = None, accumulate: bool = False):
    """``xe_linear.moe_forward_vec`` (models/qwen3_moe.py:88-105, deepseek.py:296-312): the routed
    experts' MLPs of ONE token.  ``gates`` / ``ups`` / ``downs`` are device tensors of the experts'
    weight addresses (``expert.gate_proj.weight.data_ptr()`` ...), ``topk_ids`` / ``topk_weight``
    the router's [1, top_k] outputs, read on the device.  Returns [1, hidden] in x's dtype."""
    invalidInputError(x.dtype in _TORCH_TO_DT, f"unsupported activation dtype {x.dtype}")
    invalidInputError(x.numel() == hidden, "moe_forward_vec handles one token")
    invalidInputError(topk_ids.dtype in (torch.int32, torch.int64), "topk_ids must be int32 or int64")
    invalidInputError(topk_weight.dtype in _TORCH_TO_DT, f"unsupported routing weight dtype {topk_weight.dtype}")
    invalidInputError(topk_ids.numel() == topk_weight.numel(), "topk_ids and topk_weight differ in length")
    invalidInputError(gates.numel() == ups.numel() == downs.numel() and gates.element_size() == 8,
                      "gates / ups / downs must be equally long tensors of 64-bit addresses")
    top_k = topk_ids.numel()
    x = x.contiguous()
    topk_ids, topk_weight = topk_ids.contiguous(), topk_weight.contiguous()
    if out is None:
        invalidInputError(not accumulate, "accumulate needs an output tensor")
        out = torch.empty((1, hidden), dtype=x.dtype, device=x.device)
    dev = x.device
    dt = _TORCH_TO_DT[x.dtype]
    ws = _get_workspace(dev, ggml.lowbit_moe_workspace_size(hidden, inter, top_k, dt))
    with torch.cuda.device(dev):
        ggml.lowbit_moe_cuda(ctypes.c_void_p(x.data_ptr()), dt, ctypes.c_void_p(topk_ids.data_ptr()),
                             1 if topk_ids.dtype == torch.int64 else 0, ctypes.c_void_p(topk_weight.data_ptr()),
                             _TORCH_TO_DT[topk_weight.dtype], top_k, ctypes.c_void_p(gates.data_ptr()),
                             ctypes.c_void_p(ups.data_ptr()), ctypes.c_void_p(downs.data_ptr()), gates.numel(),
                             hidden, inter, qtype, act, ctypes.c_void_p(out.data_ptr()), _TORCH_TO_DT[out.dtype],
                             accumulate, ctypes.c_void_p(ws.data_ptr()), ws.numel(), _stream_ptr(dev))
    return out


def ggml_matmul_src1_x_src0_t(src0: torch.Tensor, src1: torch.Tensor, src0_shape: torch.Size, src0_qtype: int):
    """LBL:494-529: result[M,N] = src1[M,K] . W[N,K]^T with W = canonical packed blocks on the
    CPU; fp32 in, fp32 out, one call across the reference C ABI (which runs on the GPU here)."""
    if src1.dtype != torch.float32:
        src1 = src1.float()
    src1 = src1.contiguous()
    if src0.dtype == torch.bfloat16:
        src0 = src0.view(torch.uint8)
    result_shape = (src1.shape[0], src0_shape[0])
    result_t = torch.empty(result_shape, dtype=torch.float32)
    src0_ne = (ctypes.c_int64 * 2)(*tuple(reversed(src0_shape)))
    src1_ne = (ctypes.c_int64 * 2)(*tuple(reversed(src1.shape)))
    ggml.ggml_compute_forward_mul_mat_q_fp32(src_0_ne=src0_ne, src_0_data=ctypes.c_void_p(src0.data_ptr()),
                                             src_0_qtype=src0_qtype, src_1_ne=src1_ne,
                                             src_1_data=ctypes.c_void_p(src1.data_ptr()),
                                             result=ctypes.c_void_p(result_t.data_ptr()))
    return result_t


class MatMulLowBit(torch.autograd.Function):
    """LBL:532-577, the autograd wrapper of the device matmul (QLoRA-style training freezes the packed
    weight: only the activation gets a gradient).  forward = the dequant-GEMV / GEMM kernels,
    backward = grad_output @ dequant(W) with W dequantised by the tiled-layout kernel."""

    @staticmethod
    def forward(ctx, A, weight, output_size):
        shape = tuple(weight._shape)
        invalidInputError(shape[0] == output_size, "output_size does not match the weight")
        A_2d = A.reshape(-1, A.shape[-1])
        result = lowbit_matmul(A_2d, weight.data, shape, weight.qtype).view(A.shape[:-1] + (output_size,))
        ctx.tensors = (A, weight) if any(ctx.needs_input_grad[:2]) else (None, None)
        return result

    @staticmethod
    def backward(ctx, grad_output):
        A, weight = ctx.tensors
        grad_A = None
        if ctx.needs_input_grad[0]:
            dequant_weight = dequantize_device_layout(weight.data, tuple(weight._shape), weight.qtype, dtype=grad_output.dtype)
            grad_A = torch.matmul(grad_output, dequant_weight)
        return grad_A, None, None


class MatMulLowBitCPU(torch.autograd.Function):
    """LBL:580-602: the host-pointer variant (canonical packed weight in CPU memory); this is the statement
    of the "mode B" semantics the kernels are checked against: A @ dequant(W).T"""

    @staticmethod
    def forward(ctx, A, weight):
        x0_fp32 = ggml_convert_fp32(weight.data, weight._shape, weight._shape[0] * weight._shape[1], weight.qtype)
        result = torch.matmul(A, x0_fp32.T.to(A.dtype))
        ctx.tensors = (A, weight) if any(ctx.needs_input_grad[:2]) else (None, None)
        return result

    @staticmethod
    def backward(ctx, grad_output):
        A, weight = ctx.tensors
        grad_A = None
        if ctx.needs_input_grad[0]:
            x0_fp32 = ggml_convert_fp32(weight.data, weight._shape, weight._shape[0] * weight._shape[1], weight.qtype)
            grad_A = torch.matmul(grad_output, x0_fp32.to(grad_output.dtype))
        return grad_A, None, None


class FP4Params(torch.nn.Parameter):
    """LBL:332-491.  Holds the packed weight; quantises on the first ``.to()``.

    Layout rule: data on "cpu"/"meta" is canonical (the on-disk format, LBL §save_low_bit);
    data on "cuda" is the tiled device layout.
    """

    def __new__(cls, data=None, requires_grad=False, quantized=False, _shape=None, convert_shape_only=False,
                qtype=None, imatrix=None, in_features=None, enable_scale_search=False, torch_fp8_scale=None):
        if data is None:
            data = torch.empty(0)
        self = torch.Tensor._make_subclass(cls, data, requires_grad)
        self.data = data
        self.quantized = quantized
        self._shape = _shape
        self.qtype = qtype
        self.convert_shape_only = convert_shape_only
        self.imatrix = imatrix
        self.in_features = in_features
        self.enable_scale_search = enable_scale_search
        self.torch_fp8_scale = torch_fp8_scale
        return self

    def quantize(self, device=None):
        """LBL:370-423 (mixed_fp4/mixed_fp8 are out of scope).  ``device``: "cpu" / "meta" /
        a cuda device -- where the canonical packed result is produced."""
        if not self.quantized:
            w = self.data.contiguous()
            shape = w.shape
            dev = torch.device(device) if device is not None else w.device
            if dev.type == "cuda":
                w = w.to(dev)
            w = w.float()
            self.data = ggml_convert_qtype(w, self.qtype, device=dev,
                                           convert_shape_only=self.convert_shape_only, imatrix=self.imatrix,
                                           in_features=self.in_features,
                                           enable_scale_search=self.enable_scale_search)
            self.quantized = True
            self._shape = shape
        return self

    def get_shape(self):
        return self._shape

    def _like(self, data):
        return FP4Params(data, requires_grad=self.requires_grad, quantized=self.quantized, _shape=self._shape,
                         qtype=self.qtype, enable_scale_search=self.enable_scale_search,
                         convert_shape_only=self.convert_shape_only)

    def to(self, *args, **kwargs):
        device, dtype, non_blocking, convert_to_format = torch._C._nn._parse_to(*args, **kwargs)
        cur = self.data.device
        if device is None:
            if self.quantized:
                return self            # dtype-only casts never touch packed data
            return self._like(self.data.to(dtype=dtype, non_blocking=non_blocking)) if dtype else self
        if device.type == "cpu" and cur.type == "cpu":
            return self.quantize("cpu")
        if device.type == "meta" and cur.type == "meta":
            return self.quantize("meta")
        if device.type == "cuda" and cur.type in ("cpu", "cuda") and not self.quantized:
            # quantise on the target GPU (fp weights go up once), keep self canonical on its device
            if device.index is None:
                device = torch.device("cuda", torch.cuda.current_device())
            self.quantize(device)                       # canonical packed, on the GPU
            canonical = self.data
            tiled = _relayout(canonical, tuple(self._shape), self.qtype, True)
            if cur.type == "cpu":
                self.data = canonical.cpu()
            return self._like(tiled)
        if device.type == "cuda" and cur.type == "cpu":
            if device.index is None:
                device = torch.device("cuda", torch.cuda.current_device())
            canonical = self.data.to(device, non_blocking=non_blocking)
            return self._like(_relayout(canonical, tuple(self._shape), self.qtype, True))
        if device.type == "cpu" and cur.type == "cuda":
            canonical = _relayout(self.data, tuple(self._shape), self.qtype, False)
            return self._like(canonical.cpu())
        # cuda -> cuda (possibly another device), meta targets, ...: layout does not change
        return self._like(self.data.to(device=device, non_blocking=non_blocking))


class LowBitLinear(nn.Linear):
    """LBL:605-755."""

    def __init__(self, input_features, output_features, qtype, bias=True, conver_to_half=True, mp_group=None,
                 optimize_lm_head=False, act_order=False, enable_scale_search=False):
        super().__init__(input_features, output_features, bias)
        _check_qtype(qtype)
        self.weight = FP4Params(self.weight.data, requires_grad=False, quantized=False, _shape=None, qtype=qtype,
                                enable_scale_search=enable_scale_search)
        self.in_len = input_features
        self.out_len = output_features
        self.weight_shape = (self.out_len, self.in_len)
        self.weight_length = self.out_len * self.in_len
        self.qtype = qtype
        self.conver_to_half = conver_to_half
        self.mp_group = mp_group
        self.compute_dtype = None
        self.optimize_lm_head = optimize_lm_head
        self.device = None
        self.is_lm_head = self.in_len * self.out_len >= 32000 * 4096 and self.bias is None
        self.low_memory_mode = self.is_lm_head
        self.act_order = act_order
        self.next_linear = None     # optional: the LowBitLinear that runs next (L2 prefetch hint, decode only)
        if act_order:
            self.register_buffer("g_idx_map", torch.arange(self.in_len, dtype=torch.int64))

    def _all_reduce(self, result):
        """LBL:715-722 / :748-752: sum the partial outputs of a row-parallel (K-split) linear,
        after the matmul and before the bias."""
        g = self.mp_group
        if hasattr(g, "all_reduce") and not isinstance(g, torch.distributed.ProcessGroup):
            return g.all_reduce(result)             # vLLM-style group object (LBL:716-717)
        torch.distributed.all_reduce(result, group=g)
        return result

    def forward(self, x: torch.Tensor):
        invalidInputError(not (self.training and x.requires_grad),
                          "training through LowBitLinear is out of scope of this build (inference only)")
        if self.bias is not None and self.bias.dtype != x.dtype:
            self.bias.data = self.bias.data.to(x.dtype)
        if self.optimize_lm_head:
            x = reshape_lm_head_input(x)
        new_shape = x.shape[:-1] + (self.out_len,)
        if 0 in x.shape:
            return torch.empty(new_shape, dtype=x.dtype, device=x.device)   # LBL:668-670
        if self.act_order:
            x = x[..., self.g_idx_map]
        x_2d = (x if x.is_contiguous() else x.contiguous()).view(-1, x.shape[-1])
        invalidInputError(self.weight.quantized, "LowBitLinear weight is not quantised yet: move the module "
                          "to a device first (FP4Params quantises in .to())")
        wdata = self.weight.data

        if wdata.device.type == "cuda":
            if x_2d.device != wdata.device:
                invalidInputError(False, f"activation on {x_2d.device}, weight on {wdata.device}")
            if x_2d.dtype not in _TORCH_TO_DT:
                x_2d = x_2d.float()
            g = self.mp_group
            if g is not None and hasattr(g, "matmul_all_reduce") and x_2d.dtype in _TORCH_TO_DT and \
                    g.fused_supported(x_2d.shape[0], self.out_len, self.qtype):
                # row-parallel decode: the GEMV pushes its partial sums to the peers itself; the bias is added once,
                # after the sum (LBL:715-725)
                result = g.matmul_all_reduce(x_2d, wdata, self.weight_shape, self.qtype, bias=self.bias)
                return result.view(new_shape).to(x.dtype)
            if x_2d.dtype == torch.float32 and self.conver_to_half and x_2d.shape[0] > GEMV_MAX_M:
                # LBL:700-705: a prefill-sized fp32 batch is computed in half precision (tensor-core GEMM);
                # conver_to_half=False keeps fp32 activations on the exact (2^-24) GEMV path at any batch size
                x_2d = x_2d.half()
            fuse_bias = self.bias is not None and self.mp_group is None
            nxt = self.next_linear.weight.data if (self.next_linear is not None and x_2d.shape[0] <= GEMV_MAX_M) else None
            result = lowbit_matmul(x_2d, wdata, self.weight_shape, self.qtype,
                                   bias=self.bias if fuse_bias else None, next_weight=nxt)
            result = result.view(new_shape)
            if self.mp_group is not None:
                result = self._all_reduce(result)
                if self.bias is not None:
                    result += self.bias
        else:
            # weight on the CPU (canonical layout): the reference's CPU call path (LBL:726-754),
            # one call across the reference C ABI per forward; the arithmetic still runs on the GPU
            result = ggml_matmul_src1_x_src0_t(wdata, x_2d.cpu(), self.weight_shape, self.qtype)
            result = result.view(new_shape).to(x.device)
            if self.mp_group is not None:
                result = self._all_reduce(result)
            if self.bias is not None:
                result += self.bias
        return result if result.dtype == x.dtype else result.to(x.dtype)


class vLLMLowBitLinear(LowBitLinear):
    """LBL:916-926"""

    def forward(self, x: torch.Tensor):
        return super().forward(x), None
