"""LowBitEmbedding (reference transformers/embedding.py:179-241): an ``nn.Embedding`` whose table is stored
in a low-bit block format and dequantised row by row on the GPU (``xe_linear.dequantize_rows`` there,
``lowbit_dequantize_rows_cuda`` here).  The table is an ``FP4Params`` like a linear's weight: canonical layout
on the CPU / on disk, tiled device layout on CUDA."""
import ctypes
from typing import Optional

import torch
from torch import Tensor

import ipex_llm_b200.ggml.model.llama.llama_cpp as ggml
from ipex_llm_b200.transformers.low_bit_linear import FP4Params, _TORCH_TO_DT, _stream_ptr
from ipex_llm_b200.utils.common import invalidInputError


class LowBitEmbedding(torch.nn.Embedding):
    def __init__(self, num_embeddings: int, embedding_dim: int, padding_idx: Optional[int] = None,
                 max_norm: Optional[float] = None, norm_type: float = 2., scale_grad_by_freq: bool = False,
                 sparse: bool = False, _weight: Optional[Tensor] = None, _freeze: bool = False, device=None,
                 dtype=None, convert_shape_only=None, qtype=None) -> None:
        super().__init__(num_embeddings, embedding_dim, padding_idx, max_norm, norm_type, scale_grad_by_freq,
                         sparse, _weight, device=device, dtype=dtype)
        invalidInputError(embedding_dim % 64 == 0, "embedding_dim must be a multiple of the block size 64")
        self.qweight = FP4Params(self.weight.data.float(), requires_grad=False, quantized=False, _shape=None,
                                 convert_shape_only=bool(convert_shape_only), qtype=qtype, in_features=embedding_dim)
        # the dummy weight records the model's dtype and device (embedding.py:204-206)
        dummy_weight = torch.empty(0, 0, dtype=self.weight.dtype, device=self.weight.device)
        self.weight = torch.nn.Parameter(dummy_weight, requires_grad=False)
        self.embedding_dim = embedding_dim
        self.num_embeddings = num_embeddings

    def forward(self, x: Tensor):
        invalidInputError(x.device.type == "cuda", "`LowBitEmbedding` only supports GPU now.")
        invalidInputError(x.dtype in (torch.int32, torch.int64), "indices must be int32 or int64")
        invalidInputError(self.qweight.quantized and self.qweight.data.device.type == "cuda",
                          "move the module to the GPU first (quantises and re-lays out the table)")
        out_dtype = self.weight.dtype if self.weight.dtype in _TORCH_TO_DT else torch.float32
        ids = x.contiguous()
        out = torch.empty(tuple(x.shape) + (self.embedding_dim,), dtype=out_dtype, device=x.device)
        w = self.qweight.data
        w = w.view(torch.uint8) if w.dtype != torch.uint8 else w
        with torch.cuda.device(x.device):
            ggml.lowbit_dequantize_rows_cuda(ctypes.c_void_p(w.data_ptr()), self.qweight.qtype, self.num_embeddings,
                                             self.embedding_dim, ctypes.c_void_p(ids.data_ptr()),
                                             1 if ids.dtype == torch.int64 else 0, ids.numel(),
                                             ctypes.c_void_p(out.data_ptr()), _TORCH_TO_DT[out_dtype], _stream_ptr(x.device))
        return out.to(self.weight.dtype)

    @classmethod
    def from_embedding(cls, embedding: torch.nn.Embedding, convert_shape_only, qtype):
        return cls(embedding.num_embeddings, embedding.embedding_dim, embedding.padding_idx, embedding.max_norm,
                   embedding.norm_type, embedding.scale_grad_by_freq, embedding.sparse, embedding.weight.data, True,
                   embedding.weight.device, embedding.weight.dtype, convert_shape_only, qtype)
