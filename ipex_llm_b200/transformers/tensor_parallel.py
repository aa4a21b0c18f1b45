"""Tensor-parallel sharding of low-bit linears, the way the reference's DeepSpeed-AutoTP recipe
does it (example/GPU/Deepspeed-AutoTP/deepspeed_autotp.py:74-106, convert.py:217-228):

* column-parallel (split the output rows N): qkv / gate / up.  In the canonical layout a row is
  ``K/QK`` consecutive block records, so a row range is one contiguous byte slice (the reference
  relies on the same fact when it concatenates packed q, k, v: models/common.py:395-410);
* row-parallel (split K): o_proj / down_proj.  Each rank keeps a block range of every row; its
  partial output is summed across ranks by ONE all-reduce placed after the matmul and before the
  bias (low_bit_linear.py:715-725) -- ``LowBitLinear._all_reduce`` with ``mp_group`` set.

Both slicings commute with quantisation: blocks never straddle a shard boundary (K/world must be
a multiple of QK), so slicing packed bytes is bit-identical to quantising the fp16 shard, which is
what the reference does (it shards first, then calls optimize_model).  All functions here are
byte/index manipulation on the CANONICAL layout; nothing is computed.
"""
import torch

import ipex_llm_b200.ggml.model.llama.llama_cpp as ggml
from ipex_llm_b200.utils.common import invalidInputError
from ipex_llm_b200.ggml.quantize import B200_SUPPORTED_QTYPE_IDS


def _as_bytes(packed):
    return packed.view(torch.uint8) if packed.dtype != torch.uint8 else packed


def shard_bounds(total, world, rank, unit=1):
    """[begin, end) of an even split of ``total`` (a multiple of ``unit * world``)."""
    invalidInputError(total % (unit * world) == 0,
                      f"cannot split {total} into {world} shards of a multiple of {unit}")
    per = total // world
    return rank * per, (rank + 1) * per


def shard_packed_column(packed: torch.Tensor, shape, qtype: int, rank: int, world: int):
    """rows [rank*N/world, (rank+1)*N/world) of a canonical packed [N, K] weight -> (packed, shape)"""
    n, k = shape
    qk, bs = ggml.ggml_qk_size(qtype), ggml.ggml_type_size(qtype)
    r0, r1 = shard_bounds(n, world, rank)
    row_bytes = (k // qk) * bs
    b = _as_bytes(packed)
    invalidInputError(b.numel() == n * row_bytes, "packed size does not match shape")
    out = b[r0 * row_bytes:r1 * row_bytes].clone()
    return (out.view(packed.dtype) if packed.dtype != torch.uint8 else out), (r1 - r0, k)


def shard_packed_row(packed: torch.Tensor, shape, qtype: int, rank: int, world: int):
    """columns [rank*K/world, (rank+1)*K/world) (whole blocks) of a canonical packed [N, K] weight"""
    n, k = shape
    qk, bs = ggml.ggml_qk_size(qtype), ggml.ggml_type_size(qtype)
    k0, k1 = shard_bounds(k, world, rank, unit=qk)
    nb = k // qk
    b = _as_bytes(packed)
    invalidInputError(b.numel() == n * nb * bs, "packed size does not match shape")
    out = b.view(n, nb, bs)[:, k0 // qk:k1 // qk, :].contiguous().view(-1)
    return (out.view(packed.dtype) if packed.dtype != torch.uint8 else out), (n, k1 - k0)


def shard_qkv_column(packed: torch.Tensor, shape, qtype: int, rank: int, world: int, q_rows: int, kv_rows: int):
    """merged qkv_proj ([q | k | v] rows, models/common.py:23-53): every rank takes its slice of the
    heads of q, of k and of v (AutoTP shards q, k, v separately, before the merge)."""
    n, k = shape
    invalidInputError(n == q_rows + 2 * kv_rows, "qkv row counts do not add up")
    qk, bs = ggml.ggml_qk_size(qtype), ggml.ggml_type_size(qtype)
    row_bytes = (k // qk) * bs
    b = _as_bytes(packed)
    parts, rows = [], 0
    for base, cnt in ((0, q_rows), (q_rows, kv_rows), (q_rows + kv_rows, kv_rows)):
        r0, r1 = shard_bounds(cnt, world, rank)
        parts.append(b[(base + r0) * row_bytes:(base + r1) * row_bytes])
        rows += r1 - r0
    out = torch.cat(parts)
    return (out.view(packed.dtype) if packed.dtype != torch.uint8 else out), (rows, k)


def make_parallel_linear(packed, shape, qtype, mode, rank, world, mp_group=None, bias=None, device=None):
    """Build the rank-local ``LowBitLinear`` of a column- ("column") or row-parallel ("row") linear
    from the unsharded canonical packed weight.  Row-parallel modules carry ``mp_group`` (their
    forward all-reduces); the bias of a row-parallel linear is added once, after the all-reduce."""
    from ipex_llm_b200.transformers.low_bit_linear import LowBitLinear, FP4Params
    invalidInputError(mode in ("column", "row"), "mode must be 'column' or 'row'")
    if mode == "column":
        p, shp = shard_packed_column(packed, shape, qtype, rank, world)
        b = None if bias is None else bias[shard_bounds(shape[0], world, rank)[0]:shard_bounds(shape[0], world, rank)[1]].clone()
        group = None
    else:
        p, shp = shard_packed_row(packed, shape, qtype, rank, world)
        b = None if bias is None else bias.clone()
        group = mp_group
    lin = LowBitLinear(shp[1], shp[0], qtype, bias=b is not None, mp_group=group)
    lin.weight = FP4Params(p, quantized=True, _shape=torch.Size(shp), qtype=qtype)
    if b is not None:
        lin.bias = torch.nn.Parameter(b, requires_grad=False)
    if device is not None:
        lin = lin.to(device)
    return lin


class PeerAllReduce:
    """``mp_group`` object for row-parallel ``LowBitLinear`` modules (their forward calls ``mp_group.all_reduce``,
    low_bit_linear.py:715-717 in the reference): the sum over the ranks of one node is computed by this library's
    kernel (csrc/comm.cu), which pushes every rank's vector into the peers' buffers over NVLink, instead of by NCCL -- for
    the small vectors of decode (M x hidden, tens of KiB) the cost is one launch and one NVLink crossing.  Tensors larger than
    ``lowbit_comm_max_bytes()`` (prefill), or whose byte size is not a multiple of 16, go to
    ``torch.distributed.all_reduce`` on ``group``.

    One communicator (buffer + mapped peer buffers) is created per payload size on first use; creation is a
    collective over ``group`` (handles travel by ``all_gather_object``), so every rank must reduce the same sizes in
    the same order -- which tensor-parallel decoding does.  Results are bit-identical on all ranks.
    """

    def __init__(self, group=None, device=None):
        import torch.distributed as dist
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        invalidInputError(self.world <= 8, "PeerAllReduce serves the GPUs of one node (world size <= 8)")
        self._comms = {}                               # payload bytes -> (own pointer, [peer pointers])

    def _comm(self, nbytes):
        import torch.distributed as dist
        c = self._comms.get(nbytes)
        if c is None:
            with torch.cuda.device(self.device):
                own, handle = ggml.lowbit_comm_create(nbytes)
                handles = [None] * self.world
                dist.all_gather_object(handles, handle, group=self.group)
                peers = [own if r == self.rank else ggml.lowbit_comm_open(handles[r]) for r in range(self.world)]
            dist.barrier(group=self.group)             # every rank has mapped every buffer before the first use
            c = (own, peers)
            self._comms[nbytes] = c
        return c

    def all_reduce(self, tensor):
        import torch.distributed as dist
        nbytes = tensor.numel() * tensor.element_size()
        if (self.world == 1 or not tensor.is_cuda or nbytes % 16 or nbytes > ggml.lowbit_comm_max_bytes() or not tensor.is_contiguous()
                or tensor.dtype not in (torch.float32, torch.float16, torch.bfloat16) or tensor.data_ptr() % 16):
            if self.world > 1:
                dist.all_reduce(tensor, group=self.group)
            return tensor
        own, peers = self._comm(nbytes)
        with torch.cuda.device(tensor.device):
            ggml.lowbit_allreduce_sum(peers, self.rank, nbytes, tensor, tensor, torch.cuda.current_stream(tensor.device).cuda_stream)
        return tensor

    def fused_supported(self, m, n_rows, qtype):
        """can ``matmul_all_reduce`` take this row-parallel matmul?  (decode sizes: M <= 32, 64-row groups, fp32 partials
        within the communicator limit)"""
        return (self.world > 1 and 1 <= m <= 32 and n_rows % 64 == 0 and m * n_rows * 4 <= ggml.lowbit_comm_max_bytes()
                and qtype in B200_SUPPORTED_QTYPE_IDS)

    def matmul_all_reduce(self, x_2d, weight, weight_shape, qtype, bias=None, out=None, accumulate=False):
        """``all_reduce(x_2d @ dequant(W).T) + bias`` of a row-parallel linear in one call: the GEMV pushes its fp32
        partial sums to the peers from its epilogue (``lowbit_matmul_allreduce_cuda``)."""
        from ipex_llm_b200.transformers import low_bit_linear as lbl
        n_rows, k = weight_shape
        m = x_2d.shape[0]
        invalidInputError(self.fused_supported(m, n_rows, qtype), "matmul_all_reduce: shape outside the fused kernel's range")
        x_2d = x_2d.contiguous()
        if out is None:
            out = torch.empty((m, n_rows), dtype=x_2d.dtype, device=x_2d.device)
        if bias is not None and bias.dtype != out.dtype:
            bias = bias.to(out.dtype)
        own, peers = self._comm(m * n_rows * 4)
        dt = lbl._TORCH_TO_DT[x_2d.dtype]
        ws = lbl._get_workspace(x_2d.device, ggml.lowbit_matmul_workspace_size(n_rows, k, m, dt))
        w = weight.view(torch.uint8) if weight.dtype != torch.uint8 else weight
        with torch.cuda.device(x_2d.device):
            ggml.lowbit_matmul_allreduce(w, qtype, n_rows, k, x_2d, out, bias, accumulate, peers, self.rank, m * n_rows * 4, ws,
                                         torch.cuda.current_stream(x_2d.device).cuda_stream)
        return out

    def timed_out(self):
        """True if a wait inside a kernel ran into its bound (a peer did not take part).  Synchronises the device."""
        return any(ggml.lowbit_comm_status(own) for own, _ in self._comms.values())

    def check(self):
        """raise if a wait inside a kernel ever ran out on one of the communicators (their results are NaN from then
        on).  Synchronises the device: call it at a point where the host waits for results anyway."""
        invalidInputError(not self.timed_out(),
                          "PeerAllReduce: a rank did not take part in an all-reduce within LOWBIT_COMM_TIMEOUT_S seconds; "
                          "the results since then are NaN and the communicator is unusable")

    def close(self):
        dead = self.timed_out() if self._comms else False
        self._close()
        invalidInputError(not dead, "PeerAllReduce: a wait ran out during this communicator's life (results were NaN)")

    def _close(self):
        for own, peers in self._comms.values():
            for r, p in enumerate(peers):
                if r != self.rank:
                    ggml.lowbit_comm_close(p)
            ggml.lowbit_comm_destroy(own)
        self._comms = {}
