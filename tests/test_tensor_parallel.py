"""Tensor-parallel host logic on CPU: sharding of packed weights commutes with
quantisation, and a world_size-2 gloo run reproduces the unsharded result with the all-reduce
where the reference puts it (after the matmul, before the bias; low_bit_linear.py:715-725).

The product has no CPU arithmetic, so inside these CPU tests the native matmul of the CPU branch of
LowBitLinear.forward is replaced by the oracle (tests may use the oracle); what is under test is the
sharding, the module wiring and the placement of the collective.
"""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle

QT = {"sym_int4": 2, "asym_int4": 3, "nf4": 10, "fp8_e4m3": 15, "fp6": 29}


@pytest.mark.parametrize("qtype", list(QT))
def test_sharding_commutes_with_quantisation(native_lib, qtype):
    from ipex_llm_b200.transformers import tensor_parallel as tp
    rng = np.random.default_rng(1)
    N, K, world = 96, 512, 4
    w = (rng.standard_normal((N, K)) * 0.02).astype(np.float32)
    packed = torch.from_numpy(oracle.quantize(w, qtype))
    if qtype == "nf4":
        packed = packed.view(torch.bfloat16)
    for rank in range(world):
        col, cshape = tp.shard_packed_column(packed, (N, K), QT[qtype], rank, world)
        r0, r1 = rank * N // world, (rank + 1) * N // world
        assert cshape == (N // world, K) and col.dtype == packed.dtype
        np.testing.assert_array_equal(col.view(torch.uint8).numpy(), oracle.quantize(w[r0:r1], qtype))
        row, rshape = tp.shard_packed_row(packed, (N, K), QT[qtype], rank, world)
        k0, k1 = rank * K // world, (rank + 1) * K // world
        assert rshape == (N, K // world)
        np.testing.assert_array_equal(row.view(torch.uint8).numpy(), oracle.quantize(np.ascontiguousarray(w[:, k0:k1]), qtype))
    # merged qkv: q, k and v are sharded by heads separately
    qkv, shp = tp.shard_qkv_column(packed, (N, K), QT[qtype], 1, 2, q_rows=64, kv_rows=16)
    want = np.concatenate([oracle.quantize(w[32:64], qtype), oracle.quantize(w[72:80], qtype), oracle.quantize(w[88:96], qtype)])
    assert shp == (48, K)
    np.testing.assert_array_equal(qkv.view(torch.uint8).numpy(), want)
    # K shards must be whole blocks
    with pytest.raises(RuntimeError):
        tp.shard_packed_row(packed, (N, K), QT[qtype], 0, 16)        # 512 / 16 = 32 < QK


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _tp_worker(rank, world, port, w, x, bias, qtype, out, peer_group=False):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from ipex_llm_b200.transformers import low_bit_linear as lbl, tensor_parallel as tp
        # CPU stand-in for the native matmul (the product refuses to compute without a GPU)
        lbl.ggml_matmul_src1_x_src0_t = lambda w_, x_, shape, qt: torch.from_numpy(
            oracle.matmul(w_.view(torch.uint8).numpy(), tuple(shape), qt, x_.float().numpy()))
        N, K = w.shape
        I = 2 * N
        rng = np.random.default_rng(5)
        w_up = (rng.standard_normal((I, K)) * 0.02).astype(np.float32)       # column-parallel
        w_down = (rng.standard_normal((N, I)) * 0.02).astype(np.float32)     # row-parallel
        p_up = torch.from_numpy(oracle.quantize(w_up, qtype))
        p_down = torch.from_numpy(oracle.quantize(w_down, qtype))
        up = tp.make_parallel_linear(p_up, (I, K), QT[qtype], "column", rank, world)
        # the group object of a row-parallel linear: a torch process group, or the library's PeerAllReduce (which hands
        # tensors that are not on a GPU -- everything here -- to torch.distributed)
        group = tp.PeerAllReduce(device="cpu") if peer_group else dist.group.WORLD
        down = tp.make_parallel_linear(p_down, (N, I), QT[qtype], "row", rank, world, mp_group=group,
                                       bias=torch.from_numpy(bias))
        assert up.mp_group is None and down.mp_group is not None
        xt = torch.from_numpy(x)
        h = up(xt)                                        # [M, I/world], no communication
        y = down(h)                                       # partial sums all-reduced, then bias once
        # unsharded reference computed locally with the oracle
        h_full = oracle.matmul(p_up.numpy(), (I, K), qtype, x)
        y_full = oracle.matmul(p_down.numpy(), (N, I), qtype, h_full) + bias
        np.testing.assert_allclose(h.numpy(), h_full[:, rank * I // world:(rank + 1) * I // world], rtol=0, atol=0)
        np.testing.assert_allclose(y.numpy(), y_full, rtol=1e-5, atol=1e-6)
        out[rank] = 1
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("qtype,peer_group", [("sym_int4", False), ("fp6", False), ("sym_int4", True)])
def test_row_parallel_allreduce_gloo_world2(native_lib, qtype, peer_group):
    rng = np.random.default_rng(3)
    N, K, M = 64, 256, 3
    w = (rng.standard_normal((N, K)) * 0.02).astype(np.float32)
    x = rng.standard_normal((M, K)).astype(np.float32)
    bias = rng.standard_normal(N).astype(np.float32)
    world = 2
    ctx = mp.get_context("spawn")
    out = ctx.Manager().dict()
    port = _free_port()
    procs = [ctx.Process(target=_tp_worker, args=(r, world, port, w, x, bias, qtype, out, peer_group)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=120)
    assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
    assert sorted(out.keys()) == list(range(world))


class LinearLayer(torch.nn.Module):
    """stand-in for deepspeed.module_inject.layers.LinearLayer: a column-parallel shard, weight [out / world, in]"""

    def __init__(self, weight, bias=None):
        super().__init__()
        self.weight = torch.nn.Parameter(weight, requires_grad=False)
        self.bias = None if bias is None else torch.nn.Parameter(bias, requires_grad=False)

    def forward(self, x):
        y = x @ self.weight.t()
        return y if self.bias is None else y + self.bias


class LinearAllreduce(LinearLayer):
    """stand-in for deepspeed's LinearAllreduce: a row-parallel shard, weight [out, in / world]; the partial outputs
    are summed over ``mp_group`` and the bias is added once"""

    def __init__(self, weight, bias=None, mp_group=None):
        super().__init__(weight, bias)
        self.mp_group = mp_group

    def forward(self, x):
        y = x @ self.weight.t()
        dist.all_reduce(y, group=self.mp_group)
        return y if self.bias is None else y + self.bias


class ShardedMLP(torch.nn.Module):
    def __init__(self, up, down):
        super().__init__()
        self.up_proj, self.down_proj = up, down

    def forward(self, x):
        return self.down_proj(self.up_proj(x))


def _autotp_worker(rank, world, port, w_up, w_down, bias, x, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import ctypes
        import ipex_llm_b200.ggml.model.llama.llama_cpp as ggml
        from ipex_llm_b200 import optimize_model
        from ipex_llm_b200.transformers import low_bit_linear as lbl
        # CPU stand-ins for the two native calls (the product refuses to compute without a GPU)
        lbl.ggml_matmul_src1_x_src0_t = lambda w_, x_, shape, qt: torch.from_numpy(
            oracle.matmul(w_.view(torch.uint8).numpy(), tuple(shape), "sym_int4", x_.float().numpy()))

        def quantize_on_cpu(src, dst, qtype, n, k, hist, scale_search):
            a = np.ctypeslib.as_array(src, shape=(n // k, k))
            packed = oracle.quantize(np.ascontiguousarray(a), "sym_int4")
            ctypes.memmove(dst, packed.ctypes.data, packed.nbytes)
            return packed.nbytes

        ggml.ggml_quantize_tensor = quantize_on_cpu
        I, K = w_up.shape
        N = w_down.shape[0]
        i0, i1 = rank * I // world, (rank + 1) * I // world
        # the AutoTP recipe (example/GPU/Deepspeed-AutoTP/deepspeed_autotp.py:81-90): the fp model is sharded first
        # (deepspeed.init_inference replaces the linears), THEN optimize_model quantises every local shard
        model = ShardedMLP(LinearLayer(torch.from_numpy(w_up[i0:i1].copy())),
                           LinearAllreduce(torch.from_numpy(np.ascontiguousarray(w_down[:, i0:i1])), torch.from_numpy(bias.copy()),
                                           mp_group=dist.group.WORLD))
        model = optimize_model(model, low_bit="sym_int4")
        assert isinstance(model.up_proj, lbl.LowBitLinear) and model.up_proj.mp_group is None
        assert isinstance(model.down_proj, lbl.LowBitLinear) and model.down_proj.mp_group is dist.group.WORLD
        assert model.up_proj.weight_shape == (I // world, K) and model.down_proj.weight_shape == (N, I // world)
        # the local shards were quantised like any other weight: same bytes as slicing the unsharded packed weight
        np.testing.assert_array_equal(model.up_proj.weight.data.numpy(), oracle.quantize(w_up[i0:i1], "sym_int4"))
        y = model(torch.from_numpy(x))
        h_full = oracle.matmul(oracle.quantize(w_up, "sym_int4"), (I, K), "sym_int4", x)
        y_full = oracle.matmul(oracle.quantize(w_down, "sym_int4"), (N, I), "sym_int4", h_full) + bias
        np.testing.assert_allclose(y.numpy(), y_full, rtol=1e-5, atol=1e-6)
        out[rank] = 1
    finally:
        dist.destroy_process_group()


def test_autotp_sharded_model_through_optimize_model_gloo_world2(native_lib):
    rng = np.random.default_rng(11)
    K, I, N, M = 128, 256, 64, 2
    w_up = (rng.standard_normal((I, K)) * 0.05).astype(np.float32)
    w_down = (rng.standard_normal((N, I)) * 0.05).astype(np.float32)
    bias = rng.standard_normal(N).astype(np.float32)
    x = rng.standard_normal((M, K)).astype(np.float32)
    world = 2
    ctx = mp.get_context("spawn")
    out = ctx.Manager().dict()
    port = _free_port()
    procs = [ctx.Process(target=_autotp_worker, args=(r, world, port, w_up, w_down, bias, x, out)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=120)
    assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
    assert sorted(out.keys()) == list(range(world))
