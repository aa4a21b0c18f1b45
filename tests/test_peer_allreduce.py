"""The library's own all-reduce for tensor-parallel decode (csrc/comm.cu; the reference calls
torch.distributed / DeepSpeed all_reduce at low_bit_linear.py:715-725).  2, 4 and 8 ranks, one process per GPU
(each case is skipped on a box with fewer GPUs).  Expected values are the rank-ordered fp32 sums, rounded once to the
tensor's dtype: the kernel must match them bit for bit, on every rank.
"""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _rank_input(rank, n, dtype, seed, dev):
    gen = torch.Generator(device=dev).manual_seed(1000 * seed + rank)
    return torch.randn(n, device=dev, generator=gen).to(dtype)


def _expected(world, n, dtype, seed, dev):
    acc = torch.zeros(n, device=dev, dtype=torch.float32)
    for r in range(world):
        acc = acc + _rank_input(r, n, dtype, seed, dev).float()              # rank order, fp32
    return acc.to(dtype)


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("cpu:gloo,cuda:nccl", rank=rank, world_size=world, device_id=dev)
    try:
        from ipex_llm_b200.transformers import low_bit_linear as lbl, tensor_parallel as tp
        from ipex_llm_b200.ggml.quantize import ggml_tensor_qtype
        group = tp.PeerAllReduce()
        seed = 0
        # every dtype, sizes from one chunk to many, back to back (both slots reused many times)
        for dtype in (torch.float32, torch.float16, torch.bfloat16):
            for n in (8, 4096, 8192, 8192, 5120, 40960):
                for _ in range(6):
                    seed += 1
                    x = _rank_input(rank, n, dtype, seed, dev)
                    y = group.all_reduce(x)
                    assert y.data_ptr() == x.data_ptr()
                    assert torch.equal(y, _expected(world, n, dtype, seed, dev)), (dtype, n, seed)
        # a captured graph replays (the call counters live in device memory)
        n = 8192
        bufs = [torch.zeros(n, device=dev, dtype=torch.float16) for _ in range(3)]
        stream = torch.cuda.Stream(device=dev)
        with torch.cuda.stream(stream):
            for b in bufs:
                group.all_reduce(b)                      # warm-up outside capture (communicator creation is a collective)
            torch.cuda.synchronize()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=stream):
                for b in bufs:
                    group.all_reduce(b)
            for it in range(5):
                for i, b in enumerate(bufs):
                    b.copy_(_rank_input(rank, n, torch.float16, 500 + 10 * it + i, dev))
                graph.replay()
                torch.cuda.synchronize()
                for i, b in enumerate(bufs):
                    assert torch.equal(b, _expected(world, n, torch.float16, 500 + 10 * it + i, dev)), (it, i)
        # sizes the kernel does not take go to torch.distributed: 2 MiB, and a byte count that is not a multiple of 16
        for n, dtype in ((1 << 19, torch.float32), (1001, torch.float16)):
            seed += 1
            x = _rank_input(rank, n, dtype, seed, dev)
            want = _expected(world, n, dtype, seed, dev)
            y = group.all_reduce(x)
            assert float((y.float() - want.float()).abs().max()) <= 2e-2
        # a row-parallel LowBitLinear with this group against the unsharded linear
        qt = ggml_tensor_qtype["sym_int4"]
        N, K = 4096, 8192
        gen = torch.Generator(device=dev).manual_seed(7)
        w = torch.empty(N, K, device=dev).normal_(0.0, 0.02, generator=gen)
        xin = torch.randn(1, K, device=dev, generator=gen).half()
        packed = lbl.ggml_convert_qtype(w, qt).cpu()
        full = lbl.LowBitLinear(K, N, qt, bias=False)
        full.weight = lbl.FP4Params(packed.clone(), quantized=True, _shape=torch.Size((N, K)), qtype=qt)
        full = full.to(dev)
        part = tp.make_parallel_linear(packed, (N, K), qt, "row", rank, world, mp_group=group, device=dev)
        k0, k1 = tp.shard_bounds(K, world, rank, unit=64)
        y_tp = part(xin[:, k0:k1].contiguous())
        y_full = full(xin)
        assert float((y_tp.float() - y_full.float()).abs().max()) <= 2e-3 * float(y_full.float().abs().max())
        gathered = [torch.empty_like(y_tp) for _ in range(world)]
        dist.all_gather(gathered, y_tp)
        assert all(torch.equal(g, y_tp) for g in gathered)                   # the same bits on every rank
        # the fused form: the GEMV pushes its fp32 partial sums itself (LowBitLinear uses it for M <= 32); the sum over
        # ranks is taken before the rounding to fp16, so it is compared with the unsharded fp32 result
        assert isinstance(part.mp_group, tp.PeerAllReduce) and group.fused_supported(1, N, qt)
        for M in (1, 5, 16):
            xm = torch.randn(M, K, device=dev, generator=gen).half()
            bias = torch.randn(N, device=dev, generator=gen).half()
            want32 = lbl.lowbit_matmul(xm, full.weight.data, (N, K), qt, out_dtype=torch.float32)
            got = group.matmul_all_reduce(xm[:, k0:k1].contiguous(), part.weight.data, part.weight_shape, qt, bias=bias)
            ref = (want32 + bias.float())
            assert float((got.float() - ref).abs().max()) <= 1.5e-3 * float(ref.abs().max()), M
            res = torch.randn(M, N, device=dev, generator=gen).half()
            acc = res.clone()
            for _ in range(3):                           # back to back: both slots, same buffers
                acc.copy_(res)
                group.matmul_all_reduce(xm[:, k0:k1].contiguous(), part.weight.data, part.weight_shape, qt, out=acc, accumulate=True)
            assert float((acc.float() - (want32 + res.float())).abs().max()) <= 1.5e-3 * float((want32 + res.float()).abs().max()), M
            gathered = [torch.empty_like(got) for _ in range(world)]
            dist.all_gather(gathered, got)
            assert all(torch.equal(g, got) for g in gathered)
        # ... and interleaved with plain all-reduces on a communicator of the same size, inside a graph
        yb = torch.empty(1, N, device=dev, dtype=torch.float16)
        t32 = torch.zeros(N, device=dev, dtype=torch.float32)              # N * 4 bytes: the fused call's communicator
        xs = xin[:, k0:k1].contiguous()
        with torch.cuda.stream(stream):
            group.all_reduce(t32)
            torch.cuda.synchronize()
            graph2 = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph2, stream=stream):
                for _ in range(3):
                    group.matmul_all_reduce(xs, part.weight.data, part.weight_shape, qt, out=yb)
                    group.all_reduce(t32)
            for it in range(4):
                t32.fill_(float(rank + 1))
                graph2.replay()
                torch.cuda.synchronize()
                assert float((yb.float() - y_full.float()).abs().max()) <= 2e-3 * float(y_full.float().abs().max())
                base = float(sum(range(1, world + 1)))
                assert torch.equal(t32, torch.full_like(t32, base * world * world)), it     # three all-reduces of the running value
        assert not group.timed_out()
        group.close()
        out[rank] = 1
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4, 8])
def test_peer_allreduce_ranks(native_lib, world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    ctx = mp.get_context("spawn")
    out = ctx.Manager().dict()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, out)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=240)
    for p in procs:
        if p.is_alive():
            p.kill()
    assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
    assert sorted(out.keys()) == list(range(world))


def _expiry_worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), LOWBIT_COMM_TIMEOUT_S="0.3")
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("cpu:gloo,cuda:nccl", rank=rank, world_size=world, device_id=dev)
    try:
        from ipex_llm_b200.transformers import tensor_parallel as tp
        group = tp.PeerAllReduce()
        x = torch.ones(4096, device=dev, dtype=torch.float16)
        group.all_reduce(x)                                   # a normal call: every rank takes part
        torch.cuda.synchronize()
        assert torch.equal(x, torch.full_like(x, float(world))) and not group.timed_out()
        dist.barrier()
        if rank == 0:
            # the peer never makes this call: the wait runs out, the result is NaN (not a partial sum), the communicator
            # reports it, and every later call on it gives NaN at once
            y = torch.ones(4096, device=dev, dtype=torch.float16)
            group.all_reduce(y)
            torch.cuda.synchronize()
            assert bool(torch.isnan(y).all())
            assert group.timed_out()
            with pytest.raises(RuntimeError):
                group.check()
            z = torch.ones(4096, device=dev, dtype=torch.float16)
            group.all_reduce(z)
            torch.cuda.synchronize()
            assert bool(torch.isnan(z).all())
            with pytest.raises(RuntimeError):
                group.close()
        dist.barrier()
        if rank != 0:
            group.close()
        out[rank] = 1
    finally:
        dist.destroy_process_group()


def test_peer_allreduce_missing_peer_fails_loudly(native_lib):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    ctx = mp.get_context("spawn")
    out = ctx.Manager().dict()
    port = _free_port()
    procs = [ctx.Process(target=_expiry_worker, args=(r, 2, port, out)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=120)
    for p in procs:
        if p.is_alive():
            p.kill()
    assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
    assert sorted(out.keys()) == [0, 1]
