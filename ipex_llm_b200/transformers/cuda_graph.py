"""CUDA-graph replay of a decode step built from ``LowBitLinear`` calls.

At B200 speeds one decoded token is ~1 ms of GPU work spread over 161 ``LowBitLinear.forward`` calls;
issuing them from Python costs several times that.  ``capture_decode_step`` records the calls once
(after a warm-up that sizes every workspace) and replays them with one launch; inputs and outputs live
in static tensors the caller copies into / out of.  Programmatic dependent launch and the L2
next-weight hints set on the modules are recorded with the kernels.

(The reference has no counterpart: its XPU path issues every op from Python.  This is plumbing around
the hot path, not part of it.)
"""
import torch


class GraphedStep:
    def __init__(self, fn, graph, outputs, stream):
        self.fn, self.graph, self.outputs, self.stream = fn, graph, outputs, stream

    def replay(self):
        """enqueue the recorded step on the capture stream; returns the static output tensor(s)"""
        with torch.cuda.stream(self.stream):
            self.graph.replay()
        return self.outputs


def capture_decode_step(fn, warmup=2, stream=None):
    """``fn()`` must read its inputs from, and write its result to, tensors that stay alive (static
    buffers); it is run ``warmup`` times eagerly, then captured.  Returns a ``GraphedStep``."""
    stream = stream or torch.cuda.Stream()
    with torch.cuda.stream(stream):
        for _ in range(warmup):
            fn()
        stream.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=stream):
            outputs = fn()
    torch.cuda.synchronize()
    return GraphedStep(fn, graph, outputs, stream)
