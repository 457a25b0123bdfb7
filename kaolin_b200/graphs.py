"""CUDA-graph helper for launch-bound use of the DIB-R path.

Every entry point of libdibr_b200.so only enqueues work on the caller's stream (no allocation, no
synchronisation, no host round trip), so a whole step — host->device copies from pinned memory, forward,
backward — can be captured once and replayed with a single launch.  For small renders (config 1: ~0.1 ms of GPU
work behind ~0.25 ms of Python / autograd dispatch) this removes the dispatch cost; for large ones it lets
several view chunks be pipelined across streams without paying the dispatch per chunk.

    step = GraphedStep(fn)      # fn() runs the step on STATIC tensors and returns a tensor or a tuple
    out = step()                # replay; `out` are the same static output tensors every time
"""
import torch


class GraphedStep:
    def __init__(self, fn, warmup=3, device=None):
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else device
        self.graph = torch.cuda.CUDAGraph()
        side = torch.cuda.Stream(self.device)
        side.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(side):              # warm-up off the default stream, as torch's graph recipe asks
            for _ in range(warmup):
                fn()
        torch.cuda.current_stream(self.device).wait_stream(side)
        torch.cuda.synchronize(self.device)
        with torch.cuda.graph(self.graph):
            self.outputs = fn()

    def __call__(self):
        self.graph.replay()
        return self.outputs
