"""CUDA-graph capture of a launch-bound step (the FI training step is ~460 small launches: Cayley adjoints of nine
layers, per-layer backbone adjoints, the fused FI loss).  Every launch in this package goes to
``torch.cuda.current_stream()`` through the C ABI and every buffer comes from torch's caching allocator, so a whole
``compute_loss -> backward -> all-reduce`` step is capturable with ``torch.cuda.graph``; replaying it removes the
host-side launch cost.  Inputs are static device buffers: copy the next batch into them before ``replay()``."""
from __future__ import annotations

import torch


class GraphedStep:
    def __init__(self, fn, warmup=3):
        """``fn()`` must read its inputs from fixed device tensors, do no host synchronisation (no .item() / .cpu())
        and return tensors (or None).  It is run ``warmup`` times on a side stream, then captured once."""
        cur = torch.cuda.current_stream()
        side = torch.cuda.Stream()
        side.wait_stream(cur)
        with torch.cuda.stream(side):
            for _ in range(warmup):
                fn()
        cur.wait_stream(side)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.out = fn()

    def __call__(self):
        self.graph.replay()
        return self.out
