"""CUDA-graph capture of a launch-bound step (the FI training step is ~460 small launches: Cayley adjoints of nine
layers, per-layer backbone adjoints, the fused FI loss).  Every launch in this package goes to
``torch.cuda.current_stream()`` through the C ABI and every buffer comes from torch's caching allocator, so a whole
``compute_loss -> backward -> all-reduce`` step is capturable with ``torch.cuda.graph``; replaying it removes the
host-side launch cost.  Inputs are static device buffers: copy the next batch into them before ``replay()``."""
from __future__ import annotations

import os

import torch


class GraphedStep:
    def __init__(self, fn, warmup=3, key_fn=None):
        """``fn()`` must read its inputs from fixed device tensors, do no host synchronisation (no .item() / .cpu())
        and return tensors (or None).  It is run ``warmup`` times on a side stream, then captured once.
        ``key_fn()`` (optional) runs on the host before every replay -- it is where schedule values are written into
        device scalars (``LyapunovLearning.prepare_step``) -- and returns a hashable key of whatever the capture has
        baked in as Python constants (sampler counts, scale_nominal, dropout p): a changed key re-captures."""
        self.fn, self.warmup, self.key_fn = fn, warmup, key_fn
        self.key = key_fn() if key_fn else None
        self.captures = 0
        self._stream = None
        self._capture()

    def _capture(self):
        fn, warmup = self.fn, self.warmup
        cur = torch.cuda.current_stream()
        side = torch.cuda.Stream()
        side.wait_stream(cur)
        with torch.cuda.stream(side):
            for _ in range(warmup):
                fn()
        cur.wait_stream(side)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        # captured on a high-priority stream: kernel nodes keep the priority of the stream they were captured on, so the step's
        # main chain (and the library's branch streams, created with the same priority) is scheduled ahead of the Cayley
        # transforms' side chains whenever both have thread blocks waiting for an SM
        if self._stream is None:
            self._stream = torch.cuda.Stream(priority=-1) if os.environ.get('FIODE_GRAPH_PRIO', '1') != '0' else torch.cuda.Stream()
        with torch.cuda.graph(self.graph, stream=self._stream):
            self.out = fn()
        self.captures += 1

    def __call__(self):
        if self.key_fn is not None:
            key = self.key_fn()
            if key != self.key:
                self.key = key
                self._capture()
        self.graph.replay()
        return self.out
