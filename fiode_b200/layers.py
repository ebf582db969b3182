"""Cayley-orthogonal layers with the reference's names, constructor arguments and state_dict keys
(libs/ortho_conv/layers.py), running on libfiode_b200.so.

Differences a maintainer should know about:
  * the Cayley matrices are cached per *weight version* (the reference recomputes the conv ones on
    every forward and never invalidates the linear ones in eval mode, layers.py:72-86, 103-105);
  * in ``.eval()`` mode (inference / certification) the layers run forward only on the fused kernels;
    in ``.train()`` mode with gradients enabled every layer goes through ``fiode_b200.autograd`` -- one C-ABI
    call forward, one hand-written adjoint backward (csrc/train.cu, csrc/cayley.cu) -- so that
    ``loss.backward()`` fills ``weight.grad / bias.grad / alpha.grad`` like the reference's autograd graph.
"""
from __future__ import annotations

import os

import torch
from torch import nn

from . import ops
from . import autograd as ag


def _differentiable(module, x):
    """The reference's graph reaches the weights only in train mode (layers.py:103-105: eval uses Q.detach()); the input
    gradient (PGD in validation_step, pl_modules.py:189-199) flows in either mode."""
    if not torch.is_grad_enabled():
        return False
    return x.requires_grad or (module.training and any(p.requires_grad for p in module.parameters(recurse=False)))


def cayley(W):
    """libs/ortho_conv/layers.py:50-60 for a real 2-D weight."""
    assert W.dim() == 2 and not W.is_complex(), 'fiode_b200.layers.cayley handles real (cout, cin) matrices'
    return ops.cayley_linear(W, W.norm())


class GroupSort(nn.Module):
    """layers.py:127-131."""
    def forward(self, x):
        if torch.is_grad_enabled() and x.requires_grad:
            return ag.GroupSortFn.apply(x)
        return ops.group_sort(x)


def _version(*ts):
    return tuple((t.data_ptr(), t._version) for t in ts)


def _take_prefetch(module, n=None):
    """The Q that ``prefetch_orthogonal`` prepared for this forward, if it is still the Q of the current parameters."""
    pf, module._q_prefetch = module._q_prefetch, None
    if pf is None:
        return None
    key, pn, Q, st = pf
    if st is not None:                 # prepared on a side stream: the consumer joins it here, not the whole step up front
        torch.cuda.current_stream().wait_stream(st)
    if key != _version(module.weight, module.alpha) or pn != n:
        return None
    return Q


def finish_prefetch(*roots):
    """Join the side streams of prepared layers that no forward has picked up yet (a stream forked inside a CUDA-graph
    capture has to come back to the capturing stream; eager code loses nothing by it)."""
    for root in roots:
        for m in root.modules():
            pf = getattr(m, '_q_prefetch', None)
            if pf is not None and pf[3] is not None:
                torch.cuda.current_stream().wait_stream(pf[3])
                m._q_prefetch = pf[:3] + (None,)


_PREFETCH = os.environ.get('FIODE_PREFETCH', '1') != '0'
_side_streams = {}


def _side_stream(device, i):
    key = (torch.device(device).index, i)
    if key not in _side_streams:
        _side_streams[key] = torch.cuda.Stream(device=device)
    return _side_streams[key]


def prefetch_orthogonal(*roots):
    """Every Cayley transform of a training step up front.  The reference evaluates ``cayley(alpha W / ||W||)`` inside each
    layer's forward (libs/ortho_conv/layers.py:74-82, :103-104); the transforms depend on the weights only, so here the small
    layers' go out as ONE launch (``CayleySmallBatchQ``, one CTA per layer, one more launch for all their adjoints at the end of
    backward) and each large layer's chain of ~25 dependent launches (eight for the Gauss-Jordan steps of a 512 x 512 matrix alone) runs on its
    own side stream next to the others; autograd runs each node's backward on the stream of its forward, so the adjoints
    overlap the activation backward the same way.  Values and gradients are those of the per-layer path, bit for bit.
    Each layer's forward picks its Q up and joins that layer's stream there (``_take_prefetch``), so the first layers' activations
    run while the last layers' matrices are still being inverted; ``finish_prefetch`` joins what no forward reached.
    ``FIODE_PREFETCH=0`` turns this off.  Returns the number of layers prepared."""
    if not _PREFETCH or not torch.is_grad_enabled():
        return 0
    small, big = [], []
    for root in roots:
        for m in root.modules():
            if not isinstance(m, (CayleyLinear, CayleyConv)) or not m.training:
                continue
            m._q_prefetch = None
            if not (m.weight.is_cuda and m.weight.requires_grad):
                continue
            if isinstance(m, CayleyConv):
                if m.alpha is None or m._last_n is None:
                    continue                     # first step: alpha and the spatial size appear in the layer's own forward
                n = m._last_n
                if m.kernel_size[0] == 1 and ops.cayley_small_ok(m.weight.shape[0], m.weight.shape[1]):
                    small.append((m, n, float(n * (n // 2 + 1)) ** 0.5))
                else:
                    big.append((m, n))
            elif ops.cayley_small_ok(*m.weight.shape):
                small.append((m, None, 1.0))
            else:
                big.append((m, None))
    if not small and not big:
        return 0
    cur = torch.cuda.current_stream()
    big.sort(key=lambda e: -e[0].weight.numel())
    for i, (m, n) in enumerate(big):
        st = _side_stream(m.weight.device, i)
        st.wait_stream(cur)
        with torch.cuda.stream(st):
            Q = ag.CayleyLinearQ.apply(m.weight, m.alpha) if n is None else ag.CayleyConvQ.apply(m.weight, m.alpha, n)
        m._q_prefetch = (_version(m.weight, m.alpha), n, Q, st)
    # small layers: the linear ones (dynamics, fc -- their dL/dQ all arrive early in backward) as one node, each stride-2 conv's
    # matrix as its own (its dL/dQ arrives late; sharing a node would hold every small adjoint back to the end of the step)
    groups = [[e for e in small if e[1] is None]] + [[e] for e in small if e[1] is not None]
    for j, grp in enumerate(g for g in groups if g):
        st = _side_stream(grp[0][0].weight.device, len(big) + j)
        st.wait_stream(cur)
        flat = []
        for m, _, _ in grp:
            flat += [m.weight, m.alpha]
        with torch.cuda.stream(st):
            Qs = ag.CayleySmallBatchQ.apply(tuple(mu for _, _, mu in grp), *flat)
        for (m, n, _), Q in zip(grp, Qs):
            m._q_prefetch = (_version(m.weight, m.alpha), n, Q, st)
    return len(small) + len(big)


class CayleyLinear(nn.Linear):
    """layers.py:88-105.  ``Q`` holds the cached orthogonal matrix like the reference attribute."""
    def __init__(self, in_features, out_features, bias=True):
        super().__init__(in_features, out_features, bias)
        self.alpha = nn.Parameter(torch.ones(1, dtype=torch.float32))
        self.alpha.data = self.weight.norm()
        self._q_key = None
        self._q_prefetch = None

    def reset_parameters(self):
        std = 1 / self.weight.shape[1] ** 0.5
        nn.init.uniform_(self.weight, -std, std)
        if self.bias is not None:
            self.bias.data.uniform_(-std, std)
        self.Q = None

    def orthogonal_weight(self):
        key = _version(self.weight, self.alpha)
        if self.Q is None or self._q_key != key:
            self.Q = ops.cayley_linear(self.weight, self.alpha)
            self._q_key = key
        return self.Q

    def orthogonal_weight_autograd(self):
        """Q as a node of the autograd graph (its backward is fiode_cayley_linear_bwd); the detached value also
        refreshes the cache that ``orthogonal_weight`` / ``dyn_weights`` read."""
        Q = _take_prefetch(self)
        if Q is None:
            Q = ag.CayleyLinearQ.apply(self.weight, self.alpha)
        self.Q, self._q_key = Q.detach(), _version(self.weight, self.alpha)
        return Q

    def forward(self, X, group_sort=False):
        if _differentiable(self, X):
            Q = self.orthogonal_weight_autograd() if self.training else self.orthogonal_weight()
            y = ag.Linear.apply(X, Q, self.bias)
            return ag.GroupSortFn.apply(y) if group_sort else y
        return ops.linear_fwd(X, self.orthogonal_weight(), self.bias, group_sort)


class CayleyConv(nn.Conv2d):
    """layers.py:24-48 (StridedConv: stride 2 is emulated by space-to-depth, 4x in_channels,
    kernel // 2) + :62-86 (CayleyConv).  ``alpha`` is created lazily at the first forward, as in the
    reference (:80-81), or loaded from a checkpoint that already has it."""
    def __init__(self, in_channels, out_channels, kernel_size, stride=1, bias=True):
        self.striding = stride == 2
        if self.striding:
            in_channels, kernel_size = 4 * in_channels, max(1, kernel_size // 2)
        super().__init__(in_channels, out_channels, kernel_size, stride=1, padding=kernel_size // 2, bias=bias)
        self.register_parameter('alpha', None)
        self.Qf = None
        self._q_key = None
        self._q_prefetch = None
        self._last_n = None            # spatial size of the last training forward: what prefetch_orthogonal prepares for

    def _load_from_state_dict(self, state_dict, prefix, *args, **kwargs):
        k = prefix + 'alpha'
        if k in state_dict and self.alpha is None:
            self.alpha = nn.Parameter(torch.zeros((), dtype=torch.float32, device=self.weight.device))
        super()._load_from_state_dict(state_dict, prefix, *args, **kwargs)

    def orthogonal_spectrum(self, n):
        if self.alpha is None:
            self.alpha = nn.Parameter(ops.conv_alpha_init(self.weight, n))
        key = _version(self.weight, self.alpha) + (n,)
        if self.Qf is None or self._q_key != key:
            self.Qf = ops.cayley_conv(self.weight, self.alpha, n)
            self._q_key = key
        return self.Qf

    def forward(self, x, group_sort=False):
        n = x.shape[-1] // (2 if self.striding else 1)
        if _differentiable(self, x):
            if self.alpha is None:
                self.alpha = nn.Parameter(ops.conv_alpha_init(self.weight, n))
            if self.training:
                Qf = _take_prefetch(self, n)
                if Qf is None:
                    Qf = ag.CayleyConvQ.apply(self.weight, self.alpha, n)
                self._last_n = n
            else:
                Qf = self.orthogonal_spectrum(n)
            y = (ag.Conv1x1S2d if self.striding else ag.CayleyConvLayer).apply(x, Qf, self.bias)
            return ag.GroupSortFn.apply(y) if group_sort else y
        Qf = self.orthogonal_spectrum(n)
        if self.striding:
            assert self.kernel_size[0] == 1
            return ops.conv1x1_s2d_fwd(x, Qf, self.bias, group_sort)
        return ops.cayley_conv_fwd(x, Qf, self.bias, group_sort)
