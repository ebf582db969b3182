"""torch.autograd glue for the FI training step: every Function's forward AND backward is one C-ABI call
into libfiode_b200.so (fiode_b200/csrc/train.cu, cayley.cu); autograd only chains them in the order the
reference's graph does (pl_modules.py:504-506 `loss.backward()` through models_ortho.py:12-28 and
libs/ortho_conv/layers.py:50-105).  No torch arithmetic on the path."""
from __future__ import annotations

import os

import torch
from torch.autograd import Function

from . import ops


_SAVE_MINV = os.environ.get('FIODE_SAVE_MINV', '1') != '0'     # A/B switch: 0 = the adjoint inverts (I + A) again


class CayleyLinearQ(Function):
    """Q = cayley(alpha * W / ||W||), layers.py:103-104."""
    @staticmethod
    def forward(ctx, weight, alpha):
        if not _SAVE_MINV:
            ctx.save_for_backward(weight, alpha)
            return ops.cayley_linear(weight, alpha)
        Q, kept = ops.cayley_linear(weight, alpha, save_ctx=True)
        ctx.save_for_backward(weight, alpha, kept)       # scale, (I + A)^-1, staged matrices: the adjoint starts at dL/dQ
        return Q

    @staticmethod
    def backward(ctx, gQ):
        weight, alpha, *rest = ctx.saved_tensors
        gW, ga = ops.cayley_linear_bwd(weight, alpha, gQ, ctx=rest[0] if rest else None)
        return gW, ga.reshape(alpha.shape)


class CayleyConvQ(Function):
    """Per-frequency Q of a CayleyConv (layers.py:74-82); real (cout, cin) when k = 1."""
    @staticmethod
    def forward(ctx, weight, alpha, n):
        ctx.save_for_backward(weight, alpha)
        ctx.n = n
        return ops.cayley_conv(weight, alpha, n)

    @staticmethod
    def backward(ctx, gQ):
        weight, alpha = ctx.saved_tensors
        gw, ga = ops.cayley_conv_bwd(weight, alpha, ctx.n, gQ)
        return gw, ga.reshape(alpha.shape), None


class CayleySmallBatchQ(Function):
    """Q of several small Cayley layers (``ops.cayley_small_ok``) as ONE node: forward is one launch with a CTA per layer,
    backward -- which autograd runs once every layer's dL/dQ has arrived, i.e. at the end of the step's backward pass -- is one
    launch too.  ``mults``: 1 for a CayleyLinear, sqrt(n (n/2+1)) for a k = 1 CayleyConv; tensors = w0, a0, w1, a1, ..."""
    @staticmethod
    def forward(ctx, mults, *wa):
        ws, als = wa[0::2], wa[1::2]
        Qs, Ms = ops.cayley_small_batch(ws, als, mults)
        ctx.mults, ctx.k = tuple(mults), len(ws)
        ctx.save_for_backward(*ws, *als, *Ms)
        return tuple(Qs)

    @staticmethod
    def backward(ctx, *gQs):
        k = ctx.k
        sv = ctx.saved_tensors
        ws, als, Ms = sv[:k], sv[k:2 * k], sv[2 * k:]
        live = [i for i in range(k) if gQs[i] is not None]
        out = [None] * (2 * k)
        if live:
            gWs, gas = ops.cayley_small_batch_bwd([ws[i] for i in live], [als[i] for i in live], [ctx.mults[i] for i in live],
                                                  [gQs[i] for i in live], [Ms[i] for i in live])
            for j, i in enumerate(live):
                out[2 * i], out[2 * i + 1] = gWs[j], gas[j].reshape(als[i].shape)
        return (None, *out)


class Linear(Function):
    """F.linear(X, Q, b), layers.py:105."""
    @staticmethod
    def forward(ctx, X, Q, bias):
        ctx.save_for_backward(X, Q)
        ctx.has_bias = bias is not None
        return ops.linear_fwd(X, Q, bias, False)

    @staticmethod
    def backward(ctx, gY):
        X, Q = ctx.saved_tensors
        gX, gQ, gb = ops.linear_bwd(X, Q, gY, need_gx=ctx.needs_input_grad[0], need_gb=ctx.has_bias)
        return gX, gQ, gb


class Conv1x1S2d(Function):
    """Emulated stride-2 CayleyConv: space-to-depth + real 1x1 mixing (layers.py:24-48)."""
    @staticmethod
    def forward(ctx, x, Q, bias):
        ctx.save_for_backward(x, Q)
        ctx.has_bias = bias is not None
        return ops.conv1x1_s2d_fwd(x, Q, bias, False)

    @staticmethod
    def backward(ctx, gy):
        x, Q = ctx.saved_tensors
        gx, gQ, gb = ops.conv1x1_s2d_bwd(x, Q, gy, need_gx=ctx.needs_input_grad[0], need_gb=ctx.has_bias)
        return gx, gQ, gb


class CayleyConvLayer(Function):
    """y = irfft2(Qf rfft2(x)) + b, layers.py:78-85."""
    @staticmethod
    def forward(ctx, x, Qf, bias):
        ctx.save_for_backward(x, Qf)
        ctx.has_bias = bias is not None
        return ops.cayley_conv_fwd(x, Qf, bias, False)

    @staticmethod
    def backward(ctx, gy):
        x, Qf = ctx.saved_tensors
        gx, gQf, gb = ops.cayley_conv_layer_bwd(x, Qf, gy, need_gx=ctx.needs_input_grad[0], need_gb=ctx.has_bias)
        return gx, gQf, gb


class GroupSortFn(Function):
    """layers.py:127-131."""
    @staticmethod
    def forward(ctx, x):
        ctx.save_for_backward(x)
        return ops.group_sort(x)

    @staticmethod
    def backward(ctx, gy):
        (x,) = ctx.saved_tensors
        return ops.group_sort_bwd(x, gy)


class NormalizeFn(Function):
    """Normalize.forward (models.py:23-26); backward g / std (the PGD attack of validation_step differentiates w.r.t. x)."""
    @staticmethod
    def forward(ctx, x, mu, std):
        ctx.save_for_backward(std)
        return ops.normalize(x, mu, std)

    @staticmethod
    def backward(ctx, g):
        (std,) = ctx.saved_tensors
        return ops.normalize(g, torch.zeros_like(std), std), None, None
