"""barrier_projection/barrier_projection.py of the reference: autograd.Function factories
``FastBarrierProjectionNoUpper(max_iter, tol, verbose) -> apply(lower, nominal)`` (:217-313) and
``FastBarrierProjection(...) -> apply(lower, upper, nominal)`` (:108-214), forward and backward."""
import torch

from . import ops


def FastBarrierProjectionNoUpper(max_iter=100, tol=1e-8, verbose=False):
    class BarrierProjectionFn(torch.autograd.Function):
        @staticmethod
        def forward(ctx, lower, nominal):
            v, mu, _ = ops.barrier_project_fwd(lower, nominal, None, max_iter, tol)
            ctx.save_for_backward(v, mu, nominal.detach())
            return v

        @staticmethod
        def backward(ctx, grad_v):
            v, mu, nominal = ctx.saved_tensors
            d_lower, d_nominal = ops.barrier_project_bwd(grad_v, v, mu, nominal)
            return d_lower, d_nominal

    return BarrierProjectionFn.apply


def FastBarrierProjection(max_iter=100, tol=1e-8, verbose=False):
    class BarrierProjectionFn(torch.autograd.Function):
        @staticmethod
        def forward(ctx, lower, upper, nominal):
            v, mu, _ = ops.barrier_project_fwd(lower, nominal, upper, max_iter, tol)
            ctx.save_for_backward(v, mu, nominal.detach())
            return v

        @staticmethod
        def backward(ctx, grad_v):
            v, mu, nominal = ctx.saved_tensors
            return ops.barrier_project_bwd_upper(grad_v, v, mu, nominal)

    return BarrierProjectionFn.apply
