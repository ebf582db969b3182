"""robustness/certify_crown.py of the reference, restricted to the part that is not auto_LiRPA: the interval step after
the CROWN bound (:139-150).  ``bounds(eta, static_state) -> (lb, ub)`` is the reference's
``bounded_model.compute_bounds(x=(BoundedTensor(eta, ptb), static_state), method='CROWN')`` and stays there."""
from __future__ import annotations

import torch

from .. import ops


def perturbed_vdot(f_lb, f_ub, label, runner_up):
    """certify_crown.py:29-34."""
    return ops.perturbed_vdot(f_lb, f_ub, label, runner_up)


def crown_chunk_violation(dynamics, eta, eps, lb, ub, label, kappa):
    """One chunk of the loop body certify_crown.py:139-150: returns (violation (N,), max violation as a 1-element device
    tensor).  ``eta`` (N,10) grid chunk, ``lb / ub`` the CROWN bounds on the raw MLP output."""
    if dynamics.scale_nominal:
        lb, ub = dynamics.ibp_sigmoid(lb, ub, eta - eps, eta + eps)
    f_lb, f_ub = dynamics.ibp_cbf_qp(eta, eps, lb, ub)
    ind_wrong = ops.runner_up_mask(eta, label, 2 * eps)
    y = torch.full((1,), int(label), dtype=torch.int32, device=eta.device)
    vdot, vmax = ops.perturbed_vdot(f_lb, f_ub, y, ind_wrong, return_max=True)
    return vdot + kappa, vmax + kappa
