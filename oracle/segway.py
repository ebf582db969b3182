"""oracle/segway.py -- TEST INFRASTRUCTURE ONLY.

CPU restatement of the Segway closed loop and its forward-invariance grid check.

Follows:
  * ``libs/core/core/systems/segway_system.py:19-29``      Segway.forward
  * ``libs/core/core/controllers/nn_controller.py:6-34``   NNController (3->32->32->1, ReLU)
  * ``control/control_utils.py:21-28``    AutoLirpaModel  f(x) = Segway(x, ctrl(x, 0), 0)
  * ``control/control_utils.py:215-234``  LyaQuadratic  V = ||P x||^2, h_dot
  * ``control/control_utils.py:97-106``   SegwaySingleBarrierModel (args swapped, App. B-6)
  * ``control/grid_sampler.py:91-98, 109-117`` reject_sampling, grid_uniform_3d
  * ``control/certify_segway.py:35-59``   level-set shell + pointwise evaluation
  * ``libs/core/core/dynamics/system_dynamics.py:30-91`` step / simulate (ZOH + dopri5 1e-6)
"""
from __future__ import annotations

import math
import numpy as np
import torch
import torch.nn.functional as F

from .dopri5 import odeint_dopri5


def nn_controller(x, ctrl):
    """ctrl: dict with 'model.0.weight', 'model.0.bias', 'model.2.*', 'model.4.*'."""
    z = torch.relu(F.linear(x, ctrl['model.0.weight'], ctrl['model.0.bias']))
    z = torch.relu(F.linear(z, ctrl['model.2.weight'], ctrl['model.2.bias']))
    return F.linear(z, ctrl['model.4.weight'], ctrl['model.4.bias'])


def segway_f(x, u):
    phi, v, phid = x[:, 0:1], x[:, 1:2], x[:, 2:3]
    c, s = torch.cos(phi), torch.sin(phi)
    return torch.cat([
        phid,
        (c * (-1.8 * u + 11.5 * v + 9.8 * s) - 10.9 * u + 68.4 * v - 1.2 * phid * phid * s) / (c - 24.7),
        ((9.3 * u - 58.8 * v) * c + 38.6 * u - 234.5 * v - s * (208.3 + phid * phid * c)) / (c * c - 24.7),
    ], dim=1)


def closed_loop(x, ctrl):
    return segway_f(x, nn_controller(x, ctrl))


def lya_V(x, P):
    Px = F.linear(x, P)
    return (Px * Px).sum(dim=1, keepdim=True)


def barrier_hdot(x, ctrl, P, pt_detached=False):
    """SegwaySingleBarrierModel.forward: h_dot(f, x) = sum(P^T (P f) * x).  ``pt_detached``: the autograd graph of the
    reference -- LyaQuadratic.h_dot refreshes its transposed copy through ``.data`` (control_utils.py:230), so the gradient
    w.r.t. P flows through the inner ``self.P(.)`` only (App. B-6: it is applied to f because the call swaps the arguments)."""
    f = closed_loop(x, ctrl)
    return (F.linear(F.linear(f, P), P.t().detach() if pt_detached else P.t()) * x).sum(dim=1, keepdim=True)


def grid_axes(sizes, r):
    """np.arange(-sizes[i], sizes[i], r[i]) exactly as certify_segway.py:55-57 calls it, i.e. with
    0-dim float32 torch tensors: numpy then takes start, start+step and the length
    ceil((stop-start)/step) from float32 tensor arithmetic, and fills start + i*(next-start) in
    float64 (PyArray_ArangeObj).  Returns float64 axes (cast to float32 by the caller)."""
    axes = []
    for i in range(3):
        size, step = np.float32(sizes[i]), np.float32(r[i])
        start, stop = np.float32(-size), size
        nxt = np.float32(start + step)
        length = int(np.ceil(np.float32(np.float32(stop - start) / step)))
        delta = np.float64(nxt) - np.float64(start)
        axes.append(np.float64(start) + np.arange(length, dtype=np.float64) * delta)
    return axes


def grid_uniform_3d(sizes, r):
    """grid_sampler.py:109-117 (np.meshgrid default 'xy' indexing: axis order (d1, d0, d2))."""
    d0, d1, d2 = grid_axes(sizes, r)
    phi, v, phid = np.meshgrid(d0, d1, d2)
    g = np.concatenate((phi.reshape(-1, 1), v.reshape(-1, 1), phid.reshape(-1, 1)), axis=1)
    return torch.from_numpy(g).float()


def shell_levels(P, level=0.15, r=0.01):
    """certify_segway.py:35-42."""
    smax = torch.linalg.svdvals(P).max().item()
    ub = (math.sqrt(level) + math.sqrt(3) / 2 * r * smax) ** 2
    lb = (math.sqrt(level) - math.sqrt(3) / 2 * r * smax) ** 2
    return lb, ub


@torch.no_grad()
def certify_grid(ctrl, P, sizes, r, level=0.15):
    """certify_segway.py:55-59: returns (eta (M,3), val (M,1), mask (N,))."""
    grid = grid_uniform_3d(sizes, r)
    lb, ub = shell_levels(P, level, float(r[0]))
    val = lya_V(grid, P)
    mask = ((val >= lb) & (val <= ub)).squeeze(1)
    eta = grid[mask]
    return eta, barrier_hdot(eta, ctrl, P), mask


@torch.no_grad()
def simulate(x0, ctrl, ts, rtol=1e-6, atol=1e-6):
    """system_dynamics.py:54-91: zero-order hold, one dopri5 call per sample interval.
    Returns (xs (B, N, 3), us (B, N-1, 1))."""
    xs, us = [x0], []
    for j in range(len(ts) - 1):
        x = xs[j]
        u = nn_controller(x, ctrl)
        us.append(u)
        sol = odeint_dopri5(lambda t, xx: segway_f(xx, u), x, torch.tensor([ts[j], ts[j + 1]]), rtol, atol)
        xs.append(sol[-1])
    return torch.stack(xs, dim=1), torch.stack(us, dim=1)


# ---- adversarial training (control/train_segway.py:119-158) ---------------------------------------------------------
def pgd_linf(x, hdot_fn, eps=0.02, nb_iter=7, eps_iter=None, clip_min=-2 * math.pi, clip_max=2 * math.pi, delta0=None):
    """advertorch 0.2.x ``perturb_iterative`` (ord = inf, minimize = False) as ``LinfPGDAttack(model.forward_adv,
    loss_fn=mean, eps, nb_iter, eps_iter, rand_init, clip_min, clip_max, targeted=False).perturb(x)`` runs it
    (train_segway.py:119-123; advertorch is not in this image: restated from its published source).  ``delta0``: the
    rand_init draw, uniform(-eps, eps) in advertorch; ``hdot_fn(x) -> (N,1)`` differentiable w.r.t. x."""
    eps_iter = 2.5 * eps / nb_iter if eps_iter is None else eps_iter
    delta = torch.zeros_like(x) if delta0 is None else delta0.clone()
    delta = torch.clamp(x + delta, clip_min, clip_max) - x
    for _ in range(nb_iter):
        d = delta.clone().requires_grad_(True)
        loss = hdot_fn(x + d).mean()
        g, = torch.autograd.grad(loss, d)
        delta = delta + eps_iter * g.sign()
        delta = torch.clamp(delta, -eps, eps)
        delta = torch.clamp(x + delta, clip_min, clip_max) - x
    return torch.clamp(x + delta, clip_min, clip_max)


def train_loss(x, ctrl, P, margin=0.01):
    """train_segway.py:147 / :150: sum(relu(model(x) + margin)); differentiable w.r.t. ctrl tensors and P."""
    return torch.relu(barrier_hdot(x, ctrl, P, pt_detached=True) + margin).sum()
