"""oracle/dynamics.py -- TEST INFRASTRUCTURE ONLY.

CPU restatement of the ODE right-hand side, the barrier QP projection (forward and
its hand-derived backward), the DecisionBoundary candidate and the FI loss.

Follows:
  * ``dynamics/classification.py:96-126``  _h_dot_raw / eval_dot / eval_dot_light
  * ``barrier_projection/barrier_projection.py:217-313`` FastBarrierProjectionNoUpper
  * ``barrier_projection/barrier_projection.py:108-214`` FastBarrierProjection (with upper)
  * ``lya_cands.py:72-94`` DecisionBoundary
  * ``pl_modules.py:390-484`` LyapunovLearning.compute_loss (order 1, act relu)
  * ``models.py:203-242`` IVP.forward, ``dynamics/init_coordinates.py:25-44``
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

from .layers import cayley_linear_Q, backbone_forward
from .dopri5 import odeint_dopri5

DYN_LAYERS = ('hidden_to_mlp', 'mlp_to_mlp', 'mlp_to_hidden', 'U_x')


def dyn_Q(sd, prefix=''):
    """Cayley-transformed matrices of the four dynamics linears (layers.py:103-104)."""
    return {n: cayley_linear_Q(sd[f'{prefix}{n}.weight'], sd[f'{prefix}{n}.alpha']) for n in DYN_LAYERS}


def h_dot_raw(h, s, sd, Q=None, prefix='', masks=None):
    """classification.py:96-102.  ``masks`` = optional (m1, m2) dropout masks already
    scaled by 1/(1-p) (train mode); None = eval mode."""
    Q = Q or dyn_Q(sd, prefix)
    b = lambda n: sd[f'{prefix}{n}.bias']
    z = F.linear(h, Q['hidden_to_mlp'], b('hidden_to_mlp')) + F.linear(s, Q['U_x'], b('U_x'))
    if masks is not None:
        z = z * masks[0]
    z = torch.relu(z)
    z = F.linear(z, Q['mlp_to_mlp'], b('mlp_to_mlp'))
    if masks is not None:
        z = z * masks[1]
    z = torch.relu(z)
    return F.linear(z, Q['mlp_to_hidden'], b('mlp_to_hidden'))


def project_no_upper(lower, nominal, max_iter=30, tol=1e-4):
    """barrier_projection.py:232-269.  Returns (v, mu, n_iter) where n_iter is the number
    of bisection midpoints evaluated (batch-global stop, App. B-3)."""
    mu_hi = (nominal - lower).max(dim=-1, keepdim=True).values.clone()
    mu_lo = nominal.min(dim=-1, keepdim=True).values.clone()
    mu = torch.zeros_like(mu_hi)
    v = torch.zeros_like(nominal)
    n_iter = 0
    for _ in range(max_iter):
        mu = (mu_hi - mu_lo) / 2 + mu_lo
        v = torch.maximum(nominal - mu, lower)
        eps = v.sum(dim=-1, keepdim=True)
        n_iter += 1
        if eps.abs().max() < tol:
            break
        mu_lo = torch.where(eps > 0, mu, mu_lo)
        mu_hi = torch.where(eps < 0, mu, mu_hi)
    return v, mu, n_iter


def project_with_upper(lower, upper, nominal, max_iter=30, tol=1e-4):
    """barrier_projection.py:123-160 (clamp to [lower, upper]; mu_floor from upper)."""
    mu_hi = (nominal - lower).max(dim=-1, keepdim=True).values.clone()
    mu_lo = (nominal - upper).min(dim=-1, keepdim=True).values.clone()
    mu = torch.zeros_like(mu_hi)
    v = torch.zeros_like(nominal)
    n_iter = 0
    for _ in range(max_iter):
        mu = (mu_hi - mu_lo) / 2 + mu_lo
        v = torch.minimum(torch.maximum(nominal - mu, lower), upper)
        eps = v.sum(dim=-1, keepdim=True)
        n_iter += 1
        if eps.abs().max() < tol:
            break
        mu_lo = torch.where(eps > 0, mu, mu_lo)
        mu_hi = torch.where(eps < 0, mu, mu_hi)
    return v, mu, n_iter


def project_no_upper_backward(g, v, mu, lower, nominal):
    """barrier_projection.py:272-311 in closed form.
    Returns (d_lower, d_nominal)."""
    active = (v - nominal + mu) > 0            # lower bound active
    free = ~active
    m = free.sum(dim=-1, keepdim=True).to(g.dtype)
    # m == 0 (every slot on its lower bound): the reference's masks are empty, so the 1/0 is never
    # scattered -> d_lower = g, d_nominal = 0 (barrier_projection.py:296-309)
    mean_free = torch.where(m > 0, (g * free).sum(dim=-1, keepdim=True) / m.clamp(min=1), torch.zeros_like(m))
    d_nom = torch.where(free, g - mean_free, torch.zeros_like(g))
    d_low = torch.where(active, g - mean_free, torch.zeros_like(g))
    return d_low, d_nom


def project_with_upper_backward(g, v, mu, nominal):
    """barrier_projection.py:164-214 in closed form.  lambda = v - nominal + mu; L = {lambda > 0}, U = {lambda < 0},
    F = neither.  The three Jacobians of the reference average over DIFFERENT sets: d_nominal over F, d_lower over the
    complement of L, d_upper over the complement of U (`recp_card(not_lower_active)`, `recp_card(not_upper_active)`).
    Returns (d_lower, d_upper, d_nominal)."""
    lam = v - nominal + mu
    up, lo = lam < 0, lam > 0
    free = ~(up | lo)

    def centred(mask_out, mask_avg):
        m = mask_avg.sum(dim=-1, keepdim=True).to(g.dtype)
        mean = torch.where(m > 0, (g * mask_avg).sum(dim=-1, keepdim=True) / m.clamp(min=1), torch.zeros_like(m))
        return torch.where(mask_out, g - mean, torch.zeros_like(g))
    return centred(lo, ~lo), centred(up, ~up), centred(free, free)


class _ProjFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, lower, nominal):
        v, mu, _ = project_no_upper(lower.detach(), nominal.detach())
        ctx.save_for_backward(v, mu, lower, nominal)
        return v

    @staticmethod
    def backward(ctx, g):
        v, mu, lower, nominal = ctx.saved_tensors
        return project_no_upper_backward(g, v, mu, lower, nominal)


def eval_dot(h, s, sd, cfg, Q=None, prefix='', masks=None, differentiable=False, return_iters=False):
    """classification.py:104-126.  cfg: dict(alpha_1, alpha_2, sigma_1, scale_nominal)."""
    f_t = h_dot_raw(h, s, sd, Q, prefix, masks)
    lower = -cfg['alpha_1'] * (torch.exp(cfg['sigma_1'] * h) - 1)
    upper = cfg['alpha_2'] * (1 - h)
    if cfg['scale_nominal']:
        f_t = (upper - lower) * torch.sigmoid(f_t) + lower
    if differentiable:
        return _ProjFn.apply(lower, f_t)
    # classification.py:63-64 fixes max_iter=30, tol=1e-4; tests may run the bisection to convergence
    # (proj_tol=0) to take its ~1e-5 quantisation of mu out of a solver comparison
    v, _mu, n_iter = project_no_upper(lower, f_t, cfg.get('proj_max_iter', 30), cfg.get('proj_tol', 1e-4))
    return (v, n_iter) if return_iters else v


def decision_boundary_V(h, y, log_mode=False):
    """lya_cands.py:79-94 with on_simplex=True."""
    hy = h.gather(1, y[:, None])[:, 0]
    wrong = h.masked_fill(F.one_hot(y, h.shape[1]).bool(), -float('inf')).max(dim=-1).values
    out = 1 + wrong - hy
    return torch.log(out) if log_mode else out


def ivp_forward(x, sd, cfg, t_max=1.0, tol=1e-3, t_steps=2, return_stats=False,
                prefix='model.', return_traj=False):
    """pl_modules.py:322-325 -> models.py:203-242 with UniformInitFun (init_coordinates.py:38-44)."""
    s = backbone_forward(x, sd, prefix + 'init_coordinates.param_map.')
    h0 = sd[prefix + 'init_coordinates.h0_0'][None, :].repeat(x.shape[0], 1)
    dp = prefix + 'dyn_fun.'
    Q = dyn_Q(sd, dp)
    ts = torch.linspace(0., t_max, t_steps)
    sol, stats = odeint_dopri5(lambda t, hh: eval_dot(hh[0], s, sd, cfg, Q, dp), (h0,), ts, tol, tol,
                               return_stats=True)
    out = sol[-1] if return_traj else sol[-1][-1]
    return (out, s, stats) if return_stats else out


def fi_loss(h, s_rows, y_rows, sd, cfg, kappa, prefix='', Q=None, masks=None):
    """pl_modules.py:435-466 in the closed form of SURVEY App. C-5 (jvp of the
    DecisionBoundary candidate == f[argmax wrong] - f[y]).  Differentiable w.r.t. anything
    ``s_rows``/``sd`` depend on.  Returns (loss, V, Vdot, effective_batch_size)."""
    f = eval_dot(h, s_rows, sd, cfg, Q, prefix, masks, differentiable=True)
    onehot = F.one_hot(y_rows, h.shape[1]).bool()
    w = h.masked_fill(onehot, -float('inf')).argmax(dim=-1)
    hy = h.gather(1, y_rows[:, None])[:, 0]
    V = 1 + h.gather(1, w[:, None])[:, 0] - hy
    Vdot = f.gather(1, w[:, None])[:, 0] - f.gather(1, y_rows[:, None])[:, 0]
    viol = torch.relu(Vdot + kappa * V.detach())
    return viol.mean(), V, Vdot, int((viol > 0).sum())


def mean_active_constraints(h, f, cfg):
    """pl_modules.py:474-483 (note: linear ``lower`` in the log statistic, App. B-8)."""
    lower = -cfg['alpha_1'] * h
    upper = cfg['alpha_2'] * (1 - h)
    return (((f - lower).abs() <= 1e-6) | ((f - upper).abs() <= 1e-6)).float().mean()
