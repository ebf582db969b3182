"""oracle/dopri5.py -- TEST INFRASTRUCTURE ONLY.

Restatement of torchdiffeq==0.2.2 ``odeint(..., method='dopri5')`` (third-party,
pinned in the reference's ``env.yml:251``; NOT present under the reference tree
nor installable here => "parity unpinned" for the solver arithmetic).

Reference call sites this stands in for:
  * ``models.py:236-241``  odeint(self.h_dot, state_tuple, ts, method, rtol, atol)
  * ``libs/core/core/dynamics/system_dynamics.py:48-52``  odeint(x_dot, x_0, [t0, tf], atol, rtol)

Semantics restated:
  * tuple state is flattened to one 1-D tensor; RMS norm over all elements
    (batch-global);
  * state / stages / error in y0.dtype (fp32); t, dt and the controller in fp64;
  * initial step = Hairer's heuristic evaluated in fp32 (2 RHS calls incl. f0);
  * 7-stage FSAL Dormand-Prince; stage weights formed as fp32(beta)*fp32(dt)
    then contracted over the stage axis; accept iff rms(err/tol) <= 1;
  * dt <- dt * min(10, max(0.9 / ratio^(1/5), dfactor)), dfactor=1 when ratio<1,
    x10 when ratio == 0; no clamping of the step to the output time;
  * output = quartic interpolant evaluated inside the last accepted step.
"""
from __future__ import annotations

import torch

# Dormand-Prince(5,4) tableau, Shampine's dense-output mid weights.
ALPHA = [1 / 5, 3 / 10, 4 / 5, 8 / 9, 1.0, 1.0]
BETA = [
    [1 / 5],
    [3 / 40, 9 / 40],
    [44 / 45, -56 / 15, 32 / 9],
    [19372 / 6561, -25360 / 2187, 64448 / 6561, -212 / 729],
    [9017 / 3168, -355 / 33, 46732 / 5247, 49 / 176, -5103 / 18656],
    [35 / 384, 0.0, 500 / 1113, 125 / 192, -2187 / 6784, 11 / 84],
]
C_SOL = [35 / 384, 0.0, 500 / 1113, 125 / 192, -2187 / 6784, 11 / 84, 0.0]
C_ERR = [
    35 / 384 - 1951 / 21600,
    0.0,
    500 / 1113 - 22642 / 50085,
    125 / 192 - 451 / 720,
    -2187 / 6784 - -12231 / 42400,
    11 / 84 - 649 / 6300,
    -1.0 / 60.0,
]
C_MID = [
    6025192743 / 30085553152 / 2,
    0.0,
    51252292925 / 65400821598 / 2,
    -2691868925 / 45128329728 / 2,
    187940372067 / 1594534317056 / 2,
    -1776094331 / 19743644256 / 2,
    11237099 / 235043384 / 2,
]

SAFETY, IFACTOR, DFACTOR, ORDER = 0.9, 10.0, 0.2, 5


def _rms(x: torch.Tensor) -> torch.Tensor:
    return x.pow(2).mean().sqrt()


class SolveStats:
    """Counters the CUDA solver must reproduce exactly."""

    def __init__(self):
        self.n_accept = 0
        self.n_reject = 0
        self.nfe = 0
        self.dts = []       # every attempted dt (python float, fp64)
        self.ratios = []    # fp32 error ratio of every attempted step
        self.t_final = 0.0  # right edge of the last accepted step

    def as_dict(self):
        return dict(n_accept=self.n_accept, n_reject=self.n_reject, nfe=self.nfe,
                    dts=list(self.dts), ratios=list(self.ratios), t_final=self.t_final)


def _initial_step(func, t0, y0, f0, rtol, atol, stats):
    # torchdiffeq misc._select_initial_step with order argument 4 (= solver order - 1);
    # everything here is y0.dtype (fp32) 0-dim tensor arithmetic.
    dtype = y0.dtype
    t0c = t0.to(dtype)
    scale = atol + y0.abs() * rtol            # fp64 0-dim (x) fp32 tensor -> fp32
    d0 = _rms(y0 / scale)
    d1 = _rms(f0 / scale)
    if d0 < 1e-5 or d1 < 1e-5:
        h0 = torch.tensor(1e-6, dtype=dtype)
    else:
        h0 = 0.01 * d0 / d1
    y1 = y0 + h0 * f0
    f1 = func(t0c + h0, y1)
    stats.nfe += 1
    d2 = _rms((f1 - f0) / scale) / h0
    if d1 <= 1e-15 and d2 <= 1e-15:
        h1 = torch.max(torch.tensor(1e-6, dtype=dtype), h0 * 1e-3)
    else:
        h1 = (0.01 / max(d1, d2)) ** (1.0 / float(ORDER))   # order arg 4 -> exponent 1/(4+1)
    return torch.min(100 * h0, h1).to(torch.float64)


def _rk_step(func, y0, f0, t0, dt, stats):
    dtype = y0.dtype
    t0c, dtc, t1c = t0.to(dtype), dt.to(dtype), (t0 + dt).to(dtype)
    k = torch.empty(*f0.shape, 7, dtype=dtype)
    k[..., 0] = f0
    yi = y0
    for i, (a_i, b_i) in enumerate(zip(ALPHA, BETA)):
        ti = t1c if a_i == 1.0 else t0c + a_i * dtc
        w = torch.tensor(b_i, dtype=dtype) * dtc             # fp32(beta) * fp32(dt)
        yi = y0 + k[..., : i + 1].matmul(w).view_as(f0)
        k[..., i + 1] = func(ti, yi)
        stats.nfe += 1
    y1 = yi                                                  # FSAL: c_sol == last beta row
    f1 = k[..., 6]
    err = k.matmul(dtc * torch.tensor(C_ERR, dtype=dtype))
    return y1, f1, err, k


def _interp_fit(y0, y1, k, dt):
    dtc = dt.to(y0.dtype)
    y_mid = y0 + k.matmul(dtc * torch.tensor(C_MID, dtype=y0.dtype)).view_as(y0)
    f0, f1 = k[..., 0], k[..., 6]
    a = 2 * dtc * (f1 - f0) - 8 * (y1 + y0) + 16 * y_mid
    b = dtc * (5 * f0 - 3 * f1) + 18 * y0 + 14 * y1 - 32 * y_mid
    c = dtc * (f1 - 4 * f0) - 11 * y0 - 5 * y1 + 16 * y_mid
    d = dtc * f0
    return [y0, d, c, b, a]


def _interp_eval(coef, t0, t1, t):
    x = ((t - t0) / (t1 - t0)).to(coef[0].dtype)
    total = coef[0] + x * coef[1]
    xp = x
    for c in coef[2:]:
        xp = xp * x
        total = total + xp * c
    return total


def _next_dt(dt, ratio):
    # torchdiffeq misc._optimal_step_size (fp64 controller fed by the fp32 ratio)
    if ratio == 0:
        return dt * IFACTOR
    dfac = 1.0 if ratio < 1 else DFACTOR
    r64 = ratio.to(torch.float64)
    factor = torch.clamp(SAFETY / r64 ** (1.0 / ORDER), min=dfac, max=IFACTOR)
    return dt * factor


def odeint_dopri5(func, y0, t, rtol, atol, max_num_steps=2 ** 31 - 1, return_stats=False, differentiable=False,
                  forced_dts=None):
    """``differentiable=True`` keeps the autograd graph of the accepted stages (torchdiffeq's use_adjoint=False behaviour).
    ``forced_dts`` (test aid, not torchdiffeq): a list of ACCEPTED step sizes to take instead of running the controller --
    step sizes are constants of torchdiffeq's graph, so two solvers fed the same sizes compute the same function of
    (y0, parameters) and their gradients can be compared even where the controller's own sequence is chaotic."""
    with torch.set_grad_enabled(bool(differentiable) and torch.is_grad_enabled()):
        return _odeint_dopri5(func, y0, t, rtol, atol, max_num_steps, return_stats, forced_dts)


def _odeint_dopri5(func, y0, t, rtol, atol, max_num_steps=2 ** 31 - 1, return_stats=False, forced_dts=None):
    """func(t, y) -> dy/dt with y a Tensor or a tuple of Tensors (then func returns a
    Tensor/tuple of the same total size).  Returns (len(t), *y0.shape) (tuple-wrapped
    if y0 was a tuple).  Raises AssertionError exactly where torchdiffeq does."""
    is_tuple = isinstance(y0, (tuple, list))
    if is_tuple:
        shapes = [tuple(v.shape) for v in y0]
        numels = [v.numel() for v in y0]
        flat0 = torch.cat([v.reshape(-1) for v in y0])

        def unflat(v):
            out, o = [], 0
            for s, n in zip(shapes, numels):
                out.append(v[o:o + n].view(s)); o += n
            return tuple(out)

        def f(tt, yy):
            r = func(tt, unflat(yy))
            if isinstance(r, torch.Tensor):
                return r.reshape(-1)
            return torch.cat([v.reshape(-1) for v in r])
    else:
        flat0, f = y0, func
    t = t.to(torch.float64)
    rtol_t = torch.as_tensor(rtol, dtype=torch.float64)
    atol_t = torch.as_tensor(atol, dtype=torch.float64)
    stats = SolveStats()

    y, tcur = flat0, t[0]
    fcur = f(tcur, y)
    stats.nfe += 1
    # torchdiffeq's step-size heuristics run under torch.no_grad (misc._select_initial_step / _optimal_step_size): the
    # step sizes are constants of the autograd graph, gradients flow through the accepted stages only
    if forced_dts is None:
        dt = _initial_step(f, tcur, y.detach(), fcur.detach(), rtol_t, atol_t, stats).detach()
    else:
        forced = [torch.as_tensor(float(d), dtype=torch.float64) for d in forced_dts]
        dt = forced[0]
    t_left = tcur
    coef = [y] * 5
    sol = [flat0]
    for i in range(1, len(t)):
        nsteps = 0
        while t[i] > tcur:
            assert nsteps < max_num_steps, 'max_num_steps exceeded ({}>={})'.format(nsteps, max_num_steps)
            assert tcur + dt > tcur, 'underflow in dt {}'.format(dt.item())
            assert torch.isfinite(y).all(), 'non-finite values in state `y`: {}'.format(y)
            y1, f1, err, k = _rk_step(f, y, fcur, tcur, dt, stats)
            tol = atol_t + rtol_t * torch.max(y.abs(), y1.abs())
            ratio = _rms(err / tol).abs()
            stats.dts.append(float(dt)); stats.ratios.append(float(ratio))
            if ratio <= 1 or forced_dts is not None:
                coef = _interp_fit(y, y1, k, dt)
                t_left, tcur, y, fcur = tcur, tcur + dt, y1, f1
                stats.n_accept += 1
            else:
                stats.n_reject += 1
            if forced_dts is None:
                dt = _next_dt(dt, ratio.detach())
            elif stats.n_accept < len(forced):
                dt = forced[stats.n_accept]
            nsteps += 1
        sol.append(_interp_eval(coef, t_left, tcur, t[i]))
    stats.t_final = float(tcur)
    out = torch.stack(sol, 0)
    if is_tuple:
        res, o = [], 0
        for s, n in zip(shapes, numels):
            res.append(out[:, o:o + n].view(len(t), *s)); o += n
        out = tuple(res)
    return (out, stats) if return_stats else out


def odeint(func, y0, t, *, rtol=1e-7, atol=1e-9, method=None, options=None, **unused):
    """Drop-in for ``torchdiffeq.odeint`` restricted to dopri5 (the only method the
    reference's README commands select; torchdiffeq's default method is dopri5 too)."""
    assert method in (None, 'dopri5'), 'oracle restates dopri5 only'
    return odeint_dopri5(func, y0, t, rtol, atol)
