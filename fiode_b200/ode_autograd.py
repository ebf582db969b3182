"""Differentiable dopri5 solve: ``loss.backward()`` through ``IVP.forward`` like the reference's
``use_adjoint=False`` path (models.py:235-241; train_ode pl_modules.py:490-500, PGD :189-199).

torchdiffeq back-propagates through the arithmetic of every ACCEPTED Runge-Kutta step with the step sizes as constants
(its controller and initial-step heuristic run under ``torch.no_grad``); rejected attempts leave no trace in the graph.
Here the forward is the device-resident solver (one kernel, no graph); the backward
  1. reads the attempt log the solver kept (step sizes + error ratios, ``fiode_solve_stats``),
  2. replays the accepted steps with ``fiode_rhs_fwd_train`` (activations kept), including the quartic dense-output
     interpolant the requested times are read from,
  3. walks the stages backwards with ``fiode_rhs_bwd_data`` (one VJP per stage; FSAL links stage 7 of a step to stage 1
     of the next) and forms every weight gradient at the end from the stacked stage deltas (``fiode_rhs_bwd_params``).
All vector arithmetic is ``fiode_lincomb``; torch only owns the buffers."""
from __future__ import annotations

import torch
from torch.autograd import Function

from . import ops

# Dormand-Prince tableau (torchdiffeq dopri5.py; same constants as csrc/common.cuh)
BETA = [[1 / 5], [3 / 40, 9 / 40], [44 / 45, -56 / 15, 32 / 9], [19372 / 6561, -25360 / 2187, 64448 / 6561, -212 / 729],
        [9017 / 3168, -355 / 33, 46732 / 5247, 49 / 176, -5103 / 18656],
        [35 / 384, 0., 500 / 1113, 125 / 192, -2187 / 6784, 11 / 84]]
C_MID = [6025192743 / 30085553152 / 2, 0., 51252292925 / 65400821598 / 2, -2691868925 / 45128329728 / 2,
         187940372067 / 1594534317056 / 2, -1776094331 / 19743644256 / 2, 11237099 / 235043384 / 2]
LOG_CAPACITY = 1024    # first size of the device attempt log; a solve with more attempts is re-run with a log that fits


def _f32c(v):
    import numpy as np
    return float(np.float32(v))


def _new(like):
    return torch.zeros_like(like)


def _axpy(dst, c, x):
    return ops.lincomb(dst, [(1.0, dst), (c, x)])


def accepted_steps(log, t0):
    """[(t_left, dt)] of the accepted attempts of an attempt log [(dt, ratio)], time in float64 like the device controller."""
    steps, t = [], float(t0)
    for dt, ratio in log:
        if ratio <= 1.0:
            steps.append((t, float(dt)))
            t = t + float(dt)
    return steps


def solve_with_log(dw, h0, s, t, rtol, atol, opts):
    """Forward solve that keeps EVERY step attempt: the log lives in a caller-owned device buffer
    (fiode_dopri5_simplex_logged); if a stiff solve outgrows it the solve -- deterministic -- is repeated with a log that
    fits.  Trained dynamics need hundreds of attempts (profiles/: 272 for stress B at tol 1e-3)."""
    cap = int(opts.get('log_capacity', LOG_CAPACITY))
    while True:
        res = ops.dopri5_simplex(dw, h0, s, t, rtol, atol, max_num_steps=opts.get('max_num_steps', 2 ** 31 - 1),
                                 rows_per_cta=opts.get('rows_per_cta', 0), log_capacity=cap)
        res.raise_for_status()
        log = res.attempt_log()
        if log is not None:
            return res, log
        st = res.stats()
        cap = st.n_accept + st.n_reject


class OdeSolve(Function):
    """y(t_1..t_k) = odeint(f(., s), h0, t); differentiable in h0, s, the four orthogonal matrices and the biases."""

    @staticmethod
    def forward(ctx, dyn, h0, s, t, rtol, atol, opts, Q1, Qx, Q2, Q3, b1, bx, b2, b3):
        dw = dyn.dyn_weights()
        res, log = solve_with_log(dw, h0, s, t, rtol, atol, opts)
        ctx.dyn, ctx.log, ctx.t = dyn, log, [float(x) for x in t.detach().cpu().double()]
        ctx.stats = res.stats()
        ctx.save_for_backward(h0.detach(), s.detach())
        return res.out

    @staticmethod
    def backward(ctx, g_out):
        h0, s = ctx.saved_tensors
        dyn, t = ctx.dyn, ctx.t
        dw = dyn.dyn_weights()
        QT = ops.dyn_transposes(dw, h0.device)
        steps = accepted_steps(ctx.log, t[0])
        g_out = g_out.contiguous().float()
        # ---- replay ----------------------------------------------------------------------------------------
        y = h0.contiguous().float().clone()
        ev0 = ops.rhs_fwd_train(dw, y, s)
        k_first = ev0
        rec = []                                  # per step: dict(y0, dt, dtf, evs[6], ys[6])
        for (tl, dt) in steps:
            dtf = _f32c(dt)
            ks = [k_first.f]
            evs, y0 = [], y
            for st in range(6):
                ys = ops.lincomb(torch.empty_like(y0), [(1.0, y0)] + [(_f32c(_f32c(BETA[st][j]) * dtf), ks[j]) for j in range(st + 1)])
                ev = ops.rhs_fwd_train(dw, ys, s)
                evs.append(ev)
                ks.append(ev.f)
            rec.append(dict(tl=tl, dt=dt, dtf=dtf, y0=y0, evs=evs, k_first=k_first))
            y, k_first = evs[5].h, evs[5]          # y1 = the point stage 7 was evaluated at; FSAL
        # ---- which step does each requested time read its interpolant from (the first whose right edge reaches it) --
        reads = [[] for _ in rec]
        g_y0_direct = _new(h0)
        _axpy(g_y0_direct, 1.0, g_out[0])                       # solution[0] = y0
        for k in range(1, len(t)):
            n = next(i for i, r in enumerate(rec) if r['tl'] + r['dt'] >= t[k])
            reads[n].append((k, (t[k] - rec[n]['tl']) / rec[n]['dt']))
        # ---- backward -----------------------------------------------------------------------------------------
        a_y = _new(h0)                            # adjoint of the state at the right edge of the current step
        a_k_last = _new(h0)                       # adjoint of stage 7 (= stage 1 of the following step)
        g_s = _new(s.float())
        done = [ev0]
        for n in range(len(rec) - 1, -1, -1):
            r = rec[n]
            dtf = r['dtf']
            a_k = [_new(h0) for _ in range(7)]
            _axpy(a_k[6], 1.0, a_k_last)
            a_y0 = _new(h0)
            a_y1 = a_y
            for (k, x) in reads[n]:
                g = g_out[k]
                p1, p2, p3, p4 = x, x * x, x ** 3, x ** 4
                w_ym = 16 * p4 - 32 * p3 + 16 * p2
                _axpy(a_y0, (-8 * p4 + 18 * p3 - 11 * p2 + 1) + w_ym, g)
                _axpy(a_y1, -8 * p4 + 14 * p3 - 5 * p2, g)
                for j in range(7):
                    if C_MID[j] != 0.:
                        _axpy(a_k[j], w_ym * _f32c(dtf * _f32c(C_MID[j])), g)
                _axpy(a_k[0], dtf * (-2 * p4 + 5 * p3 - 4 * p2 + p1), g)
                _axpy(a_k[6], dtf * (2 * p4 - 3 * p3 + p2), g)
            for st in range(5, -1, -1):           # stage st+2 = f(ys_st), ys_st = y0 + dt sum_j beta[st][j] k_j
                g_h, g_sv = ops.rhs_bwd_data(dw, r['evs'][st], a_k[st + 1], QT)
                done.append(r['evs'][st])
                _axpy(g_s, 1.0, g_sv)
                if st == 5:
                    _axpy(g_h, 1.0, a_y1)         # ys_5 is y1 itself
                _axpy(a_y0, 1.0, g_h)
                for j in range(st + 1):
                    if BETA[st][j] != 0.:
                        _axpy(a_k[j], _f32c(_f32c(BETA[st][j]) * dtf), g_h)
            a_y, a_k_last = a_y0, a_k[0]
        g_h, g_sv = ops.rhs_bwd_data(dw, ev0, a_k_last, QT)      # stage 1 of the first step: f(h0)
        _axpy(g_s, 1.0, g_sv)
        _axpy(a_y, 1.0, g_h)
        _axpy(a_y, 1.0, g_y0_direct)
        # ---- weight gradients from the stacked stage deltas -----------------------------------------------------------
        cat = lambda name: torch.cat([getattr(e, name) for e in done])
        gp = ops.rhs_bwd_params(cat('h'), s.float().repeat(len(done), 1), cat('z1'), cat('z2'), cat('dft'), cat('d2'), cat('d1'))
        return (None, a_y, g_s, None, None, None, None, gp['Q1'], gp['Qx'], gp['Q2'], gp['Q3'], gp['b1'], gp['b1'].clone(),
                gp['b2'], gp['b3'])
