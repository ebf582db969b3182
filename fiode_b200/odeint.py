"""Drop-in for ``torchdiffeq.odeint(func, y0, t, *, rtol, atol, method='dopri5')`` as the reference
calls it (models.py:236-241): when ``func`` is the bound ``IVP.h_dot`` / ``dyn.ode_forward`` of a
fiode_b200 dynamics module, the whole adaptive solve runs in ONE persistent CUDA kernel with no
host sync per step (fiode_dopri5_simplex).  torchdiffeq 0.2.2 semantics: SURVEY.md Appendix A."""
from __future__ import annotations

import torch

from . import ops


def _resolve_dynamics(func):
    """Find the fiode_b200 dynamics module behind a bound method (IVP.h_dot or dyn.ode_forward)."""
    owner = getattr(func, '__self__', None)
    if owner is None:
        return None
    if hasattr(owner, 'dyn_weights'):
        return owner
    dyn = getattr(owner, 'dyn_fun', None)
    return dyn if dyn is not None and hasattr(dyn, 'dyn_weights') else None


def odeint(func, y0, t, *, rtol=1e-7, atol=1e-9, method=None, options=None, return_result=False, **unused):
    if method not in (None, 'dopri5'):
        raise NotImplementedError(f'fiode_b200.odeint implements dopri5 (got method={method!r})')
    dyn = _resolve_dynamics(func)
    if dyn is None:
        raise NotImplementedError('fiode_b200.odeint fuses the solve with a fiode_b200 dynamics module; pass '
                                  'IVP.h_dot / dyn.ode_forward (use control.segway.simulate for the Segway loop)')
    is_tuple = isinstance(y0, (tuple, list))
    h0 = y0[0] if is_tuple else y0
    assert not is_tuple or len(y0) == 1, 'single-tensor state expected (models.py:219)'
    assert dyn.static_state is not None, '[ERROR] You forgot to set static state before calling forward.'
    opts = options or {}
    res = ops.dopri5_simplex(dyn.dyn_weights(), h0, dyn.static_state, t, rtol, atol,
                             max_num_steps=opts.get('max_num_steps', 2 ** 31 - 1),
                             rows_per_cta=opts.get('rows_per_cta', 0))
    if opts.get('check', True):
        res.raise_for_status()          # the single host sync of the solve
    if return_result:
        return res
    return (res.out,) if is_tuple else res.out
