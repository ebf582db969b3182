"""tests/golden/make_golden.py -- regenerate the golden vectors under tests/golden/.

Runs ONLY in the build container: it imports the reference's own Python modules from the
read-only checkout (oracle/ref_harness.py) and freezes their outputs on seeded synthetic inputs.
The one restated piece on the reference side is the solver arithmetic (torchdiffeq 0.2.2 is
not installable here; oracle/dopri5.py stands in for it -- "parity unpinned" for that piece).

    python tests/golden/make_golden.py            # all groups
    python tests/golden/make_golden.py rhs ivp    # selected groups
"""
from __future__ import annotations

import math
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle import ref_harness as rh            # noqa: E402
from oracle import dopri5 as odp                # noqa: E402
from tests import fixtures as fx                # noqa: E402

OUT = fx.GOLDEN


def save(name, obj):
    path = os.path.join(OUT, name)
    torch.save(obj, path)
    print(f'  wrote {name}: {os.path.getsize(path) / 1024:.1f} KiB')


def build_module(scale_nominal=False, alpha_1=100., alpha_2=20., t_max=1.0, tol=1e-3, dropout=0.5,
                 h_sample_size=256, sd=None):
    mod = rh.build_reference_module(seed=0, scale_nominal=scale_nominal, alpha_1=alpha_1, alpha_2=alpha_2,
                                    t_max=t_max, tol=tol, dropout=dropout, h_sample_size=h_sample_size)
    if sd is not None:
        missing, unexpected = mod.load_state_dict(sd, strict=False)
        assert not unexpected, unexpected
        reset_q(mod)
    return mod


def reset_q(mod):
    for m in mod.modules():              # CayleyLinear caches Q in eval mode (layers.py:103-105)
        if hasattr(m, 'Q'):
            m.Q = None


def gen_weights():
    print('[weights]')
    mod = rh.build_reference_module(seed=0)
    with torch.no_grad():
        for name in fx.BIG:
            w = fx.big_weight(name)
            p = dict(mod.named_parameters())
            p[name].copy_(w)
            p[name[:-len('weight')] + 'alpha'].data = w.norm()
    reset_q(mod)
    sd = {k: v.detach().clone() for k, v in mod.state_dict().items() if v is not None}
    small = {k: v for k, v in sd.items() if k not in fx.BIG and k[:-len('alpha')] + 'weight' not in fx.BIG}
    cks = {name: float(sd[name].double().abs().sum()) for name in fx.BIG}
    save('weights.pt', dict(small=small, big_checksum=cks))


def gen_layers():
    """Per-layer outputs of the reference backbone on one image + cayley() known answers."""
    print('[layers]')
    import layers as ref_layers
    sd = fx.synth_state_dict()
    mod = build_module(sd=sd)
    pm = mod.model.init_coordinates.param_map
    x = fx.synth_images(1)
    acts = {}
    with torch.no_grad():
        h = pm[0](x)
        acts['normalize'] = h.clone()
        for i, layer in enumerate(pm[1].model):
            h = layer(h)
            acts[f'model.{i}'] = h.clone()
        acts['fc'] = pm[1].fc(h)
        s8 = pm(fx.synth_images(8))
    g = torch.Generator().manual_seed(11)
    cay = {}
    for name, shape, cplx in (('real_128x10', (128, 10), False), ('real_10x128', (10, 128), False),
                              ('real_128x128', (128, 128), False), ('cplx_5x32x3', (5, 32, 3), True),
                              ('cplx_3x64x32', (3, 64, 32), True), ('cplx_2x32x128', (2, 32, 128), True)):
        W = torch.randn(shape, generator=g)
        if cplx:
            W = torch.complex(W, torch.randn(shape, generator=g))
        W = W / W.norm() * 3.0
        cay[name] = dict(W=W, Q=ref_layers.cayley(W))
    # per-frequency Q of the two k=3 conv layers and the real Q of the k=1 layers (first/last freq)
    save('layers.pt', dict(x=x, acts=acts, s8=s8, cayley=cay))


def gen_rhs():
    """eval_dot_light of the reference dynamics on seeded rows; both scale_nominal settings."""
    print('[rhs]')
    sd = fx.synth_state_dict()
    out = {}
    for sn in (False, True):
        mod = build_module(scale_nominal=sn, sd=sd)
        dyn = mod.model.dyn_fun
        h = fx.simplex_rows(777, seed=7, off_simplex=0.02)
        s_rows = torch.randn(777, 10, generator=torch.Generator().manual_seed(8)) * 0.5
        s_one = s_rows[:1].clone()
        with torch.no_grad():
            f_rows = dyn.eval_dot_light(h, s_rows)
            f_one = dyn.eval_dot_light(h, s_one)
            raw = dyn._h_dot_raw(h, s_rows)
        out[f'sn{int(sn)}'] = dict(h=h, s_rows=s_rows, s_one=s_one, f_rows=f_rows, f_one=f_one, raw=raw)
    # projection known answers, incl. backward (barrier_projection.py:217-313)
    from barrier_projection.barrier_projection import FastBarrierProjectionNoUpper, FastBarrierProjection
    g = torch.Generator().manual_seed(9)
    nominal = torch.randn(300, 10, generator=g) * 3
    lower = -torch.rand(300, 10, generator=g) * 2
    upper = lower + torch.rand(300, 10, generator=g) * 4 + 0.5
    gout = torch.randn(300, 10, generator=g)
    lo_r, no_r = lower.clone().requires_grad_(True), nominal.clone().requires_grad_(True)
    v = FastBarrierProjectionNoUpper(max_iter=30, tol=1e-4, verbose=False)(lo_r, no_r)
    v.backward(gout)
    with torch.no_grad():
        v_up = FastBarrierProjection(max_iter=30, tol=1e-4, verbose=False)(lower, upper, nominal)
    out['proj'] = dict(nominal=nominal, lower=lower, upper=upper, v=v.detach(), gout=gout,
                       d_lower=lo_r.grad.clone(), d_nominal=no_r.grad.clone(), v_upper=v_up)
    save('rhs.pt', out)


def gen_proj_upper():
    """FastBarrierProjection (lower AND upper bound, barrier_projection.py:108-214) forward + backward of the reference on
    seeded rows; the bounds are tight enough that all three sets (lower-active, upper-active, free) occur in most rows."""
    print('[proj_upper]')
    from barrier_projection.barrier_projection import FastBarrierProjection
    g = torch.Generator().manual_seed(19)
    n = 400
    nominal = torch.randn(n, 10, generator=g) * 3
    lower = -torch.rand(n, 10, generator=g) * 2
    upper = lower + torch.rand(n, 10, generator=g) * 1.5 + 0.1
    # feasibility of sum v = 0 inside the box: shift rows whose box excludes it
    bad = (lower.sum(1) > -0.05) | (upper.sum(1) < 0.05)
    lower[bad] -= 1.0
    upper[bad] += 1.0
    gout = torch.randn(n, 10, generator=g)
    lo_r, up_r, no_r = (t.clone().requires_grad_(True) for t in (lower, upper, nominal))
    captured = {}
    import torch.autograd.function as _f
    v = FastBarrierProjection(max_iter=30, tol=1e-4, verbose=False)(lo_r, up_r, no_r)
    # the saved forward state (v, mu): re-derive mu exactly as the forward does (it is not returned)
    from oracle import dynamics as od
    v2, mu, n_iter = od.project_with_upper(lower, upper, nominal, 30, 1e-4)
    assert torch.equal(v2, v.detach()), 'oracle forward must reproduce the reference bit for bit on this CPU'
    v.backward(gout)
    lam = v.detach() - nominal + mu
    print(f'  rows {n}, iterations {n_iter}, lower-active {int((lam > 0).sum())}, upper-active {int((lam < 0).sum())}, '
          f'free {int((lam == 0).sum())}')
    save('proj_upper.pt', dict(nominal=nominal, lower=lower, upper=upper, gout=gout, v=v.detach(), mu=mu, n_iter=n_iter,
                               d_lower=lo_r.grad.clone(), d_upper=up_r.grad.clone(), d_nominal=no_r.grad.clone()))


def _solve_with_stats(mod, x, t_max, tol):
    """The reference's IVP.forward with the restated solver, plus the step log."""
    model = mod.model
    with torch.no_grad():
        static_state, state = model.init_coordinates(x, model.dyn_fun)
        model.dyn_fun.static_state = static_state
        ts = torch.linspace(0., t_max, 2)
        sol, st = odp.odeint_dopri5(model.h_dot, state, ts, tol, tol, return_stats=True)
        y_via_module = mod(x)                       # ODELearning.forward -> IVP.forward (same thing)
    out = sol[-1][-1]
    assert torch.equal(out, y_via_module)
    return out, static_state, st


def gen_ivp():
    print('[ivp]')
    sd = fx.synth_state_dict()
    res = {}
    cases = {
        # BASELINE config 2: CIFAR inference, dopri5 rtol=atol=1e-3, t_max=1, batch 256
        'c2_b256_tol1e-3': dict(B=256, tol=1e-3, t_max=1.0, kw={}),
        'c5_b64_tol1e-5': dict(B=64, tol=1e-5, t_max=1.0, kw={}),
        'c5_b64_tol1e-2': dict(B=64, tol=1e-2, t_max=1.0, kw={}),
        # stress dynamics that exercise the reject path
        'stress_a': dict(B=64, tol=1e-5, t_max=5.0, kw=dict(scale_nominal=True, alpha_2=50.)),
        'stress_b': dict(B=32, tol=1e-3, t_max=5.0, kw=dict(scale_nominal=True, alpha_1=1000., alpha_2=200.)),
    }
    for name, c in cases.items():
        t0 = time.time()
        mod = build_module(t_max=c['t_max'], tol=c['tol'], sd=sd, **c['kw'])
        x = fx.synth_images(c['B'])
        y, s, st = _solve_with_stats(mod, x, c['t_max'], c['tol'])
        # The reference's own conditioning: re-solve with the static feature perturbed at the fp32
        # round-off level (relative 1e-6, what a different summation order in the backbone does).
        # The barrier projection quantises mu at ~1e-5 and early error estimates are far below that, so
        # the step sequence -- and y at tolerance level -- moves; parity tolerances are tied to this.
        sens, sens_dt = 0.0, 0.0
        for ps in (1, 2, 3):
            gp = torch.Generator().manual_seed(5000 + ps)
            s_p = s * (1 + 1e-6 * torch.randn(s.shape, generator=gp))
            with torch.no_grad():
                mod.model.dyn_fun.static_state = s_p
                h0 = mod.model.init_coordinates(x[:1], mod.model.dyn_fun)[1][0][:1].repeat(c['B'], 1)
                sol_p, st_p = odp.odeint_dopri5(mod.model.h_dot, (h0,), torch.linspace(0., c['t_max'], 2),
                                                c['tol'], c['tol'], return_stats=True)
            sens = max(sens, float((sol_p[-1][-1] - y).abs().max()))
            if (st_p.n_accept, st_p.n_reject) == (st.n_accept, st.n_reject):
                n = min(len(st_p.dts), len(st.dts))
                sens_dt = max(sens_dt, max(abs(a / b - 1) for a, b in zip(st_p.dts[:n], st.dts[:n])))
            else:
                sens_dt = float('inf')
        res[name] = dict(B=c['B'], tol=c['tol'], t_max=c['t_max'], kw=c['kw'], y=y, s=s, stats=st.as_dict(),
                         y_sensitivity=sens, dt_sensitivity=sens_dt)
        print(f'    self-sensitivity to 1e-6 relative noise in s: max|dy| = {sens:.3e}, max rel d(dt) = {sens_dt:.3e}')
        print(f'  {name}: acc={st.n_accept} rej={st.n_reject} nfe={st.nfe} ({time.time() - t0:.1f}s)')
    save('ivp.pt', res)


def gen_fi():
    """LyapunovLearning.compute_loss (the reference's own) at a small shape, dropout=0, with the
    sampler's Exp(1) variates captured so that the new sampler can be fed the same draws."""
    print('[fi]')
    sd = fx.synth_state_dict()
    res = {}
    B, S = 16, 32
    x = fx.synth_images(B)
    y = fx.synth_labels(B)
    y[0] = 9                                        # DecisionBoundarySampling needs class 9 present (App. B-5)
    from sampling.sampler import (CompositeSampler, UniformSimplexSampling, CorrectConeSampling,
                                  DecisionBoundarySampling)
    for name, sn, mix, samplers in (
            ('uniform_sn1', True, [1.0, 0.0], None),
            ('mixed_sn0', False, [0.5, 0.5], None),
            ('boundary_sn0', False, [0.5, 0.5], 'db')):
        mod = build_module(scale_nominal=sn, dropout=0.0, h_sample_size=S, sd=sd)
        mod.train()                                  # Cayley recomputed each call, dropout p=0
        if samplers == 'db':
            mod.sampler = CompositeSampler((10,), [DecisionBoundarySampling((10,)), DecisionBoundarySampling((10,))])
        mod.sampler.device_initialize(torch.device('cpu'))
        mod.sampler_scheduler.get_mixer_coefficients = lambda epoch, m=mix: list(m)
        mod.epoch_off_scale = -1                    # do not flip scale_nominal at epoch 0
        torch.manual_seed(4242)
        _, hs = mod.sampler(x, y, mod, S, B, mix)
        h = hs[0].clone()
        # replay the draws in the same order to capture the raw variates
        torch.manual_seed(4242)
        counts = mod.sampler._coefficient_to_num_samples(S, mix)
        exp = torch.distributions.exponential.Exponential(torch.tensor(1.))
        variates = []
        for smp, n in zip(mod.sampler.samplers, counts):
            if n == 0:
                variates.append(None)
            elif isinstance(smp, UniformSimplexSampling):
                variates.append(exp.sample((n, 10)))
            elif isinstance(smp, CorrectConeSampling):
                variates.append(exp.sample((B, n, 10)))
            else:
                variates.append(exp.sample((B, n, 9)))
        torch.manual_seed(4242)
        mod.zero_grad()
        loss = mod.compute_loss(x, y, B, 'relu')
        loss.backward()
        grads = {k: p.grad.clone() for k, p in mod.named_parameters()
                 if p.grad is not None and p.numel() <= 6000}
        gnorm = {k: float(p.grad.norm()) for k, p in mod.named_parameters() if p.grad is not None}
        res[name] = dict(B=B, S=S, x=x, y=y, mix=mix, scale_nominal=sn, h=h, variates=variates,
                         kinds=[type(s).__name__ for s in mod.sampler.samplers], counts=counts,
                         loss=loss.detach().clone(), logged={k: float(v) for k, v in mod.logged.items()},
                         grads=grads, grad_norms=gnorm)
        print(f'  {name}: loss={float(loss):.6f} eff={mod.logged.get("effective_batch_size")}')
    save('fi.pt', res)


def gen_fi_state():
    """Forward state the reference itself saves in compute_loss (pl_modules.py:390-466) and the gradients its own
    loss.backward() produces *at the Cayley matrices* -- so that the CUDA backward (fiode_fi_loss_bwd,
    fiode_barrier_project_bwd) can be fed the reference's bits (v, mu, nominal from BarrierProjectionFn's
    ctx.save_for_backward, barrier_projection.py:270; post-activation z1, z2; V, Vdot returned by jvp, pl_modules.py:411-414)
    and must then reproduce the reference's gradients to round-off, independent of the round-off active set
    (v - nominal + mu > 0, barrier_projection.py:288-289) that separates two forward passes."""
    print('[fi_state]')
    import pl_modules
    sd = fx.synth_state_dict()
    res = {}
    B, S = 16, 32
    x = fx.synth_images(B)
    y = fx.synth_labels(B)
    y[0] = 9
    from sampling.sampler import CompositeSampler, DecisionBoundarySampling
    for name, sn, mix, samplers in (
            ('uniform_sn1', True, [1.0, 0.0], None),
            ('mixed_sn0', False, [0.5, 0.5], None),
            ('boundary_sn0', False, [0.5, 0.5], 'db')):
        mod = build_module(scale_nominal=sn, dropout=0.0, h_sample_size=S, sd=sd)
        mod.train()
        if samplers == 'db':
            mod.sampler = CompositeSampler((10,), [DecisionBoundarySampling((10,)), DecisionBoundarySampling((10,))])
        mod.sampler.device_initialize(torch.device('cpu'))
        mod.sampler_scheduler.get_mixer_coefficients = lambda epoch, m=mix: list(m)
        mod.epoch_off_scale = -1
        dyn = mod.model.dyn_fun
        cap = dict(act=[], Q={}, proj=[])
        orig_qp = dyn.fast_qp_solver

        def qp(lower, nominal, _o=orig_qp, _c=cap):
            v = _o(lower, nominal)
            if v.grad_fn is not None:
                _c['proj'].append(tuple(t.detach().clone() for t in v.grad_fn.saved_tensors))
            return v
        dyn.fast_qp_solver = qp
        orig_jvp = pl_modules.jvp

        def jvp_cap(*a, _c=cap, **k):
            out = orig_jvp(*a, **k)
            _c['V'], _c['Vdot'] = out[0].detach().clone(), out[1].detach().clone()
            return out
        pl_modules.jvp = jvp_cap
        hooks = [dyn.activation.register_forward_hook(lambda m, i, o, _c=cap: _c['act'].append(o.detach().clone()))]
        for ln in ('hidden_to_mlp', 'U_x', 'mlp_to_mlp', 'mlp_to_hidden'):
            def hk(m, i, o, _c=cap, _n=ln):
                if _n not in _c['Q'] and m.Q.requires_grad:
                    m.Q.retain_grad()
                    _c['Q'][_n] = m.Q
            hooks.append(getattr(dyn, ln).register_forward_hook(hk))

        def pm_hook(m, i, o, _c=cap):
            if 's' not in _c:
                o.retain_grad()
                _c['s'] = o
        hooks.append(mod.model.init_coordinates.param_map.register_forward_hook(pm_hook))
        torch.manual_seed(4242)
        _, hs = mod.sampler(x, y, mod, S, B, mix)
        h = hs[0].clone()
        torch.manual_seed(4242)
        mod.zero_grad()
        try:
            loss = mod.compute_loss(x, y, B, 'relu')
            loss.backward()
        finally:
            pl_modules.jvp = orig_jvp
            for hh in hooks:
                hh.remove()
        v, mu, lower, nominal = cap['proj'][0]
        p = dict(mod.named_parameters())
        pre = 'model.dyn_fun.'
        res[name] = dict(
            B=B, S=S, y=y, scale_nominal=sn, kappa=float(mod.logged['kappa']), h=h, s=cap['s'].detach().clone(),
            f=v, mu=mu[:, 0].clone(), ft=nominal, lower=lower, z1=cap['act'][0], z2=cap['act'][1], V=cap['V'], Vdot=cap['Vdot'],
            loss=loss.detach().clone(),
            dQ1=cap['Q']['hidden_to_mlp'].grad.clone(), dQx=cap['Q']['U_x'].grad.clone(),
            dQ2=cap['Q']['mlp_to_mlp'].grad.clone(), dQ3=cap['Q']['mlp_to_hidden'].grad.clone(),
            db1=p[pre + 'hidden_to_mlp.bias'].grad.clone(), dbx=p[pre + 'U_x.bias'].grad.clone(),
            db2=p[pre + 'mlp_to_mlp.bias'].grad.clone(), db3=p[pre + 'mlp_to_hidden.bias'].grad.clone(),
            ds=cap['s'].grad.clone())
        free = ~((v - nominal + mu) > 0)
        print(f'  {name}: loss={float(loss):.6f} rows={v.shape[0]} free entries={int(free.sum())} '
              f'rows with no free entry={int((free.sum(1) == 0).sum())} |dQ2|={float(res[name]["dQ2"].norm()):.4e}')
    # stand-alone projection with the state BarrierProjectionFn saved (adds mu to the 'proj' entry of rhs.pt)
    from barrier_projection.barrier_projection import FastBarrierProjectionNoUpper
    g = torch.Generator().manual_seed(9)
    nominal = torch.randn(300, 10, generator=g) * 3
    lower = -torch.rand(300, 10, generator=g) * 2
    _ = torch.rand(300, 10, generator=g)
    gout = torch.randn(300, 10, generator=g)
    lo_r, no_r = lower.clone().requires_grad_(True), nominal.clone().requires_grad_(True)
    v = FastBarrierProjectionNoUpper(max_iter=30, tol=1e-4, verbose=False)(lo_r, no_r)
    sv_v, sv_mu, _, _ = [t.detach().clone() for t in v.grad_fn.saved_tensors]
    v.backward(gout)
    res['proj'] = dict(nominal=nominal, lower=lower, v=sv_v, mu=sv_mu[:, 0].clone(), gout=gout, d_lower=lo_r.grad.clone(),
                       d_nominal=no_r.grad.clone())
    save('fi_state.pt', res)


def gen_certify():
    """Grid functions of robustness/eval_utils.py and the certify loop body (restated around the
    reference's own h_vdot / eval_dot_light) on the T=10 grid."""
    print('[certify]')
    import eval_utils as eu
    import certify_lipschitz as cl
    sd = fx.synth_state_dict()
    res = {'counts': {(n, T): eu.count_samples_decision_boundary(n, T)
                      for (n, T) in ((10, 40), (10, 20), (10, 10), (4, 7), (3, 6), (5, 12))}}
    g_small = eu.sample_decision_boundary(n=4, T=8)
    res['grid_4_8'] = torch.from_numpy(np.rint(g_small * 8).astype(np.int64))
    grid_cache = {}

    def grids_for(T):
        if T not in grid_cache:
            g0 = eu.sample_decision_boundary(n=10, T=T)
            grid_cache[T] = [eu.get_grid_for_label(g0.copy(), i) for i in range(10)]
            if T == 10:
                res['grid_10_10_sorted'] = torch.from_numpy(np.unique(np.rint(g0 * T).astype(np.uint8), axis=0))
        return grid_cache[T]

    for T, eps, sn, shift in ((10, 0.141, False, 0.0), (10, 0.141, True, 0.0), (20, 0.02, False, 14.0),
                              (20, 0.02, False, 1.5)):
        grids = grids_for(T)
        mod = build_module(scale_nominal=sn, sd=sd)
        dyn = mod.model.dyn_fun
        if shift:                                 # 'planted' variant: push label 3 so some images certify
            with torch.no_grad():
                e = torch.zeros(10); e[3] = 1.0
                dyn.mlp_to_hidden.bias.add_(shift * (e - 0.1))
        M = 6
        x = fx.synth_images(M, seed=1236)
        labels = fx.synth_labels(M, seed=1237)
        labels[:3] = 3
        batches = 10
        std_min = min(mod.model.init_coordinates.param_map[0].std).item()
        Lfx = (dyn.alpha_1 if sn else 1.0) / std_min
        kappa = math.sqrt(2) * Lfx * eps
        dist = math.sqrt(10) / T
        eps_g = 1 / T
        outs = []
        with torch.no_grad():
            s_all = mod.model.init_coordinates(x, dyn)[0]
            for m in range(M):
                label = int(labels[m])
                ss = s_all[m:m + 1]
                G = grids[label]
                ebs = G.shape[0] // batches
                nb = batches + (1 if G.shape[0] % batches else 0)
                v1, v2 = [], []
                for bi in range(nb):
                    eta = G[bi * ebs:(bi + 1) * ebs] if (bi + 1) * ebs < G.shape[0] else G[bi * ebs:]
                    eta_ub = eta.amax(dim=1) + eps_g
                    Lf = math.sqrt(10) * (dyn.sigma_1 * dyn.alpha_1 * torch.exp(dyn.sigma_1 * eta_ub)) + 1
                    perturb = (math.sqrt(2) * Lf) * dist
                    ind = eta == eta.max(dim=-1, keepdim=True).values
                    ind[:, label] = False
                    yv = torch.ones(eta.shape[0], dtype=torch.long) * label
                    hx = cl.h_vdot(dyn, eta, ss, yv, ind)
                    v2.append((hx + kappa).max().item())
                    v1.append((hx + perturb + kappa).max().item())
                outs.append((max(v1), max(v2)))
        res[f'T{T}_sn{int(sn)}_shift{shift}'] = dict(T=T, eps=eps, batches=batches, x=x, labels=labels, s=s_all,
                                               kappa=kappa, shift=shift, scale_nominal=sn,
                                               out=torch.tensor(outs, dtype=torch.float64))
        print(f'  T={T} sn={sn} shift={shift}:', [f'{a:.4f}' for a, _ in outs])
    save('certify.pt', res)


def gen_certify40(stride=1, shift=2.1, M=8, eps=0.141, save_it=True):
    """BASELINE config 4 at full size: the reference's own h_vdot (robustness/certify_lipschitz.py:37-42) driven over
    ALL 41 320 837 rows of grid_40 in the chunks of certify_lipschitz.py:104-121 (batches = 10 -> ten chunks of
    4 132 083 rows + a 7-row tail) for M = 8 images, loop body restated from :122-143.  Rows come from the oracle's
    C enumerator (oracle/grid_enum.c) in canonical order -- the reference's O(N^2) numpy generator would need minutes
    and > 10 GB here; the set is the same (every consumer is a max over rows) and the chunk composition, which fixes the
    batch-global bisection stop, is then the one the CUDA enumerator produces too.  'planted' dynamics: label-3 images
    pushed by shift * (e_3 - 1/10) on mlp_to_hidden.bias so that both verdicts occur."""
    print('[certify40]')
    import certify_lipschitz as cl
    from oracle import certify as oc
    T, batches = 40, 10
    sd = fx.synth_state_dict()
    z = oc.boundary_grid_numerators_c(10, T)
    res = dict(T=T, eps=eps, batches=batches, shift=shift, scale_nominal=False, rows=z.shape[0],
               row_hash=oc.grid_row_hash(z), stride=stride)
    if stride > 1:
        z = z[::stride]
    mod = build_module(scale_nominal=False, sd=sd)
    dyn = mod.model.dyn_fun
    with torch.no_grad():
        e = torch.zeros(10); e[3] = 1.0
        dyn.mlp_to_hidden.bias.add_(shift * (e - 0.1))
    x = fx.synth_images(M, seed=1236)
    labels = fx.synth_labels(M, seed=1237)
    labels[:M // 2] = 3
    std_min = min(mod.model.init_coordinates.param_map[0].std).item()
    kappa = math.sqrt(2) * (1.0 / std_min) * eps
    dist = math.sqrt(10) / T
    eps_g = 1 / T
    n = z.shape[0]
    ebs = n // batches
    nb = batches + (1 if n % batches else 0)
    per_chunk = torch.zeros(M, nb, 2, dtype=torch.float64)
    t0 = time.time()
    with torch.no_grad():
        s_all = mod.model.init_coordinates(x, dyn)[0]
        for bi in range(nb):
            zc = z[bi * ebs:(bi + 1) * ebs] if (bi + 1) * ebs < n else z[bi * ebs:]
            for label in sorted(set(labels.tolist())):
                eta = oc.grid_for_label(zc, T, label)          # eval_utils.py:64-69, float64 z/T -> .float()
                eta_ub = eta.amax(dim=1) + eps_g
                Lf = math.sqrt(10) * (dyn.sigma_1 * dyn.alpha_1 * torch.exp(dyn.sigma_1 * eta_ub)) + 1
                perturb = (math.sqrt(2) * Lf) * dist
                ind = eta == eta.max(dim=-1, keepdim=True).values
                ind[:, label] = False
                yv = torch.ones(eta.shape[0], dtype=torch.long) * label
                for m in range(M):
                    if int(labels[m]) != label:
                        continue
                    hx = cl.h_vdot(dyn, eta, s_all[m:m + 1], yv, ind)
                    per_chunk[m, bi, 1] = (hx + kappa).max().item()
                    per_chunk[m, bi, 0] = (hx + perturb + kappa).max().item()
            print(f'  chunk {bi + 1}/{nb} ({zc.shape[0]} rows) done at {time.time() - t0:.0f} s', flush=True)
    out = per_chunk.max(dim=1).values
    near_zero = [(m, float(out[m, 0])) for m in range(M) if abs(float(out[m, 0])) < 3e-4]
    res.update(x=x, labels=labels, s=s_all, kappa=kappa, out=out, per_chunk=per_chunk, near_zero=near_zero,
               cpu_seconds=time.time() - t0, cpu_threads=torch.get_num_threads())
    print('  out:', [f'{a:.4f}' for a in out[:, 0].tolist()], 'labels', labels.tolist(), 'near_zero', near_zero)
    if save_it:
        save('certify40.pt', res)
    return res


def gen_segway():
    print('[segway]')
    import control_utils as cu
    import grid_sampler as gs
    from core.systems.segway_system import Segway
    from core.controllers.nn_controller import NNController
    torch.set_default_dtype(torch.float32)          # certify_segway.py:26 (core.* imports flip it to float64)
    torch.manual_seed(0)
    system = Segway()
    ctrl = NNController(system, 3, 1, 32)
    with torch.no_grad():                          # tame the random controller so trajectories stay bounded
        ctrl.model[4].weight.mul_(0.05)
    ctrl.eval()
    P = torch.eye(3) + 0.1 * torch.randn(3, 3, generator=torch.Generator().manual_seed(2))
    lya = cu.LyaQuadratic(P.clone(), torch.zeros(1, 3))
    dynamics = cu.AutoLirpaModel(system, ctrl)
    model = cu.SegwaySingleBarrierModel(dynamics, lya)
    smax = torch.linalg.svd(lya.P.weight.data)[1].max().item()
    level, r = 0.15, 0.05
    ub = (math.sqrt(level) + math.sqrt(3) / 2 * r * smax) ** 2
    lb = (math.sqrt(level) - math.sqrt(3) / 2 * r * smax) ** 2
    sizes = torch.tensor([np.pi / 12, 1.5, 1.5])
    rs = torch.ones(3) * r
    grid = gs.grid_uniform_3d(sizes, rs)[0]
    with torch.no_grad():
        eta, mask = gs.reject_sampling(grid, lya, lb, ub, return_mask=True)
        val = model(eta)
        f = dynamics(eta)
        # closed-loop ZOH simulation (system_dynamics.py:54-91) -- uses the restated dopri5
        x0 = torch.tensor([[0.1, 0.2, -0.1], [-0.05, 0.0, 0.3]])
        ts = torch.linspace(0, 0.5, 26)
        import tqdm as _tqdm
        xs, us = system.simulate(x0, ctrl, ts)
    res = dict(ctrl={k: v.clone() for k, v in ctrl.state_dict().items()}, P=P, level=level, r=r,
               sizes=sizes, n_grid=grid.shape[0], mask_count=int(mask.sum()), mask_idx=mask.nonzero()[:, 0].int(),
               val=val, f=f, val_max=float(val.max()), lb=lb, ub=ub, x0=x0, ts=ts, xs=xs, us=us)
    print(f'  grid {grid.shape[0]} pts, shell {int(mask.sum())}, max hdot {float(val.max()):.5f}')
    save('segway.pt', res)


def gen_segway_train():
    """The adversarial-training pieces of control/train_segway.py:119-158 on the reference's own modules: barrier values at
    arbitrary points, the input gradient the PGD attack uses, one PGD run (perturb_iterative restated, fixed rand_init draw)
    and the gradients of sum(relu(model(x_adv) + 0.01)) w.r.t. the controller weights and P."""
    print('[segway_train]')
    import control_utils as cu
    from core.systems.segway_system import Segway
    from core.controllers.nn_controller import NNController
    from oracle import segway as oseg
    torch.set_default_dtype(torch.float32)
    torch.manual_seed(3)
    system = Segway()
    ctrl = NNController(system, 3, 1, 32)
    with torch.no_grad():
        ctrl.model[4].weight.mul_(0.2)
        for m in (ctrl.model[0], ctrl.model[2], ctrl.model[4]):
            m.bias.normal_(0, 0.1)
    P = torch.eye(3) + 0.1 * torch.randn(3, 3, generator=torch.Generator().manual_seed(4))
    lya = cu.LyaQuadratic(P.clone(), torch.zeros(1, 3))
    model = cu.SegwaySingleBarrierModel(cu.AutoLirpaModel(system, ctrl), lya)
    g = torch.Generator().manual_seed(5)
    n = 777
    x = (torch.rand(n, 3, generator=g) * 2 - 1) * torch.tensor([math.pi / 12, 1.5, 1.5])
    delta0 = (torch.rand(n, 3, generator=g) * 2 - 1) * 0.02
    xr = x.clone().requires_grad_(True)
    val = model(xr)
    gx, = torch.autograd.grad(val.mean(), xr)
    x_adv = oseg.pgd_linf(x, lambda z: model.forward_adv(z), eps=0.02, nb_iter=7, delta0=delta0)
    for p in list(ctrl.parameters()) + list(lya.P.parameters()):
        p.grad = None
    loss = torch.relu(model(x_adv) + 0.01).sum()
    loss.backward()
    grads = {k: v.grad.clone() for k, v in ctrl.named_parameters()}
    res = dict(ctrl={k: v.detach().clone() for k, v in ctrl.state_dict().items()}, P=P, x=x, delta0=delta0, val=val.detach(),
               grad_x_of_mean=gx, x_adv=x_adv, loss=float(loss), grads=grads, grad_P=lya.P.weight.grad.clone(),
               active=int((model(x_adv) + 0.01 > 0).sum()))
    print(f'  {n} points, loss {float(loss):.4f}, active {res["active"]}, max |x_adv - x| {float((x_adv - x).abs().max()):.4f}')
    save('segway_train.pt', res)


def gen_crown():
    """Post-CROWN interval step of robustness/certify_crown.py:139-150 with the reference's own functions:
    CrownOrthoClassDynProjectSimplexLips.ibp_sigmoid / ibp_cbf_qp (dynamics/classification.py:175-242) and
    perturbed_vdot (certify_crown.py:29-34).  The CROWN bounds themselves (auto_LiRPA) stay on the reference: lb / ub are
    synthetic intervals around a random nominal."""
    print('[crown]')
    import importlib
    cls = importlib.import_module('dynamics.classification')
    g = torch.Generator().manual_seed(77)
    res = {}
    for name, N, eps, a1, a2, sn, width in [('a1_100', 600, 0.0125, 100., 20., False, 0.3), ('a1_5_sn', 400, 0.025, 5., 5., True, 1.0)]:
        dyn = cls.CrownOrthoClassDynProjectSimplexLips(alpha_1=a1, alpha_2=a2, sigma_1=0.02, scale_nominal=sn)
        e = -torch.log1p(-torch.rand(N, 10, generator=g))
        eta = e / e.sum(1, keepdim=True)
        mid = torch.randn(N, 10, generator=g) * 2.0
        half = torch.rand(N, 10, generator=g) * width
        lb, ub = mid - half, mid + half
        label = 3
        with torch.no_grad():
            lb2, ub2 = (dyn.ibp_sigmoid(lb, ub, eta - eps, eta + eps) if sn else (lb, ub))
            f_lb, f_ub = dyn.ibp_cbf_qp(eta, eps, lb2.clone(), ub2.clone())
            max_wrong = torch.max(eta, dim=-1, keepdim=True).values
            ind_wrong = eta >= (max_wrong - 2 * eps)
            ind_wrong[:, label] = False
            y = torch.ones(N, dtype=torch.long) * label
            f_y = f_lb[list(range(N)), y]
            f_wrong = torch.max(f_ub.masked_fill(~ind_wrong, -float('inf')), dim=-1).values
            vdot = -f_y + f_wrong                                  # perturbed_vdot, certify_crown.py:29-34
        res[name] = dict(eta=eta, eps=eps, alpha_1=a1, alpha_2=a2, sigma_1=0.02, scale_nominal=sn, lb=lb, ub=ub, lb_sig=lb2, ub_sig=ub2,
                         f_lb=f_lb, f_ub=f_ub, label=label, ind_wrong=ind_wrong, vdot=vdot)
        print(f'  {name}: f_lb [{float(f_lb.min()):.3f}, {float(f_lb.max()):.3f}]  vdot max {float(vdot.max()):.4f}')
    save('crown.pt', res)


GROUPS = dict(weights=gen_weights, layers=gen_layers, rhs=gen_rhs, ivp=gen_ivp, fi=gen_fi, fi_state=gen_fi_state,
              certify40=gen_certify40,
              certify=gen_certify, segway=gen_segway, crown=gen_crown, proj_upper=gen_proj_upper, segway_train=gen_segway_train)

if __name__ == '__main__':
    assert rh.available(), 'reference checkout not found: golden vectors can only be regenerated in the build container'
    rh.install()
    torch.set_num_threads(8)
    which = sys.argv[1:] or list(GROUPS)
    for name in which:
        GROUPS[name]()
