"""pl_modules.py of the reference, restricted to the hot path: make_solver_params (:24-35),
ODELearning.forward (:322-325) and LyapunovLearning.compute_loss (:390-484).  Optimisers, attacks,
validation and logging plumbing stay with Lightning on the reference side; this class is a plain
nn.Module exposing the same attributes so that it can sit under the reference's LightningModule."""
from __future__ import annotations

import torch
from torch import nn

from . import ops
from .models import IVP

ADAPTIVE_SOLVERS = ['dopri8', 'dopri5', 'bosh3', 'adaptive_heun']


def make_solver_params(solver_name, ode_tol):
    if solver_name in ADAPTIVE_SOLVERS:
        return dict(method=solver_name, rtol=ode_tol, atol=ode_tol)
    raise RuntimeError('[ERROR] Invalid Solver Name')


class _FiLoss(torch.autograd.Function):
    """loss = mean relu(Vdot + kappa V).  Differentiable inputs: the four Cayley matrices, the biases and
    the static feature; the backward is fiode_fi_loss_bwd (hand-written, no autograd graph inside)."""

    @staticmethod
    def forward(ctx, dyn, h, y, S, kappa, masks, s, Q1, Qx, Q2, Q3, b1, bx, b2, b3):
        sv = ops.fi_loss_fwd(dyn.dyn_weights(), h, s, y, S, kappa, *(masks or (None, None)))
        ctx.sv = sv
        dyn.last_fi = sv
        return sv.out4[0].clone()

    @staticmethod
    def backward(ctx, grad_out):
        g = ops.fi_loss_bwd(ctx.sv, 1.0)
        go = grad_out.to(torch.float32)
        sc = lambda t: t * go
        return (None, None, None, None, None, None, sc(g['s']), sc(g['Q1']), sc(g['Qx']), sc(g['Q2']), sc(g['Q3']),
                sc(g['b1']), sc(g['b1']), sc(g['b2']), sc(g['b3']))


class LyapunovLearning(nn.Module):
    def __init__(self, order, h_sample_size, h_dist_lim, lya_cand, sampler, sampler_scheduler, dynamics, output,
                 init_fun, n_input=3072, n_output=10, t_max=1.0, train_ode_solver='dopri5', train_ode_tol=1e-6,
                 val_ode_solver='dopri5', val_ode_tol=1e-6, act='relu', epoch_off_scale=10, **unused):
        super().__init__()
        assert order == 1, 'order-1 FI loss (the README configuration)'
        self.order, self.h_sample_size, self.h_dist_lim = order, h_sample_size, h_dist_lim
        self.lya_cand, self.sampler, self.sampler_scheduler = lya_cand, sampler, sampler_scheduler
        self.t_max, self.act, self.epoch_off_scale = t_max, act, epoch_off_scale
        self.train_ode_solver, self.train_ode_tol = train_ode_solver, train_ode_tol
        self.val_ode_solver, self.val_ode_tol = val_ode_solver, val_ode_tol
        self.use_adjoint = False
        self.keep_graph_tensors = False        # tests: expose dL/dQ and dL/ds of the last step (layer.Q.grad, last_static_state.grad)
        self.current_epoch, self.global_step = 0, 10 ** 9
        self.logged = {}
        self.model = IVP(n_input=n_input, n_output=n_output, init_coordinates=init_fun, ts=torch.linspace(0, t_max, 2),
                         ode_tol=train_ode_tol, dyn_fun=dynamics, output_fun=output)
        object.__setattr__(lya_cand, '_dyn_for_v', dynamics)

    def log(self, name, value, **kw):
        self.logged[name] = value

    @property
    def train_solver_params(self):
        return make_solver_params(self.train_ode_solver, self.train_ode_tol)

    @property
    def val_solver_params(self):
        return make_solver_params(self.val_ode_solver, self.val_ode_tol)

    def forward(self, x, t_steps=2, return_traj=False):
        return self.model(x, ts=torch.linspace(0., self.t_max, t_steps, device=x.device),
                          int_params=self.val_solver_params, use_adjoint=self.use_adjoint, return_traj=return_traj)

    def current_kappa(self):
        dyn = self.model.dyn_fun                                  # pl_modules.py:447-450
        if self.global_step < dyn.kappa_length:
            return self.global_step / dyn.kappa_length * dyn.kappa
        return dyn.kappa

    def compute_loss(self, x, y, batch_size, act='identity', variates=None, masks=None, static_state=None):
        assert act == 'relu', "act='relu' (README configuration)"
        dyn = self.model.dyn_fun
        if self.current_epoch == self.epoch_off_scale:            # pl_modules.py:392-393
            dyn.scale_nominal = False
        if static_state is None:
            static_state, _ = self.model.init_coordinates(x, dyn)
        S = self.h_sample_size
        mixing = self.sampler_scheduler.get_mixer_coefficients(self.current_epoch)
        for i, m in enumerate(mixing):
            self.log(f'mixing_weight_{i}', m)
        _, h_sample_in = self.sampler(x, y, self, S, batch_size, list(mixing), variates=variates)
        kappa = self.current_kappa()
        self.log('kappa', kappa)
        lay = (dyn.hidden_to_mlp, dyn.U_x, dyn.mlp_to_mlp, dyn.mlp_to_hidden)
        Qs = []
        for l in lay:
            if dyn.training and torch.is_grad_enabled():
                # train mode: Q is a graph node, dL/dQ continues through fiode_cayley_linear_bwd to weight / alpha
                q = l.orthogonal_weight_autograd()
                if self.keep_graph_tensors:                       # dL/dQ observable in layer.Q.grad, as in eval mode
                    q.retain_grad()
                    l.Q = q
                Qs.append(q)
                continue
            q = l.orthogonal_weight()                             # eval mode: leaf proxies, dL/dQ lands in layer.Q.grad
            if not q.requires_grad:
                q.requires_grad_(True)
            Qs.append(q)
        if not static_state.requires_grad:
            static_state = static_state.detach().requires_grad_(True)
        if masks is None and dyn.training and dyn.dropout.p > 0:
            # classification.py:97-101 dropout after each hidden activation; Bernoulli masks from torch's Philox stream
            p, N = dyn.dropout.p, batch_size * S
            masks = tuple((torch.rand(N, dyn.mlp_size, device=x.device) >= p).float() / (1 - p) for _ in range(2))
        # holding graph-attached tensors on the module keeps last step's autograd graph (and its AccumulateGrad nodes)
        # alive, which breaks CUDA-graph capture of the next step: only on request
        self.last_static_state = static_state if (self.keep_graph_tensors or not static_state.grad_fn) else None
        loss = _FiLoss.apply(dyn, h_sample_in[0], y, S, float(kappa), masks, static_state, *Qs,
                             lay[0].bias, lay[1].bias, lay[2].bias, lay[3].bias)
        out4 = dyn.last_fi.out4
        self.log('mean_active_constraints', out4[2])
        self.log('effective_batch_size', out4[1])
        return loss
