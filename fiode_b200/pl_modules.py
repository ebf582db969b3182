"""pl_modules.py of the reference, restricted to the hot path: make_solver_params (:24-35),
ODELearning.forward (:322-325) and LyapunovLearning.compute_loss (:390-484).  Optimisers, attacks,
validation and logging plumbing stay with Lightning on the reference side; this class is a plain
nn.Module exposing the same attributes so that it can sit under the reference's LightningModule."""
from __future__ import annotations

import torch
from torch import nn

from . import ops
from .layers import finish_prefetch, prefetch_orthogonal
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
    def forward(ctx, dyn, h, y, S, kappa, masks, s, Q1, Qx, Q2, Q3, b1, bx, b2, b3, opt=None):
        sv = ops.fi_loss_fwd(dyn.dyn_weights(), h, s, y, S, kappa, *(masks or (None, None)), opt=opt)
        ctx.sv = sv
        dyn.last_fi = sv
        return sv.out4[0].clone()

    @staticmethod
    def backward(ctx, grad_out):
        g = ops.fi_loss_bwd(ctx.sv, 1.0)
        go = grad_out.to(torch.float32)
        sc = lambda t: t * go
        return (None, None, None, None, None, None, sc(g['s']), sc(g['Q1']), sc(g['Qx']), sc(g['Q2']), sc(g['Q3']),
                sc(g['b1']), sc(g['b1']), sc(g['b2']), sc(g['b3']), None)


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
        # Schedule inputs.  The reference reads self.trainer.global_step (kappa ramp, pl_modules.py:447-450) and
        # self.current_epoch (sampler mixing :428, epoch_off_scale :392-393) from Lightning.  Here: attach the trainer
        # (attach_trainer) or any object with .global_step / .current_epoch, or set the two attributes every step, or pass
        # them to compute_loss.  Without any of these the module refuses to guess (no silent "epoch 0, ramp finished").
        self.trainer = None
        self._current_epoch, self._global_step = None, None
        self.dropout_in_kernel = True          # Philox masks drawn inside the FI-loss kernel (False: torch.rand masks)
        self.dropout_seed = 0x5EEDF10DE
        self._kappa_dev = None                 # device scalar kappa + Philox step counter (created on first use)
        self._drop_step = None
        self.logged = {}
        self.model = IVP(n_input=n_input, n_output=n_output, init_coordinates=init_fun, ts=torch.linspace(0, t_max, 2),
                         ode_tol=train_ode_tol, dyn_fun=dynamics, output_fun=output)

    def log(self, name, value, **kw):
        self.logged[name] = value

    # ---- schedule seam (the reference: LightningModule.trainer / .current_epoch) ----------------------------------
    def attach_trainer(self, trainer):
        """``trainer`` needs ``.global_step`` and ``.current_epoch`` (a Lightning Trainer has both)."""
        self.trainer = trainer
        return self

    @property
    def global_step(self):
        if self._global_step is not None:          # an explicit value wins; set it back to None to follow the trainer
            return self._global_step
        if self.trainer is not None:
            return self.trainer.global_step
        if self._global_step is None:
            raise RuntimeError('LyapunovLearning: no global_step -- attach_trainer(trainer), set .global_step each step, or pass '
                               'global_step= to compute_loss (the kappa warm-up of pl_modules.py:447-450 depends on it)')
        return self._global_step

    @global_step.setter
    def global_step(self, v):
        self._global_step = v

    @property
    def current_epoch(self):
        if self._current_epoch is not None:
            return self._current_epoch
        if self.trainer is not None:
            return self.trainer.current_epoch
        if self._current_epoch is None:
            raise RuntimeError('LyapunovLearning: no current_epoch -- attach_trainer(trainer), set .current_epoch each epoch, or '
                               'pass current_epoch= to compute_loss (sampler mixing and epoch_off_scale depend on it)')
        return self._current_epoch

    @current_epoch.setter
    def current_epoch(self, v):
        self._current_epoch = v

    def _device_scalars(self, device):
        if self._kappa_dev is None or self._kappa_dev.device != device:
            self._kappa_dev = torch.zeros(1, dtype=torch.float32, device=device)
            self._drop_step = torch.zeros(1, dtype=torch.int64, device=device)
        return self._kappa_dev, self._drop_step

    def prepare_step(self, device, global_step=None, current_epoch=None):
        """Host-side part of a step that must NOT be captured in a CUDA graph: writes the scheduled kappa into its device
        scalar and applies the epoch switches.  Returns the key of everything a captured graph has baked in (per-sampler
        counts, scale_nominal, dropout p): re-capture when it changes (``GraphedStep(fn, key_fn=...)`` does).  compute_loss
        calls this itself when it is not being captured."""
        if global_step is not None:
            self._global_step = global_step
        if current_epoch is not None:
            self._current_epoch = current_epoch
        dyn = self.model.dyn_fun
        if self.current_epoch == self.epoch_off_scale:            # pl_modules.py:392-393
            dyn.scale_nominal = False
        kd, _ = self._device_scalars(torch.device(device))
        kappa = self.current_kappa()
        kd.copy_(torch.tensor([kappa], dtype=torch.float32), non_blocking=False)
        self._kappa_host = kappa
        mixing = list(self.sampler_scheduler.get_mixer_coefficients(self.current_epoch))
        counts = tuple(self.sampler._coefficient_to_num_samples(self.h_sample_size, mixing))
        return (counts, bool(dyn.scale_nominal), float(dyn.dropout.p) if dyn.training else 0.0)

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

    def compute_loss(self, x, y, batch_size, act='identity', variates=None, masks=None, static_state=None,
                     global_step=None, current_epoch=None):
        assert act == 'relu', "act='relu' (README configuration)"
        assert not getattr(self.lya_cand, 'log_mode', False) and getattr(self.lya_cand, 'on_simplex', True), \
            'the fused FI loss implements on_simplex=True, log_mode=False (README configuration)'
        dyn = self.model.dyn_fun
        capturing = x.is_cuda and torch.cuda.is_current_stream_capturing()
        if not capturing:     # under capture the caller has run prepare_step() before replay (kappa lives in a device scalar)
            self.prepare_step(x.device, global_step, current_epoch)
        kappa_dev, drop_step = self._device_scalars(x.device)
        if dyn.training and torch.is_grad_enabled() and x.is_cuda:
            # all of the step's Cayley transforms first: the small layers in one launch, the large ones side by side
            init = self.model.init_coordinates
            pf_roots = (dyn, init) if static_state is None and isinstance(init, nn.Module) else (dyn,)
            prefetch_orthogonal(*pf_roots)
        if static_state is None:
            static_state, _ = self.model.init_coordinates(x, dyn)
        S = self.h_sample_size
        mixing = self.sampler_scheduler.get_mixer_coefficients(self.current_epoch)
        for i, m in enumerate(mixing):
            self.log(f'mixing_weight_{i}', m)
        _, h_sample_in = self.sampler(x, y, self, S, batch_size, list(mixing), variates=variates)
        kappa = self._kappa_host if capturing else self.current_kappa()
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
            # eval mode: the reference hands F.linear a detached Q (layers.py:105) -- no gradient reaches weight / alpha.  A
            # detached copy made a leaf lets a caller still observe dL/dQ (layer.Q_grad_proxy.grad); the cached Q itself is
            # never mutated, so nothing accumulates across calls
            q = l.orthogonal_weight().detach()
            if self.keep_graph_tensors and torch.is_grad_enabled():
                q = q.clone().requires_grad_(True)
                l.Q_grad_proxy = q
            Qs.append(q)
        if dyn.training and torch.is_grad_enabled() and x.is_cuda:
            finish_prefetch(*pf_roots)
        if not static_state.requires_grad:
            static_state = static_state.detach().requires_grad_(True)
        drop_p = 0.0
        if masks is None and dyn.training and dyn.dropout.p > 0:
            # classification.py:97-101: dropout on the pre-activations of both hidden layers
            if self.dropout_in_kernel:       # Philox masks drawn inside fiode_fi_loss_fwd, never materialised
                drop_p = float(dyn.dropout.p)
            else:                            # explicit masks from torch's generator (2 x N x 128 floats through HBM)
                p, N = dyn.dropout.p, batch_size * S
                masks = tuple((torch.rand(N, dyn.mlp_size, device=x.device) >= p).float() / (1 - p) for _ in range(2))
        opt = ops.FiOpt(kappa_dev=kappa_dev, dropout_p=drop_p, seed=self.dropout_seed, step_dev=drop_step)
        # holding graph-attached tensors on the module keeps last step's autograd graph (and its AccumulateGrad nodes)
        # alive, which breaks CUDA-graph capture of the next step: only on request
        self.last_static_state = static_state if (self.keep_graph_tensors or not static_state.grad_fn) else None
        loss = _FiLoss.apply(dyn, h_sample_in[0], y, S, float(kappa), masks, static_state, *Qs,
                             lay[0].bias, lay[1].bias, lay[2].bias, lay[3].bias, opt)
        out4 = dyn.last_fi.out4
        self.log('mean_active_constraints', out4[2])
        self.log('effective_batch_size', out4[1])
        return loss
