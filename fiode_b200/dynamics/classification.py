"""dynamics/classification.py:31-132 of the reference: OrthoClassDynProjectSimplexLips, the ODE
right-hand side on the probability simplex, on the fused CUDA kernels of libfiode_b200.so."""
from __future__ import annotations

import torch
from torch import nn

from .. import ops
from .._lib import DynWeights, ptr
from ..layers import CayleyLinear, _version
from ..barrier_projection import FastBarrierProjectionNoUpper


class OrthoClassDynProjectSimplexLips(nn.Module):
    def __init__(self, n_hidden=10, activation='ReLU', dropout=0.5, mlp_size=128, kappa=5., kappa_length=3e+4,
                 alpha_1=100., alpha_2=5., sigma_1=0.02, scale_nominal=False, x_dim=10, cayley=True):
        super().__init__()
        assert (n_hidden, mlp_size, x_dim, activation, cayley) == (10, 128, 10, 'ReLU', True), \
            'the fused B200 right-hand side covers the configuration the FI-ODE configs use (10/128/10, ReLU, Cayley)'
        self.dropout = nn.Dropout(dropout)
        self.mlp_size, self.n_hidden = mlp_size, n_hidden
        self.kappa, self.kappa_length = kappa, kappa_length
        self.register_buffer('static_state', None)
        self.alpha_1, self.alpha_2, self.sigma_1 = alpha_1, alpha_2, sigma_1
        self.max_iter, self.tol = 30, 1e-4                     # classification.py:63-64
        self.fast_qp_solver = FastBarrierProjectionNoUpper(max_iter=self.max_iter, tol=self.tol, verbose=False)
        self.scale_nominal = scale_nominal
        self.cayley = cayley
        self.hidden_to_mlp = CayleyLinear(n_hidden, mlp_size, bias=True)
        self.mlp_to_mlp = CayleyLinear(mlp_size, mlp_size)
        self.mlp_to_hidden = CayleyLinear(mlp_size, n_hidden)
        self.U_x = CayleyLinear(x_dim, mlp_size)

    # ---- weights for the C ABI -------------------------------------------------------------------
    def dyn_weights(self):
        """(fiode_dyn_weights struct, tensors it points into).  Reads alpha_1 / alpha_2 / sigma_1 /
        scale_nominal at call time because the training loop mutates them (pl_modules.py:392-393).  Cached per weight
        version and hyper-parameter tuple."""
        lays = (self.hidden_to_mlp, self.U_x, self.mlp_to_mlp, self.mlp_to_hidden)
        key = tuple(_version(l.weight, l.alpha, l.bias) for l in lays) + (
            float(self.alpha_1), float(self.alpha_2), float(self.sigma_1), bool(self.scale_nominal), self.max_iter, self.tol)
        hit = getattr(self, '_dw_cache', None)
        if hit is not None and hit[0] == key:
            return hit[1]
        t = dict(Q1=self.hidden_to_mlp.orthogonal_weight(), b1=self.hidden_to_mlp.bias,
                 Qx=self.U_x.orthogonal_weight(), bx=self.U_x.bias,
                 Q2=self.mlp_to_mlp.orthogonal_weight(), b2=self.mlp_to_mlp.bias,
                 Q3=self.mlp_to_hidden.orthogonal_weight(), b3=self.mlp_to_hidden.bias)
        t = {k: v.detach().float().contiguous() for k, v in t.items()}
        dw = DynWeights(**{k: ptr(v) for k, v in t.items()}, alpha_1=float(self.alpha_1), alpha_2=float(self.alpha_2),
                        sigma_1=float(self.sigma_1), scale_nominal=int(bool(self.scale_nominal)),
                        max_iter=self.max_iter, tol=self.tol)
        dw._keep = t
        self._dw_cache = (key, dw)
        return dw

    # ---- the reference's methods --------------------------------------------------------------------
    def _h_dot_raw(self, h, x):
        assert not (self.training and self.dropout.p > 0), 'dropout inside the RHS is only supported by the FI-loss op'
        return ops.rhs_fwd(self.dyn_weights(), h, x, return_aux=True)[1] if not self.scale_nominal else \
            self._raw_without_scale(h, x)

    def _raw_without_scale(self, h, x):
        sn, self.scale_nominal = self.scale_nominal, False
        try:
            return ops.rhs_fwd(self.dyn_weights(), h, x, return_aux=True)[1]
        finally:
            self.scale_nominal = sn

    def eval_dot(self, t, h_tuple, x):
        return self.eval_dot_light(h_tuple[0], x)

    def eval_dot_light(self, h, x):
        assert not (self.training and self.dropout.p > 0), 'dropout inside the RHS is only supported by the FI-loss op'
        return ops.rhs_fwd(self.dyn_weights(), h, x)

    def ode_forward(self, t, h_tuple):
        assert self.static_state is not None, '[ERROR] You forgot to set static state before calling forward.'
        return self.eval_dot(t, h_tuple, self.static_state)


    # ---- interval step after the CROWN bound (CrownOrthoClassDynProjectSimplexLips, classification.py:175-242) ----
    def ibp_sigmoid(self, f_lb, f_ub, h_lb, h_ub):
        """classification.py:175-181.  The reference passes (eta - eps, eta + eps); the box is recovered from them."""
        h = (h_lb + h_ub) * 0.5
        eps = float(((h_ub - h_lb) * 0.5).flatten()[0])
        return ops.ibp_sigmoid(f_lb, f_ub, h, eps, self.alpha_1, self.alpha_2, self.sigma_1)

    def ibp_cbf_qp(self, h, eps, lb, ub, upper=False):
        """classification.py:208-242 with upper=False (the call at certify_crown.py:143)."""
        assert not upper, 'upper=True needs the three-argument solver the reference class does not construct either'
        return ops.ibp_cbf_qp(h, eps, lb, ub, self.alpha_1, self.sigma_1, self.max_iter, self.tol)


class CrownOrthoClassDynProjectSimplexLips(OrthoClassDynProjectSimplexLips):
    """Name the CROWN certification config instantiates (configs/certify); the plain nn.Linear re-parametrisation the
    reference needs for auto_LiRPA (convert_cayley, classification.py:281-295) is not needed here: the Cayley matrices
    are cached, and CROWN itself stays on the reference."""
