"""tests/fixtures.py -- shared helpers for the parity tests and the golden generator.

``synth_state_dict`` builds a state_dict keyed exactly like the reference's Lightning checkpoint
(SURVEY.md section 5).  The small tensors are stored verbatim in ``tests/golden/weights.pt`` (they were
produced by the reference's own constructors under ``torch.manual_seed(0)``); the two large head
matrices (512x4096, 512x512) are regenerated from a recorded CPU generator seed so the fixture
stays small -- a checksum stored next to them guards against RNG drift.
"""
from __future__ import annotations

import math
import os

import torch

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')
PM = 'model.init_coordinates.param_map.'
DYN = 'model.dyn_fun.'
BIG = {PM + '1.model.9.weight': ((512, 4096), 9001), PM + '1.model.11.weight': ((512, 512), 9002)}


def big_weight(name):
    shape, seed = BIG[name]
    g = torch.Generator().manual_seed(seed)
    bound = 1.0 / math.sqrt(shape[1])            # nn.Linear default init range
    return (torch.rand(shape, generator=g) * 2 - 1) * bound


def load_golden(name):
    return torch.load(os.path.join(GOLDEN, name), map_location='cpu', weights_only=True)


def synth_state_dict():
    """Full state_dict (float32 CPU tensors) used by every parity test."""
    blob = load_golden('weights.pt')
    sd = dict(blob['small'])
    for name in BIG:
        w = big_weight(name)
        want = blob['big_checksum'][name]
        got = float(w.double().abs().sum())
        assert abs(got - want) <= 1e-9 * abs(want), f'RNG drift regenerating {name}: {got} vs {want}'
        sd[name] = w
        sd[name[:-len('weight')] + 'alpha'] = w.norm()       # CayleyLinear.alpha init, layers.py:92-93
    return sd


def dyn_cfg(scale_nominal=False, alpha_1=100., alpha_2=20., sigma_1=0.02):
    return dict(alpha_1=alpha_1, alpha_2=alpha_2, sigma_1=sigma_1, scale_nominal=scale_nominal)


def synth_images(B, seed=1234):
    g = torch.Generator().manual_seed(seed)
    return torch.rand(B, 3, 32, 32, generator=g)


def synth_labels(B, seed=1235):
    g = torch.Generator().manual_seed(seed)
    return torch.randint(0, 10, (B,), generator=g)


def simplex_rows(N, seed=7, off_simplex=0.0):
    """N rows on the probability simplex (Exp(1) normalised), optionally jittered off it."""
    g = torch.Generator().manual_seed(seed)
    e = -torch.log1p(-torch.rand(N, 10, generator=g))
    h = e / e.sum(1, keepdim=True)
    if off_simplex:
        h = h + off_simplex * (torch.rand(N, 10, generator=g) - 0.5)
    return h


MU = (0.485, 0.456, 0.406)
STD = (0.225, 0.225, 0.225)


def build_b200_model(sd, device='cuda', t_max=1.0, **dyn_kw):
    """The fiode_b200 module tree assembled like the reference's ``module.model`` (an IVP) and loaded
    from a reference-layout state_dict -- exercises the checkpoint-key compatibility too."""
    import torch
    from fiode_b200 import models
    from fiode_b200.dynamics.classification import OrthoClassDynProjectSimplexLips
    from fiode_b200.dynamics.init_coordinates import UniformInitFun
    kw = dict(n_hidden=10, activation='ReLU', mlp_size=128, alpha_1=100., alpha_2=20., sigma_1=0.02, x_dim=10,
              cayley=True, dropout=0.5, scale_nominal=False, kappa=2.0, kappa_length=0)
    kw.update(dyn_kw)
    param_map = models.make_ortho_KWLarge_Concat(3, 10, mu=MU, std=STD, out_dim=10, act='GroupSort')
    dyn = OrthoClassDynProjectSimplexLips(**kw)
    ivp = models.IVP(3072, 10, dyn, UniformInitFun((10,), param_map=param_map), ts=torch.linspace(0., t_max, 2))
    own = {k[len('model.'):]: v for k, v in sd.items() if k.startswith('model.')}
    missing, unexpected = ivp.load_state_dict(own, strict=False)
    assert not unexpected, unexpected
    assert all('static_state' in m for m in missing), missing
    return ivp.to(device).eval()
