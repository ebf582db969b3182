"""oracle/ref_harness.py -- TEST INFRASTRUCTURE ONLY (build container only).

Imports the reference's OWN Python modules from a read-only checkout so that the oracle
restatement can be pinned against the real code and golden vectors can be generated
(``tests/golden/make_golden.py``).  The reference checkout does not exist on the GPU box;
nothing that runs there may call this module (``available()`` gates it).

Recipe: path priority root > libs/core > libs/ortho_conv; mock the
absent packages that are unused on the path; real-class stubs for ``lconvnet`` and
``pytorch_lightning``; a module named ``torchdiffeq`` exporting the restated dopri5
(``oracle/dopri5.py``) because torchdiffeq==0.2.2 is not installable here.
"""
from __future__ import annotations

import importlib.abc
import importlib.machinery
import os
import sys
import types
from unittest import mock

REF = os.environ.get('FIODE_REFERENCE', '/root/reference')

_MOCKED = ('matplotlib', 'mpl_toolkits', 'plotly', 'advertorch', 'autoattack', 'torchattacks',
           'cvxpy', 'keras', 'hydra', 'omegaconf', 'auto_LiRPA', 'wandb', 'tqdm_', 'pandas_')


def available() -> bool:
    return os.path.isdir(os.path.join(REF, 'dynamics'))


class _MockFinder(importlib.abc.MetaPathFinder, importlib.abc.Loader):
    def find_spec(self, name, path=None, target=None):
        if name.split('.')[0] in _MOCKED:
            return importlib.machinery.ModuleSpec(name, self, is_package=True)
        return None

    def create_module(self, spec):
        m = mock.MagicMock(name=spec.name)
        m.__path__ = []
        m.__spec__ = spec
        m.__name__ = spec.name
        return m

    def exec_module(self, module):
        pass


_installed = False


def install():
    """Idempotently wire sys.path / sys.modules so ``import models`` etc. resolve to the reference."""
    global _installed
    if _installed:
        return
    assert available(), f'reference checkout not found at {REF}'
    import torch
    from torch import nn

    for p in (os.path.join(REF, 'libs', 'ortho_conv'), os.path.join(REF, 'libs', 'core'),
              os.path.join(REF, 'control'), os.path.join(REF, 'robustness'), REF):
        if p in sys.path:
            sys.path.remove(p)
        sys.path.insert(0, p)
    sys.meta_path.insert(0, _MockFinder())

    # lconvnet: classes are subclassed at libs/ortho_conv/layers.py:107-117 -> need real types
    lc = types.ModuleType('lconvnet'); lcl = types.ModuleType('lconvnet.layers')
    for n in ('BCOP', 'RKO', 'OSSN', 'SVCM'):
        setattr(lcl, n, type(n, (nn.Conv2d,), {}))
    lcl.BjorckLinear = type('BjorckLinear', (nn.Linear,), {})
    lc.layers = lcl
    sys.modules['lconvnet'] = lc; sys.modules['lconvnet.layers'] = lcl

    # pytorch_lightning stand-in
    pl = types.ModuleType('pytorch_lightning')

    class _Trainer:
        global_step = 10 ** 9
        current_epoch = 0

    class LightningModule(nn.Module):
        def __init__(self, *a, **k):
            super().__init__()
            self.logged = {}
            self.trainer = _Trainer()
            self.current_epoch = 0

        def log(self, name, value, **kw):
            self.logged[name] = value

        @property
        def device(self):
            return next(self.parameters()).device

    pl.LightningModule = LightningModule
    pl.Callback = object
    pl.Trainer = object
    sys.modules['pytorch_lightning'] = pl
    for sub in ('loggers', 'plugins', 'callbacks'):
        sys.modules[f'pytorch_lightning.{sub}'] = mock.MagicMock(name=f'pl.{sub}')

    # torchdiffeq := restated dopri5
    from . import dopri5 as _d
    td = types.ModuleType('torchdiffeq')
    td.odeint = _d.odeint
    td.odeint_adjoint = _d.odeint
    sys.modules['torchdiffeq'] = td
    _installed = True


# ---- builders over the reference's own classes --------------------------------
MU = (0.485, 0.456, 0.406)
STD = (0.225, 0.225, 0.225)


def build_reference_module(seed=0, scale_nominal=False, alpha_1=100., alpha_2=20., sigma_1=0.02,
                           t_max=1.0, tol=1e-3, h_sample_size=256, kappa=2.0, dropout=0.5):
    """Instantiate the reference LyapunovLearning exactly as the README command would
    (ExpConfig.py:94,134,193,242,292-323), random init under ``seed`` in the fixed order
    backbone -> dynamics, and materialise the lazy conv alphas."""
    install()
    import torch
    import models, lya_cands, pl_modules
    from dynamics.classification import OrthoClassDynProjectSimplexLips
    from dynamics.init_coordinates import UniformInitFun
    from dynamics.output_coordinates import DefaultOutputFun
    from sampling.sampler import CompositeSampler, UniformSimplexSampling, CorrectConeSampling
    from sampling.sampler_schedulers import CompositeSamplerScheduler, LinearScheduler

    torch.manual_seed(seed)
    param_map = models.make_ortho_KWLarge_Concat(3, 10, mu=MU, std=STD, out_dim=10, act='GroupSort')
    dyn = OrthoClassDynProjectSimplexLips(n_hidden=10, activation='ReLU', mlp_size=128, alpha_1=alpha_1,
                                          alpha_2=alpha_2, sigma_1=sigma_1, x_dim=10, cayley=True,
                                          dropout=dropout, scale_nominal=scale_nominal, kappa=kappa,
                                          kappa_length=0)
    init = UniformInitFun((10,), param_map=param_map)
    sampler = CompositeSampler((10,), [UniformSimplexSampling((10,)), CorrectConeSampling((10,))])
    sched = CompositeSamplerScheduler([
        LinearScheduler(rate=-0.02, bias=1., clamp='min', clamp_val=0.02, start=10),
        LinearScheduler(rate=0.02, bias=0., clamp='max', clamp_val=0.98, start=10)], [1., 1.])
    mod = pl_modules.LyapunovLearning(
        order=1, h_sample_size=h_sample_size, h_dist_lim=15, lya_cand=lya_cands.DecisionBoundary(on_simplex=True),
        sampler=sampler, sampler_scheduler=sched, dynamics=dyn, output=DefaultOutputFun(), init_fun=init,
        t_max=t_max, train_ode_tol=tol, val_ode_tol=tol, simplex=True, act='relu', n_input=3072, n_output=10,
        adv_train=False, val_adv=False)
    mod.eval()
    with torch.no_grad():   # one warm-up forward creates CayleyConv.alpha (layers.py:80-81)
        param_map(torch.zeros(1, 3, 32, 32))
    return mod
