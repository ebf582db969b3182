"""Interval step after the CROWN bound on the GPU against golden vectors produced by the reference's own
CrownOrthoClassDynProjectSimplexLips.ibp_sigmoid / ibp_cbf_qp and perturbed_vdot (tests/golden/make_golden.py crown)."""
import pytest
import torch

from tests import fixtures as fx

pytestmark = pytest.mark.gpu


def build_dyn(c):
    from fiode_b200.dynamics.classification import CrownOrthoClassDynProjectSimplexLips
    return CrownOrthoClassDynProjectSimplexLips(n_hidden=10, activation='ReLU', mlp_size=128, alpha_1=c['alpha_1'],
                                                alpha_2=c['alpha_2'], sigma_1=c['sigma_1'], scale_nominal=c['scale_nominal'],
                                                x_dim=10, cayley=True).cuda().eval()


@pytest.mark.parametrize('name', ['a1_100', 'a1_5_sn'])
def test_ibp_cbf_qp_matches_reference(name):
    from fiode_b200 import ops
    from fiode_b200.robustness import certify_crown as cc
    c = fx.load_golden('crown.pt')[name]
    dyn = build_dyn(c)
    eta, lb, ub, eps = c['eta'].cuda(), c['lb'].cuda(), c['ub'].cuda(), c['eps']
    if c['scale_nominal']:
        lb2, ub2 = dyn.ibp_sigmoid(lb, ub, eta - eps, eta + eps)
        torch.testing.assert_close(lb2.cpu(), c['lb_sig'], rtol=1e-5, atol=1e-6)
        torch.testing.assert_close(ub2.cpu(), c['ub_sig'], rtol=1e-5, atol=1e-6)
    # same inputs to the projection as the reference had, so that only the projection arithmetic is compared
    f_lb, f_ub, iters = ops.ibp_cbf_qp(eta, eps, c['lb_sig'].cuda(), c['ub_sig'].cuda(), c['alpha_1'], c['sigma_1'], return_iters=True)
    # the bisection stops at the first midpoint whose batch-max residual is < 1e-4: mu is quantised at (hi - lo) / 2^K
    torch.testing.assert_close(f_lb.cpu(), c['f_lb'], rtol=1e-4, atol=2e-4)
    torch.testing.assert_close(f_ub.cpu(), c['f_ub'], rtol=1e-4, atol=2e-4)
    assert bool((f_lb <= f_ub + 2e-4).all()) and 1 <= int(iters.min()) <= 30
    mask = ops.runner_up_mask(eta, c['label'], 2 * eps)
    assert torch.equal(mask.cpu().bool(), c['ind_wrong'])
    vdot = cc.perturbed_vdot(c['f_lb'].cuda(), c['f_ub'].cuda(), torch.tensor([c['label']]), c['ind_wrong'])
    torch.testing.assert_close(vdot.cpu(), c['vdot'], rtol=0, atol=1e-6)
    viol, vmax = cc.crown_chunk_violation(dyn, eta, eps, lb, ub, c['label'], kappa=0.25)
    torch.testing.assert_close(viol.cpu(), c['vdot'] + 0.25, rtol=1e-4, atol=5e-4)
    assert abs(float(vmax) - float(viol.max())) < 1e-6
