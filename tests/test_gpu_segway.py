"""Segway closed loop on the GPU against the golden vectors of the reference's own control/ modules."""
import numpy as np
import pytest
import torch

from oracle import segway as oseg
from tests import fixtures as fx

pytestmark = pytest.mark.gpu


def test_segway_grid_matches_reference():
    from fiode_b200.control import segway
    g = fx.load_golden('segway.pt')
    r = [g['r']] * 3
    eta, val, vmax = segway.certify_grid(g['ctrl'], g['P'], g['sizes'].tolist(), r, level=g['level'],
                                         levels=(g['lb'], g['ub']))
    # the shell test compares an fp32 sum of three squares with the level: a point within one ulp of the level can
    # fall either side (at most 2 such points tolerated); every other point must be the same grid point, in grid
    # order, with the same barrier derivative -- compared on the intersection of the two index sets, never skipped
    grid = oseg.grid_uniform_3d(g['sizes'].tolist(), r)
    where = {tuple(row): i for i, row in enumerate(grid.tolist())}
    got_idx = torch.tensor([where[tuple(row)] for row in eta.cpu().tolist()], dtype=torch.long)   # KeyError = not a grid point
    assert bool((got_idx[1:] > got_idx[:-1]).all()), 'compaction must keep grid order'
    want_idx = g['mask_idx'].long()
    common = np.intersect1d(got_idx.numpy(), want_idx.numpy())
    assert (len(got_idx) - len(common)) + (len(want_idx) - len(common)) <= 2, (len(got_idx), len(want_idx), len(common))
    assert len(common) >= g['mask_count'] - 2 and len(common) > 900
    gi = torch.from_numpy(np.searchsorted(got_idx.numpy(), common))
    wi = torch.from_numpy(np.searchsorted(want_idx.numpy(), common))
    torch.testing.assert_close(val.cpu()[gi], g['val'][wi], rtol=1e-5, atol=1e-5)
    want_max = float(g['val'][wi].max())
    if len(common) == len(want_idx) == len(got_idx):
        assert abs(vmax - g['val_max']) <= 1e-5 * max(1.0, abs(g['val_max']))
    else:
        assert vmax >= want_max - 1e-5 * max(1.0, abs(want_max))
    lb2, ub2 = segway.shell_levels(g['P'], g['level'], g['r'])
    assert abs(lb2 - g['lb']) < 1e-6 and abs(ub2 - g['ub']) < 1e-6


def test_segway_grid_full_size_properties():
    """BASELINE config 1: r = 0.01 -> 300 x 53 x 300 = 4.77 M grid points.  Properties: every kept point lies
    in the shell, sharded evaluation equals the one-shot evaluation."""
    from fiode_b200.control import segway
    g = fx.load_golden('segway.pt')
    sizes, r = g['sizes'].tolist(), [0.01] * 3
    eta, val, vmax = segway.certify_grid(g['ctrl'], g['P'], sizes, r, level=g['level'])
    axes = segway.grid_axes(sizes, r)
    assert [len(a) for a in axes] == [53, 300, 300]
    lb, ub = segway.shell_levels(g['P'], g['level'], 0.01)
    V = ((eta.double() @ g['P'].double().T.cuda()) ** 2).sum(1)
    assert bool(((V >= lb - 1e-6) & (V <= ub + 1e-6)).all()) and eta.shape[0] > 1000
    assert abs(float(val.max()) - vmax) < 1e-6
    n = 53 * 300 * 300
    parts = [segway.certify_grid(g['ctrl'], g['P'], sizes, r, level=g['level'], first=f, count=min(n // 3 + 1, n - f))
             for f in range(0, n, n // 3 + 1)]
    assert torch.equal(torch.cat([p[0] for p in parts]), eta)
    assert torch.equal(torch.cat([p[1] for p in parts]), val)
    # spot-check against the oracle on a slice
    want = oseg.barrier_hdot(eta[:4096].cpu(), g['ctrl'], g['P'])
    torch.testing.assert_close(val[:4096].cpu(), want, rtol=1e-5, atol=1e-5)


def test_segway_simulate_matches_reference():
    from fiode_b200.control import segway
    g = fx.load_golden('segway.pt')
    xs, us = segway.simulate(g['ctrl'], g['x0'], g['ts'])
    torch.testing.assert_close(us.cpu(), g['us'], rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(xs.cpu(), g['xs'], rtol=1e-4, atol=1e-5)
    st = segway.simulate.last_stats
    assert st['nfe'] == 2 * (len(g['ts']) - 1) + 6 * (st['n_accept'] + st['n_reject'])


def test_segway_adversarial_training_step_matches_reference():
    """control/train_segway.py:119-158 on the device: barrier values at arbitrary points, the PGD attack's input gradient, the
    whole 7-iteration attack from the same rand_init draw, and loss.backward() to the controller weights and P -- against the
    reference's own modules under autograd (tests/golden/segway_train.pt), then one Adam step through train_step."""
    from fiode_b200.control import segway
    g = fx.load_golden('segway_train.pt')
    dev = torch.device('cuda', 0)
    ctrl = segway.NNController().to(dev)
    ctrl.load_state_dict(g['ctrl'])
    lya = segway.LyaQuadratic(g['P']).to(dev)
    model = segway.SegwaySingleBarrierModel(ctrl, lya)
    x = g['x'].to(dev)
    val = model(x)
    torch.testing.assert_close(val.detach().cpu(), g['val'], rtol=1e-5, atol=2e-5)
    gx = model.input_gradient(x) / x.shape[0]
    torch.testing.assert_close(gx.cpu(), g['grad_x_of_mean'], rtol=2e-4, atol=1e-6)
    xr = x.clone().requires_grad_(True)
    gx2, = torch.autograd.grad(model(xr).mean(), xr)               # the autograd.Function route gives the same numbers
    torch.testing.assert_close(gx2, gx, rtol=1e-6, atol=1e-9)
    x_adv = segway.linf_pgd(model, x, eps=0.02, nb_iter=7, delta0=g['delta0'].to(dev))
    # a sign flip of a ~0 gradient component moves one coordinate by eps_iter: allow a handful, none beyond the eps ball
    diff = (x_adv.cpu() - g['x_adv']).abs()
    assert (diff > 1e-6).float().mean() < 0.002 and float((x_adv - x).abs().max()) <= 0.02 + 1e-6
    loss = torch.relu(model(g['x_adv'].to(dev)) + 0.01).sum()
    assert abs(float(loss.detach()) - g['loss']) < 1e-3 * abs(g['loss'])
    loss.backward()
    for k, want in g['grads'].items():
        got = dict(ctrl.named_parameters())[k].grad.cpu()
        torch.testing.assert_close(got, want, rtol=5e-4, atol=1e-3 * float(want.abs().max()))
    torch.testing.assert_close(lya.P.weight.grad.cpu(), g['grad_P'], rtol=5e-4, atol=1e-3 * float(g['grad_P'].abs().max()))
    # one epoch body: attack + loss + backward + Adam
    opt = torch.optim.Adam([{'params': list(ctrl.parameters())}, {'params': list(lya.P.parameters()), 'lr': 0.02}], lr=0.01)
    before = [p.detach().clone() for p in list(ctrl.parameters()) + list(lya.P.parameters())]
    l0 = float(segway.train_step(model, x, opt, adv_train=True, delta0=g['delta0'].to(dev)))
    assert abs(l0 - g['loss']) < 2e-3 * abs(g['loss'])
    assert all(not torch.equal(a, b.detach()) for a, b in zip(before, list(ctrl.parameters()) + list(lya.P.parameters())))
