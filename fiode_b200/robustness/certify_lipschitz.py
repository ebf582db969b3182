"""robustness/certify_lipschitz.py of the reference: h_vdot (:37-42) and the certification loop
(:57-163) on the fused grid x image kernel (fiode_certify_grid), sharded over GPUs by image."""
from __future__ import annotations

import math

import torch

from .. import ops, grid as fgrid
from .. import distributed as fdist


def h_vdot(dyn_fun, eta, x_in, label, runner_up):
    """certify_lipschitz.py:37-42 (compatibility seam; the certification loop below does not go through
    it -- the fused kernel evaluates f, the tie mask and the running maxima in one pass)."""
    f = dyn_fun.eval_dot_light(eta, x_in)
    f_label = torch.gather(f, dim=1, index=label[:, None])[:, 0]
    f_wrong = torch.where(runner_up, f, torch.full_like(f, -float('inf'))).max(dim=-1).values
    return -f_label + f_wrong


def certify_constants(dyn, std_min, eps, T, n_class=10):
    """certify_lipschitz.py:65-83."""
    Lfx = (dyn.alpha_1 if dyn.scale_nominal else 1.0) / std_min
    return dict(Lfx=Lfx, kappa=math.sqrt(2) * Lfx * eps, dist=math.sqrt(n_class) / T, eps_grid=1.0 / T)


@torch.no_grad()
def certify(model, images, labels, T=40, eps=0.141, batches=10, gridz=None, images_per_launch=296,
            with_accuracy=True, ode_tol=1e-3, t_max=5.0, margin_tol=3e-4, evaluate=None, backbone=None, std_min=None,
            solve=None):
    """The loop of certify_lipschitz.py:104-163 for ``images`` (n,3,32,32) / ``labels`` (n,).

    Images are sharded round-robin over the ranks of the default process group; each rank holds the
    canonical uint8 grid (413 MB at T=40) and evaluates ``images_per_launch`` images per kernel launch.
    Returns a dict with the reference's counters (count_correct, count_certify, count_certify_larger_T),
    certified_idx, violations_store (list of per-image max violations, what the reference torch.saves) and
    ``near_zero``: images whose margin lies within ``margin_tol`` of 0 (verdict not guaranteed to match a
    run with a different summation order).
    ``std_min``: min of the model's Normalize.std (certify_lipschitz.py:66); read from the model when None.
    ``evaluate(s, labels) -> (m,2)``, ``backbone(x) -> s`` and ``solve(x) -> (m,10)`` can be injected (CPU tests of the
    sharding)."""
    dyn = model.dyn_fun
    n = images.shape[0]
    rank, world = fdist.world()
    mine = fdist.shard_indices(n, rank, world)
    dev = images.device
    if std_min is None:
        std_min = float(torch.as_tensor(model.init_coordinates.param_map[0].std).min())
    c = certify_constants(dyn, std_min, eps, T)
    if evaluate is None:
        if gridz is None:
            gridz = fgrid.boundary_grid_numerators(10, T, device=dev)
        dw = dyn.dyn_weights()
        evaluate = lambda s, lab: ops.certify_grid(dw, gridz, T, s, lab, c['kappa'], batches=batches)
    backbone = backbone or model.init_coordinates.param_map
    viol = torch.zeros(len(mine), 2, dtype=torch.float32, device=dev)
    correct = torch.zeros(len(mine), dtype=torch.float32, device=dev)
    idx_t = torch.as_tensor(mine, dtype=torch.long, device=dev)
    for o in range(0, len(mine), images_per_launch):
        sel = idx_t[o:o + images_per_launch]
        x, lab = images[sel].contiguous(), labels[sel]
        s = backbone(x)
        viol[o:o + sel.numel()] = evaluate(s, lab)
        if with_accuracy:
            # net_out = module(image) (:111-112) is a batch-1 solve per image: all of this launch's images go through ONE
            # forward_batches call as G = m independent solves of B = 1 row (each keeps its own error norm and projection
            # stop rule, exactly like the per-image calls), one status read, no per-image host sync
            if solve is not None:
                y = solve(x)
            else:
                y = model.forward_batches(x[:, None], ts=torch.linspace(0., t_max, 2),
                                          int_params=dict(method='dopri5', rtol=ode_tol, atol=ode_tol))[:, 0]
            correct[o:o + sel.numel()] = (y.argmax(-1) == lab).float()
    if bool(torch.isnan(viol).any()):
        from .._lib import FiodeError
        raise FiodeError('certify: an activation left the fp16x2 range of the tensor-core path (|x| >= 1023); '
                         'rerun with FIODE_MLP_PATH=ffma or fiode_set_mlp_path(1)')
    counts = torch.stack([correct.sum(), (viol[:, 0] < 0).float().sum(), (viol[:, 1] < 0).float().sum()]).double()
    fdist.allreduce_sum_(counts)
    v_all = fdist.gather_by_index(n, mine, viol[:, 0])
    v_T_all = fdist.gather_by_index(n, mine, viol[:, 1])
    certified_idx = (v_all < 0).nonzero()[:, 0].tolist()
    return dict(count_correct=int(counts[0]), count_certify=int(counts[1]), count_certify_larger_T=int(counts[2]),
                clean_acc=float(counts[0]) / n, certify_acc=float(counts[1]) / n, larger_T_certify_acc=float(counts[2]) / n,
                certified_idx=certified_idx, violations_store=v_all.tolist(), violations_larger_T=v_T_all.tolist(),
                near_zero=(v_all.abs() < margin_tol).nonzero()[:, 0].tolist(), kappa=c['kappa'], n_images=n, world=world)
