"""Segway NN-controller forward invariance on libfiode_b200.so: the pointwise part of
control/certify_segway.py:35-59 (level-set shell, grid, barrier derivative) and the closed-loop
zero-order-hold simulation of libs/core/core/dynamics/system_dynamics.py:54-91.  The CROWN bound and
the matplotlib figures of certify_segway.py stay on the reference (auto_LiRPA / no graphics stack)."""
from __future__ import annotations

import ctypes as C
import math

import numpy as np
import torch

from .._lib import SegwayWeights, lib, check, ptr, stream_ptr


def grid_axes(sizes, r):
    """np.arange(-sizes[i], sizes[i], r[i]) exactly as control/grid_sampler.py:111-113 evaluates it
    when called with 0-dim float32 torch tensors (certify_segway.py:55-57): length and step come
    from float32 arithmetic, values are filled as start + i * (next - start) in float64, and the grid
    is cast to float32 (grid_sampler.py:116).  Host-side scalar logic."""
    axes = []
    for i in range(3):
        size, step = np.float32(sizes[i]), np.float32(r[i])
        start, stop = np.float32(-size), size
        nxt = np.float32(start + step)
        length = int(np.ceil(np.float32(np.float32(stop - start) / step)))
        delta = np.float64(nxt) - np.float64(start)
        axes.append((np.float64(start) + np.arange(length, dtype=np.float64) * delta).astype(np.float32))
    return axes


def shell_levels(P, level=0.15, r=0.01):
    """certify_segway.py:35-42."""
    smax = float(torch.linalg.svdvals(P.detach().cpu().double()).max())
    smax = float(np.float32(smax))
    ub = (math.sqrt(level) + math.sqrt(3) / 2 * r * smax) ** 2
    lb = (math.sqrt(level) - math.sqrt(3) / 2 * r * smax) ** 2
    return lb, ub


def _weights(ctrl_sd, P, device):
    keep = dict(W0=ctrl_sd['model.0.weight'], b0=ctrl_sd['model.0.bias'], W1=ctrl_sd['model.2.weight'],
                b1=ctrl_sd['model.2.bias'], W2=ctrl_sd['model.4.weight'], b2=ctrl_sd['model.4.bias'], P=P)
    keep = {k: v.detach().to(device=device, dtype=torch.float32).contiguous() for k, v in keep.items()}
    assert keep['W0'].shape == (32, 3) and keep['W1'].shape == (32, 32) and keep['W2'].shape == (1, 32), \
        'NNController(system, 3, 1, 32) as in certify_segway.py:45'
    w = SegwayWeights(**{k: ptr(v) for k, v in keep.items()})
    w._keep = keep
    return w


def certify_grid(ctrl_sd, P, sizes, r, level=0.15, device='cuda', levels=None, first=0, count=None):
    """Returns (eta (M,3), val (M,1), max val) for the shell points of grid_uniform_3d(sizes, r)."""
    w = _weights(ctrl_sd, P, device)
    axes = [torch.from_numpy(a).to(device) for a in grid_axes(sizes, r)]
    n0, n1, n2 = (a.numel() for a in axes)
    total = n0 * n1 * n2
    count = total - first if count is None else count
    lb, ub = levels if levels is not None else shell_levels(P, level, float(r[0]))
    eta = torch.empty(count, 3, dtype=torch.float32, device=device)
    val = torch.empty(count, dtype=torch.float32, device=device)
    stats = torch.empty(2, dtype=torch.float32, device=device)
    nb = lib().fiode_segway_grid_workspace_bytes()
    ws = torch.empty(nb, dtype=torch.uint8, device=device)
    with torch.cuda.device(eta.device):
        check(lib().fiode_segway_grid(C.byref(w), ptr(axes[0]), n0, ptr(axes[1]), n1, ptr(axes[2]), n2,
                                      float(np.float32(lb)), float(np.float32(ub)), first, count, count, ptr(eta), ptr(val),
                                      ptr(stats), ptr(ws), nb, stream_ptr()), 'fiode_segway_grid')
    m, vmax = stats.tolist()
    m = int(m)
    return eta[:m], val[:m, None], vmax


def simulate(ctrl_sd, x0, ts, rtol=1e-6, atol=1e-6, device='cuda'):
    """SystemDynamics.simulate(x0, controller, ts) for the Segway + NNController closed loop.
    Returns (xs (B,N,3), us (B,N-1,1))."""
    w = _weights(ctrl_sd, torch.eye(3), device)
    x0 = x0.detach().to(device=device, dtype=torch.float32).contiguous()
    ts = ts.detach().to(device=device, dtype=torch.float32).contiguous()
    B, N = x0.shape[0], ts.numel()
    assert x0.shape[1] == 3
    xs = torch.empty(B, N, 3, dtype=torch.float32, device=device)
    us = torch.empty(B, N - 1, 1, dtype=torch.float32, device=device)
    stats = torch.zeros(4, dtype=torch.int32, device=device)
    with torch.cuda.device(xs.device):
        check(lib().fiode_segway_simulate(C.byref(w), ptr(x0), B, ptr(ts), N, float(rtol), float(atol), ptr(xs), ptr(us),
                                          ptr(stats), stream_ptr()), 'fiode_segway_simulate')
    st = stats.tolist()
    if st[0]:
        raise AssertionError({1: 'underflow in dt', 2: 'non-finite values in state `y`', 3: 'max_num_steps exceeded'}[st[0]])
    simulate.last_stats = dict(n_accept=st[1], n_reject=st[2], nfe=st[3])
    return xs, us


def slab(total, rank, world):
    """Contiguous slab [first, first + count) of the flattened grid for ``rank`` (phi is the slowest axis, so a slab is
    a range of phi planes plus a partial one: SURVEY 8e "slab the phi axis")."""
    per = (total + world - 1) // world
    first = min(total, rank * per)
    return first, min(per, total - first)


def certify_grid_sharded(ctrl_sd, P, sizes, r, level=0.15, device='cuda', levels=None, evaluate=None):
    """control/certify_segway.py:55-59 across the ranks of the default process group: every rank evaluates one slab of
    the grid (grid points are independent), then ONE all-reduce carries (max barrier derivative, shell-point count).
    Returns (eta_local, val_local, global max val, global shell count).  ``evaluate(first, count) -> (eta, val, vmax)``
    can be injected (CPU tests of the sharding); default is ``certify_grid`` on this rank's device."""
    import torch.distributed as dist
    from .. import distributed as fdist
    rank, world = fdist.world()
    total = 1
    for a in grid_axes(sizes, r):
        total *= len(a)
    first, count = slab(total, rank, world)
    if evaluate is None:
        evaluate = lambda f, c: certify_grid(ctrl_sd, P, sizes, r, level=level, device=device, levels=levels, first=f, count=c)
    if count > 0:
        eta, val, vmax = evaluate(first, count)
    else:
        eta, val, vmax = torch.empty(0, 3), torch.empty(0, 1), float('-inf')
    m = eta.shape[0]
    red_dev = eta.device if eta.is_cuda else 'cpu'
    vm = torch.tensor([vmax if m else float('-inf')], dtype=torch.float32, device=red_dev)
    cnt = torch.tensor([m], dtype=torch.int64, device=red_dev)
    if world > 1:
        dist.all_reduce(vm, op=dist.ReduceOp.MAX)
        dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
    return eta, val, float(vm.item()), int(cnt.item())


# ---- adversarial training of the controller and the Lyapunov matrix (control/train_segway.py:119-158) -------------------
class NNController(torch.nn.Module):
    """libs/core/core/controllers/nn_controller.py:6-34: 3 -> 32 -> 32 -> 1, ReLU, weights ~ N(0, 6 / fan_in), zero biases; same
    state_dict keys (``model.{0,2,4}.{weight,bias}``)."""
    def __init__(self, dynamics=None, in_channels=3, out_channels=1, hidden_channels=32):
        super().__init__()
        assert (in_channels, out_channels, hidden_channels) == (3, 1, 32), 'the fused Segway kernels cover NNController(system, 3, 1, 32)'
        nn = torch.nn
        self.model = nn.Sequential(nn.Linear(3, 32), nn.ReLU(), nn.Linear(32, 32), nn.ReLU(), nn.Linear(32, 1))
        for m in self.model.modules():
            if isinstance(m, nn.Linear):
                m.weight.data.normal_(0, math.sqrt(6. / m.in_features))
                m.bias.data.zero_()


class LyaQuadratic(torch.nn.Module):
    """control/control_utils.py:215-234: V = ||P x||^2 with ``P.weight`` trainable (goal = 0)."""
    def __init__(self, P, goal=None):
        super().__init__()
        self.P = torch.nn.Linear(3, 3, bias=False)
        self.P.weight.data = P.detach().clone().float()


class _SegwayHdot(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, W0, b0, W1, b1, W2, b2, P):
        w = SegwayWeights(W0=ptr(W0), b0=ptr(b0), W1=ptr(W1), b1=ptr(b1), W2=ptr(W2), b2=ptr(b2), P=ptr(P))
        x = x.detach().float().contiguous()
        N = x.shape[0]
        val = torch.empty(N, dtype=torch.float32, device=x.device)
        with torch.cuda.device(x.device):
            check(lib().fiode_segway_hdot(C.byref(w), ptr(x), N, ptr(val), stream_ptr()), 'fiode_segway_hdot')
        ctx.save_for_backward(x, W0, b0, W1, b1, W2, b2, P)
        return val[:, None]

    @staticmethod
    def backward(ctx, gout):
        x, W0, b0, W1, b1, W2, b2, P = ctx.saved_tensors
        w = SegwayWeights(W0=ptr(W0), b0=ptr(b0), W1=ptr(W1), b1=ptr(b1), W2=ptr(W2), b2=ptr(b2), P=ptr(P))
        N = x.shape[0]
        g = gout.detach().float().reshape(N).contiguous()
        need_x, need_p = ctx.needs_input_grad[0], any(ctx.needs_input_grad[1:])
        gx = torch.empty(N, 3, dtype=torch.float32, device=x.device) if need_x else None
        gp = torch.empty(lib().fiode_segway_param_count(), dtype=torch.float32, device=x.device) if need_p else None
        nb = lib().fiode_segway_hdot_bwd_workspace_bytes(N)
        ws = torch.empty(nb, dtype=torch.uint8, device=x.device)
        with torch.cuda.device(x.device):
            check(lib().fiode_segway_hdot_bwd(C.byref(w), ptr(x), N, ptr(g), ptr(gx), ptr(gp), ptr(ws), nb, stream_ptr()),
                  'fiode_segway_hdot_bwd')
        if gp is None:
            return (gx,) + (None,) * 7
        parts = torch.split(gp, [96, 32, 1024, 32, 32, 1, 9])
        return gx, parts[0].view(32, 3), parts[1], parts[2].view(32, 32), parts[3], parts[4].view(1, 32), parts[5], parts[6].view(3, 3)


class SegwaySingleBarrierModel(torch.nn.Module):
    """control/control_utils.py:97-111 over AutoLirpaModel(Segway(), controller) (:21-28) and LyaQuadratic: forward(x) (N,3) ->
    (N,1) barrier derivative sum(P^T (P f(x, u(x))) * x), differentiable w.r.t. x, the controller weights and P."""
    def __init__(self, controller, lya):
        super().__init__()
        self.controller, self.barrier = controller, lya

    def _params(self):
        m = self.controller.model
        return [t.float().contiguous() for t in (m[0].weight, m[0].bias, m[2].weight, m[2].bias, m[4].weight, m[4].bias,
                                                 self.barrier.P.weight)]

    def forward(self, x):
        return _SegwayHdot.apply(x, *self._params())

    forward_adv = forward

    def input_gradient(self, x):
        """d mean(val) / d x without building a graph (what LinfPGDAttack's loss_fn = mean needs)."""
        ps = [p.detach() for p in self._params()]
        w = SegwayWeights(**{k: ptr(v) for k, v in zip(('W0', 'b0', 'W1', 'b1', 'W2', 'b2', 'P'), ps)})
        x = x.detach().float().contiguous()
        N = x.shape[0]
        gx = torch.empty(N, 3, dtype=torch.float32, device=x.device)
        with torch.cuda.device(x.device):
            check(lib().fiode_segway_hdot_bwd(C.byref(w), ptr(x), N, None, ptr(gx), None, None, 0, stream_ptr()), 'fiode_segway_hdot_bwd')
        return gx                     # gradient of sum(val); sign(.) is what the attack uses, the 1/N of the mean drops out


def linf_pgd(model, x, eps=0.02, nb_iter=7, eps_iter=None, clip_min=-2 * math.pi, clip_max=2 * math.pi, rand_init=True,
             delta0=None, generator=None):
    """advertorch.attacks.LinfPGDAttack(model.forward_adv, loss_fn=mean, eps, nb_iter, eps_iter, rand_init, clip_min, clip_max,
    targeted=False).perturb(x) as train_segway.py:119-123 configures it (perturb_iterative, ord = inf: maximise the mean barrier
    derivative).  ``delta0`` replaces the uniform(-eps, eps) start (parity runs)."""
    eps_iter = 2.5 * eps / nb_iter if eps_iter is None else eps_iter
    x = x.detach().float().contiguous()
    if delta0 is None:
        delta0 = (torch.rand(x.shape, device=x.device, generator=generator) * 2 - 1) * eps if rand_init else torch.zeros_like(x)
    x_adv = torch.clamp(x + delta0.to(x), clip_min, clip_max).contiguous()
    for _ in range(nb_iter):
        g = model.input_gradient(x_adv)
        with torch.cuda.device(x.device):
            check(lib().fiode_segway_pgd_step(ptr(x_adv), ptr(x), ptr(g), x.numel(), float(eps), float(eps_iter), float(clip_min),
                                              float(clip_max), stream_ptr()), 'fiode_segway_pgd_step')
    return x_adv


def train_step(model, eta, optimizer, adv_train=True, eps=0.02, margin=0.01, delta0=None):
    """One epoch body of train_segway.py:138-152: PGD on the shell points (model in eval / no parameter gradients), then
    loss = sum(relu(model(eta_adv) + margin)) (or sum(relu(model(eta))) without the attack), backward, optimizer step."""
    optimizer.zero_grad()
    if adv_train:
        eta_adv = linf_pgd(model, eta, eps=eps, delta0=delta0)
        loss = torch.relu(model(eta_adv) + margin).sum()
    else:
        loss = torch.relu(model(eta)).sum()
    loss.backward()
    optimizer.step()
    return loss.detach()
