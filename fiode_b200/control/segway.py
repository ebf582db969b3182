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
