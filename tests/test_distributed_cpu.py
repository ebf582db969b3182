"""world_size-2 gloo tests (CPU) of the host-side multi-GPU logic: image sharding + count reduction of
the certification loop, and the flat-buffer gradient all-reduce.  The per-image evaluator is injected
(the oracle), because the CUDA kernels cannot run here -- what is under test is the plumbing."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from tests import fixtures as fx


def _free_port():
    with socket.socket() as s:
        s.bind(('127.0.0.1', 0))
        return s.getsockname()[1]


def _certify_worker(rank, world, port, out_q):
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    from fiode_b200.robustness import certify_lipschitz as cl
    from oracle import certify as oc, dynamics as od
    sd = fx.synth_state_dict()
    c = fx.load_golden('certify.pt')['T10_sn0_shift0.0']
    ivp = fx.build_b200_model(sd, device='cpu')
    z = oc.boundary_grid_numerators(10, c['T'])
    Q = od.dyn_Q(sd, fx.DYN)
    cfg = fx.dyn_cfg()
    s_all = {tuple(torch.round(c['x'][i, 0, 0, :4] * 1e6).tolist()): c['s'][i:i + 1] for i in range(c['x'].shape[0])}

    def backbone(x):
        return torch.cat([s_all[tuple(torch.round(xi[0, 0, :4] * 1e6).tolist())] for xi in x])

    def evaluate(s, lab):
        rows = []
        for m in range(s.shape[0]):
            label = int(lab[m])
            rows.append(oc.certify_image(oc.grid_for_label(z, c['T'], label), s[m:m + 1], label, sd, cfg, eps=c['eps'],
                                         T=c['T'], batches=c['batches'], prefix=fx.DYN, Q=Q))
        return torch.tensor(rows, dtype=torch.float32)

    res = cl.certify(ivp, c['x'], c['labels'], T=c['T'], eps=c['eps'], batches=c['batches'], images_per_launch=2,
                     with_accuracy=False, evaluate=evaluate, backbone=backbone)
    if rank == 0:
        out_q.put(res)
    dist.destroy_process_group()


def test_certify_sharding_and_count_reduction():
    c = fx.load_golden('certify.pt')['T10_sn0_shift0.0']
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_certify_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = q.get(timeout=300)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    want = c['out'][:, 0]
    assert res['world'] == 2 and res['n_images'] == want.shape[0]
    got = torch.tensor(res['violations_store'], dtype=torch.float64)
    assert (got - want).abs().max() < 2e-4                      # every image landed in its own slot
    assert res['count_certify'] == int((want < 0).sum())
    assert res['certified_idx'] == (want < 0).nonzero()[:, 0].tolist()


def _grad_worker(rank, world, port, out_q):
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    from fiode_b200.distributed import FlatGradAllReduce, shard_indices
    torch.manual_seed(0)
    net = torch.nn.Sequential(torch.nn.Linear(7, 5), torch.nn.Linear(5, 3))
    for i, p in enumerate(net.parameters()):
        p.grad = torch.full_like(p, float(rank + 1) * (i + 1))
    red = FlatGradAllReduce(net.parameters())
    flat = red.step()
    ok = all(torch.allclose(p.grad, torch.full_like(p, 1.5 * (i + 1))) for i, p in enumerate(net.parameters()))
    alias = all(p.grad.data_ptr() >= flat.data_ptr() for p in net.parameters())
    if rank == 0:
        out_q.put((ok, alias, red.numel, shard_indices(7)))
    dist.destroy_process_group()


def test_flat_gradient_allreduce():
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_grad_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    ok, alias, numel, shard = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert ok and alias and numel == 7 * 5 + 5 + 5 * 3 + 3 and shard == [0, 2, 4, 6]


def _segway_worker(rank, world, port, out_q):
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    from fiode_b200.control import segway
    from oracle import segway as oseg
    g = fx.load_golden('segway.pt')
    sizes, r = g['sizes'].tolist(), [g['r']] * 3
    grid = oseg.grid_uniform_3d(sizes, r)

    def evaluate(first, count):           # the oracle as the per-slab evaluator: what is under test is the sharding
        pts = grid[first:first + count]
        V = ((pts @ g['P'].T) ** 2).sum(1)
        keep = (V >= g['lb']) & (V <= g['ub'])
        eta = pts[keep]
        val = oseg.barrier_hdot(eta, g['ctrl'], g['P'])
        return eta, val.reshape(-1, 1), float(val.max()) if eta.shape[0] else float('-inf')

    eta, val, vmax, count = segway.certify_grid_sharded(g['ctrl'], g['P'], sizes, r, level=g['level'],
                                                        levels=(g['lb'], g['ub']), evaluate=evaluate)
    out_q.put((rank, eta.shape[0], vmax, count))
    dist.barrier()
    dist.destroy_process_group()


def test_segway_grid_sharding_and_max_reduction():
    """slabs of the grid per rank, one max / count reduction; equals the reference's one-shot result."""
    g = fx.load_golden('segway.pt')
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_segway_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert sum(m for _, m, _, _ in res) == res[0][3]
    assert abs(res[0][3] - g['mask_count']) <= 2
    for _, _, vmax, _ in res:
        assert abs(vmax - g['val_max']) <= 1e-5 * max(1.0, abs(g['val_max']))
    from fiode_b200.control import segway
    assert segway.slab(10, 0, 3) == (0, 4) and segway.slab(10, 2, 3) == (8, 2) and segway.slab(2, 2, 3) == (2, 0)
