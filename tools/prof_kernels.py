"""Small driver for ncu captures (run under `ncu --profile-from-start off`): the profiled region holds
one certify_grid call (grid_20 x 296 images) and one bench step (16 batch-256 solves: backbone over 4096
images + the cluster-per-solve dopri5 launch), each after un-profiled warm-up calls."""
import math
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from fiode_b200 import ops, grid as fgrid  # noqa: E402

dev = torch.device('cuda', 0)
ivp = bench.build_model(dev)
which = sys.argv[1] if len(sys.argv) > 1 else 'all'
prof = torch.cuda.profiler


def region(fn, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    prof.start()
    out = fn()
    torch.cuda.synchronize()
    prof.stop()
    return out


if which in ('certify', 'all'):
    T, M = 20, 296
    gz = fgrid.boundary_grid_numerators(10, T, device=dev)
    g = torch.Generator().manual_seed(1236)
    s = (torch.randn(M, 10, generator=g) * 0.5).to(dev)
    labels = torch.randint(0, 10, (M,), generator=g).to(dev)
    dw = ivp.dyn_fun.dyn_weights()
    out = region(lambda: ops.certify_grid(dw, gz, T, s, labels, math.sqrt(2) / 0.225 * 0.141, batches=10), warm=1)
    print('certify', out[:2].tolist())
if which in ('step', 'solve', 'all'):
    G = 16
    x = torch.rand(G, 256, 3, 32, 32, generator=torch.Generator().manual_seed(1)).to(dev)
    ts = torch.linspace(0., 1., 2)
    ip = dict(method='dopri5', rtol=1e-3, atol=1e-3, options=dict(check=False))
    y = region(lambda: ivp.forward_batches(x, ts=ts, int_params=ip))
    print('step', y[0, 0].tolist())
torch.cuda.synchronize()
