"""Small driver for ncu captures: one certify_grid call (grid_20 x 296 images) and a few b=256 solves."""
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
if which in ('certify', 'all'):
    T, M = 20, 296
    gz = fgrid.boundary_grid_numerators(10, T, device=dev)
    g = torch.Generator().manual_seed(1236)
    s = (torch.randn(M, 10, generator=g) * 0.5).to(dev)
    labels = torch.randint(0, 10, (M,), generator=g).to(dev)
    out = ops.certify_grid(ivp.dyn_fun.dyn_weights(), gz, T, s, labels, math.sqrt(2) / 0.225 * 0.141, batches=10)
    print('certify', out[:2].tolist())
if which in ('solve', 'all'):
    x = torch.rand(256, 3, 32, 32, generator=torch.Generator().manual_seed(1)).to(dev)
    ts = torch.linspace(0., 1., 2)
    for _ in range(3):
        y = ivp(x, ts=ts, int_params=dict(method='dopri5', rtol=1e-3, atol=1e-3))
    print('solve', y[0].tolist())
torch.cuda.synchronize()
