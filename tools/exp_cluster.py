"""Dev tool: cluster size / solves-per-launch sweep of the batched dopri5 launch (bench inputs, tol 1e-3)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from fiode_b200 import ops

dev = torch.device('cuda', 0)
ivp = bench.build_model(dev)
B = 256
t = torch.tensor([0., 1.], dtype=torch.float64, device=dev)
dw = ivp.dyn_fun.dyn_weights()
x = torch.rand(64, B, 3, 32, 32, generator=torch.Generator().manual_seed(1234)).to(dev)
pm = ivp.init_coordinates.param_map
s_all = torch.cat([pm(x[i * 16:(i + 1) * 16].reshape(-1, 3, 32, 32)) for i in range(4)])
h_all = torch.full((64 * B, 10), 0.1, device=dev)


def timed(fn, it=20):
    fn(); fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it


for C in (8, 4, 2):
    if 256 // C > 64:
        continue
    n = ops.dopri5_max_concurrent_solves(B, C)
    for G in sorted({1, n // 2, n - 1, n, n + 1, 2 * n, 16, 32, 64}):
        if G < 1 or G > 64:
            continue
        ms = timed(lambda: ops.dopri5_simplex_batched(dw, h_all[:G * B], s_all[:G * B], G, t, 1e-3, 1e-3, cluster_size=C))
        st = ops.dopri5_simplex_batched(dw, h_all[:G * B], s_all[:G * B], G, t, 1e-3, 1e-3, cluster_size=C).stats()[0]
        print(f'cluster {C} (max co-resident {n:3d})  G={G:3d}: {ms * 1e3:8.1f} us  -> {G / ms:8.1f} solves/ms   nfe={st.nfe}')
