"""Dev tool: certify kernel on the tcgen05 path vs the FFMA path (same inputs), with timings.
    python tools/cmp_certify.py [T] [M]"""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tests import fixtures as fx
from fiode_b200 import ops, grid as fgrid
from fiode_b200._lib import lib

T = int(sys.argv[1]) if len(sys.argv) > 1 else 20
M = int(sys.argv[2]) if len(sys.argv) > 2 else 148
dev = 'cuda:0'
sd = fx.synth_state_dict()
ivp = fx.build_b200_model(sd, device=dev)
dw = ivp.dyn_fun.dyn_weights()
gz = fgrid.boundary_grid_numerators(10, T, device=dev)
g = torch.Generator().manual_seed(5)
s = (torch.randn(M, 10, generator=g) * 0.5).to(dev)
lab = torch.randint(0, 10, (M,), generator=g).to(dev)
res = {}
for path in (1, 0):
    lib().fiode_set_mlp_path(path)
    out, it = ops.certify_grid(dw, gz, T, s, lab, 0.1, batches=10, return_iters=True)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        out, it = ops.certify_grid(dw, gz, T, s, lab, 0.1, batches=10, return_iters=True)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 3
    res[path] = (out.cpu(), it.cpu())
    print(f'path {"ffma" if path else "tc  "}: {ms:8.3f} ms  {gz.shape[0] * M / ms * 1e3:.3e} pts/s   out[0]={out[0].tolist()} iters[0]={it[0].tolist()[:4]}')
L = lib()
if hasattr(L, 'fiode_debug_tc_prof'):
    import ctypes
    buf = (ctypes.c_longlong * 32)()
    L.fiode_debug_tc_prof(buf, 1)
    L.fiode_set_mlp_path(0)
    ops.certify_grid(dw, gz, T, s, lab, 0.1, batches=10); torch.cuda.synchronize()
    L.fiode_debug_tc_prof(buf, 0)
    names = ['row: wait D1', 'row: S1', 'row: wait D2', 'row: S2', 'row: wait D3', 'row: D3 read + S0(next)', 'row: wait ring empty',
             'row: hand-over', 'proj: wait ring full', 'proj: S3', 'mma: idle', 'mma: issue L1', 'mma: issue L2', 'mma: issue L3']
    ntile = (gz.shape[0] // 10 + 127) // 128 * 10 / 2
    for i, nme in enumerate(names):
        print(f'  {nme:28s} {buf[i] / ntile:9.1f} cyc/tile')
d = (res[0][0] - res[1][0]).abs().max().item()
print(f'grid {gz.shape[0]} pts x {M} images; max |tc - ffma| = {d:.3e}; iters equal: {bool((res[0][1] == res[1][1]).all())}; '
      f'nan: {bool(torch.isnan(res[0][0]).any())}')
sys.exit(0 if d < 5e-5 else 1)
