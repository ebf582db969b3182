import time, torch, sys
sys.path.insert(0, '/root/repo')
from fiode_b200 import ops
dev = 'cuda:0'
W = (torch.rand(512, 4096, device=dev) * 2 - 1) / 64
a = W.norm() * 1.1
gQ = torch.randn(512, 4096, device=dev)
for save in (False, True):
    for _ in range(3):
        r = ops.cayley_linear(W, a, save_minv=save)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(20):
        r = ops.cayley_linear(W, a, save_minv=save)
    t1 = time.perf_counter(); torch.cuda.synchronize(); t2 = time.perf_counter()
    print('fwd save', save, 'host ms/call', (t1 - t0) / 20 * 1e3, 'total', (t2 - t0) / 20 * 1e3)
Q, minv = ops.cayley_linear(W, a, save_minv=True)
for mv in (None, minv):
    for _ in range(3):
        ops.cayley_linear_bwd(W, a, gQ, minv=mv)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(20):
        ops.cayley_linear_bwd(W, a, gQ, minv=mv)
    t1 = time.perf_counter(); torch.cuda.synchronize(); t2 = time.perf_counter()
    print('bwd saved', mv is not None, 'host ms/call', (t1 - t0) / 20 * 1e3, 'total', (t2 - t0) / 20 * 1e3)
