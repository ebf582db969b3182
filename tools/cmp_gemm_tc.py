"""Dev tool: tensor-core GEMM path vs FFMA path (linear layers and the whole backbone), with timings."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tests import fixtures as fx
from fiode_b200 import ops
from fiode_b200._lib import lib

dev = 'cuda:0'
L = lib()


def timed(fn, it=10):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it * 1e3

bad = 0
g = torch.Generator().manual_seed(3)
for (M, N, K, gs) in [(128, 64, 64, 0), (300, 512, 512, 1), (1000, 512, 4096, 1), (4096, 512, 4096, 1), (77, 128, 128, 0), (513, 32, 128, 1)]:
    X = torch.randn(M, K, generator=g).to(dev)
    Q = (torch.randn(N, K, generator=g) / K ** 0.5).to(dev)
    b = torch.randn(N, generator=g).to(dev)
    out = {}
    for path in (1, 0):
        L.fiode_set_mlp_path(path)
        out[path] = ops.linear_fwd(X, Q, b, group_sort=bool(gs))
        tt = timed(lambda: ops.linear_fwd(X, Q, b, group_sort=bool(gs)))
        out[path, 't'] = tt
    ref = torch.nn.functional.linear(X.double(), Q.double(), b.double())
    if gs:
        a_, b_ = ref[:, :N // 2], ref[:, N // 2:]
        ref = torch.cat([torch.maximum(a_, b_), torch.minimum(a_, b_)], 1)
    e_tc = (out[0].double() - ref).abs().max().item(); e_ff = (out[1].double() - ref).abs().max().item()
    print(f'linear M={M:5d} N={N:4d} K={K:5d} gs={gs}: err tc {e_tc:.2e} ffma {e_ff:.2e} | tc {out[0, "t"]:8.1f} us  ffma {out[1, "t"]:8.1f} us '
          f'({2 * M * N * K / out[0, "t"] / 1e6:.1f} TFLOP/s)')
    bad += e_tc > 1e-5
sd = fx.synth_state_dict()
ivp = fx.build_b200_model(sd, device=dev)
bw, keep = ivp.init_coordinates.param_map.backbone_weights()
for n in (256, 4096):
    x = torch.rand(n, 3, 32, 32, generator=g).to(dev)
    res = {}
    for path in (1, 0):
        L.fiode_set_mlp_path(path)
        res[path] = ops.backbone_fwd(bw, x)
        res[path, 't'] = timed(lambda: ops.backbone_fwd(bw, x), 5)
    d = (res[0] - res[1]).abs().max().item()
    print(f'backbone {n:5d} images: max |tc - ffma| = {d:.2e} | tc {res[0, "t"]:9.1f} us  ffma {res[1, "t"]:9.1f} us')
    bad += d > 2e-5
L.fiode_set_mlp_path(0)
print('FAILED' if bad else 'OK')
sys.exit(1 if bad else 0)
