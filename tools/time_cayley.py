"""Dev tool: CUDA-event times of the Cayley weight preparation and its adjoint at the layer shapes of the CIFAR model."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from fiode_b200 import ops

dev = 'cuda'
for cout, cin in [(128, 10), (128, 128), (10, 128), (512, 4096), (512, 512), (10, 512), (32, 128), (64, 256)]:
    W = (torch.rand(cout, cin, device=dev) * 2 - 1) / cin ** 0.5
    a = W.norm()
    g = torch.randn(cout, cin, device=dev)
    f = bench.timed(lambda: ops.cayley_linear(W, a), 20) / 20
    b = bench.timed(lambda: ops.cayley_linear_bwd(W, a, g), 20) / 20
    print(f'cayley_linear {cout:4d} x {cin:4d}: fwd {f * 1e3:8.1f} us   bwd {b * 1e3:8.1f} us')
for cout, cin, k, n in [(32, 3, 3, 32), (64, 32, 3, 16)]:
    w = torch.randn(cout, cin, k, k, device=dev) * (cin * k * k) ** -0.5
    a = ops.conv_alpha_init(w, n)
    F = n * (n // 2 + 1)
    g = torch.randn(F, cout, cin, 2, device=dev)
    f = bench.timed(lambda: ops.cayley_conv(w, a, n), 20) / 20
    b = bench.timed(lambda: ops.cayley_conv_bwd(w, a, n, torch.view_as_complex(g)), 20) / 20
    print(f'cayley_conv {cout}x{cin} k{k} n{n}: fwd {f * 1e3:8.1f} us   bwd {b * 1e3:8.1f} us')
