"""model.0 on the tensor cores (dft_tc.cu) against the CUDA-core FFT kernel: run once per FIODE_CONV0 setting, each run saves the
backbone output of 512 seeded images; the second run prints the difference.  Also prints the stage times of an 8192-image pass."""
import ctypes
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from fiode_b200 import ops  # noqa: E402
from fiode_b200._lib import lib  # noqa: E402

dev = torch.device('cuda', 0)
ivp = bench.build_model(dev)
pm = ivp.init_coordinates.param_map
x = torch.rand(8192, 3, 32, 32, generator=torch.Generator().manual_seed(7)).to(dev)
bw, keep = pm.backbone_weights()
s = ops.backbone_fwd(bw, x[:515])
torch.cuda.synchronize()
tag = os.environ.get('FIODE_CONV0', 'tc')
out = 'gpurun_out/conv0_%s.pt' % tag
os.makedirs('gpurun_out', exist_ok=True)
torch.save(s.cpu(), out)
other = 'gpurun_out/conv0_%s.pt' % ('fft' if tag != 'fft' else 'tc')
if os.path.exists(other):
    o = torch.load(other)
    print('max |s_%s - s_other| = %.3e   (max |s| = %.3f)' % (tag, float((s.cpu() - o).abs().max()), float(o.abs().max())))
L = lib()
buf = (ctypes.c_float * 16)()
L.fiode_backbone_enable_stage_timing(1)
acc = [0.0] * 11
for i in range(7):
    ops.backbone_fwd(bw, x)
    L.fiode_backbone_stage_times(buf, 16)
    if i >= 2:
        for j in range(11):
            acc[j] += buf[j] / 5
L.fiode_backbone_enable_stage_timing(0)
print(tag, 'stage ms per 8192 images:', ' '.join('%.3f' % a for a in acc), ' total %.3f' % sum(acc))
