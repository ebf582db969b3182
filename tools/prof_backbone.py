"""Dev tool: backbone forward over N images a few times (for ncu launch lists)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tests import fixtures as fx
from fiode_b200 import ops
N = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
it = int(sys.argv[2]) if len(sys.argv) > 2 else 3
ivp = fx.build_b200_model(fx.synth_state_dict(), device='cuda:0')
bw, keep = ivp.init_coordinates.param_map.backbone_weights()
x = torch.rand(N, 3, 32, 32, device='cuda:0')
for _ in range(it):
    s = ops.backbone_fwd(bw, x)
torch.cuda.synchronize()
print('ok', float(s.sum()))
