"""Dev tool: time the pieces of a step of G batch-256 solves (backbone over G*256 images, batched dopri5)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tests import fixtures as fx
from fiode_b200 import ops

G = int(sys.argv[1]) if len(sys.argv) > 1 else 16
B = 256
dev = 'cuda:0'
sd = fx.synth_state_dict()
ivp = fx.build_b200_model(sd, device=dev)
pm = ivp.init_coordinates.param_map
bw, keep = pm.backbone_weights()
dw = ivp.dyn_fun.dyn_weights()
x = torch.rand(G * B, 3, 32, 32, device=dev)
t = torch.tensor([0., 1.], dtype=torch.float64, device=dev)
h0 = torch.full((G * B, 10), 0.1, device=dev)


def timed(fn, it=20):
    fn(); fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it

for n in (256, 1024, G * B):
    print(f'backbone {n:5d} images: {timed(lambda: ops.backbone_fwd(bw, x[:n])) * 1e3:9.1f} us')
s = ops.backbone_fwd(bw, x)
print(f'dopri5 single b=256      : {timed(lambda: ops.dopri5_simplex(dw, h0[:B], s[:B], t, 1e-3, 1e-3)) * 1e3:9.1f} us')
for g in (1, 4, G):
    print(f'dopri5 batched G={g:3d}     : {timed(lambda: ops.dopri5_simplex_batched(dw, h0[:g * B], s[:g * B], g, t, 1e-3, 1e-3)) * 1e3:9.1f} us')
r = ops.dopri5_simplex_batched(dw, h0, s, G, t, 1e-3, 1e-3)
st = r.stats()
print('stats[0]:', st[0].n_accept, st[0].n_reject, st[0].nfe)
