import ctypes as C, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from fiode_b200._lib import lib, ptr, stream_ptr
dev = torch.device('cuda', 0)
ivp = bench.build_model(dev)
pm = ivp.init_coordinates.param_map
B = int(sys.argv[1]) if len(sys.argv) > 1 else 5
x = torch.rand(B, 3, 32, 32, generator=torch.Generator().manual_seed(7)).to(dev)
bw, keep = pm.backbone_weights()
L = lib()
nb = L.fiode_backbone_workspace_bytes(B)
ws = torch.empty(nb, dtype=torch.uint8, device=dev)
outs = []
for path in (0, 1, 2):
    a0 = torch.zeros(B, 32, 32, 32, device=dev)
    rc = L.fiode_debug_conv0(C.byref(bw), ptr(x), B, path, ptr(a0), ptr(ws), C.c_size_t(nb), stream_ptr())
    torch.cuda.synchronize()
    assert rc == 0, L.fiode_last_error()
    outs.append(a0.cpu())
for path in (1, 2):
    d = (outs[path] - outs[0]).abs()
    print('path', path, 'max diff', float(d.max()), 'mean', float(d.mean()))
    bad = (d.amax(dim=(1, 2, 3)) > 1e-4).nonzero().flatten().tolist()
    ns = min(37, B)
    print(' bad images', len(bad), [(b, 'slice', b % ns, 'n', b // ns) for b in bad[:12]])
    for b in bad[:4]:
        ch = (d[b].amax(dim=(1, 2)) > 1e-4).nonzero().flatten().tolist()
        print('  img', b, 'bad channels', ch)
        c0 = ch[0]
        rows = (d[b, c0].amax(dim=1) > 1e-4).nonzero().flatten().tolist()
        cols = (d[b, c0].amax(dim=0) > 1e-4).nonzero().flatten().tolist()
        print('   ch', c0, 'bad rows', rows, 'bad cols', cols, 'max', float(d[b, c0].max()))
