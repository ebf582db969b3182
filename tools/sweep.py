"""BASELINE config 5: dopri5 batch / tolerance sweep on the CIFAR dynamics (batch 64..8192, rtol = atol 1e-2..1e-5),
reporting accepted / rejected steps, NFE, device time per solve, rows/s, achieved FLOP/s (NFE x 37 888 x B) and the
algorithmic HBM GB/s (120 B per row).  Because the README hyper-parameters at random init never reject a
step, two stress dynamics that do exercise the reject path are swept too.

    python tools/sweep.py [--out profiles/rNN_sweep.md]
"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from fiode_b200 import ops  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument('--out', default='')
ap.add_argument('--iters', type=int, default=5)
args = ap.parse_args()
dev = torch.device('cuda', 0)
ivp = bench.build_model(dev)
dyn = ivp.dyn_fun
pm = ivp.init_coordinates.param_map
CASES = [
    ('README (alpha_1=100, alpha_2=20, t_max=1)', dict(alpha_1=100., alpha_2=20., scale_nominal=False), 1.0),
    ('stress A (alpha_2=50, scale_nominal, t_max=5)', dict(alpha_1=100., alpha_2=50., scale_nominal=True), 5.0),
    ('stress B (alpha_1=1000, alpha_2=200, scale_nominal, t_max=5)', dict(alpha_1=1000., alpha_2=200., scale_nominal=True), 5.0),
]
BATCHES = [64, 128, 256, 512, 1024, 2048, 4096, 8192]
TOLS = [1e-2, 1e-3, 1e-4, 1e-5]
x = torch.rand(8192, 3, 32, 32, generator=torch.Generator().manual_seed(1234)).to(dev)
s_all = torch.cat([pm(x[i:i + 4096]) for i in range(0, 8192, 4096)])
h_all = torch.full((8192, 10), 0.1, device=dev)
rows = []
for name, cfg, t_max in CASES:
    for k, v in cfg.items():
        setattr(dyn, k, v)
    dw = dyn.dyn_weights()
    t = torch.tensor([0., t_max], dtype=torch.float64, device=dev)
    for B in BATCHES:
        for tol in TOLS:
            run = lambda: ops.dopri5_simplex(dw, h_all[:B], s_all[:B], t, tol, tol)
            r = run()
            st = r.stats()
            if st.status:
                rows.append(dict(case=name, B=B, tol=tol, status=st.status))
                continue
            ms = bench.timed(run, args.iters) / args.iters
            ok = bool(((r.out[-1].sum(1) - 1).abs() < 1e-3).all())
            rows.append(dict(case=name, B=B, tol=tol, n_accept=st.n_accept, n_reject=st.n_reject, nfe=st.nfe, us=ms * 1e3,
                             solves_per_s=1e3 / ms, rows_per_s=B * 1e3 / ms, tflops=st.nfe * bench.RHS_FLOP * B / (ms * 1e-3) / 1e12,
                             hbm_gbs=120.0 * B / (ms * 1e-3) / 1e9, on_simplex=ok))
lines = ['# dopri5 batch / tolerance sweep (BASELINE config 5), one B200, `fiode_dopri5_simplex` (one cooperative launch per solve)',
         '', 'Device time = CUDA events around %d back-to-back solves, after a warm-up solve.  FLOP = NFE x 37 888 x B; '
         'GB/s = 120 B per row (h0 in, s in, y out) over the solve time: the solver is latency-bound, not HBM-bound.' % args.iters, '']
for name, _, _ in CASES:
    lines += [f'## {name}', '', '| B | tol | accepted | rejected | NFE | us / solve | rows/s | TFLOP/s | GB/s | on simplex |', '|---|---|---|---|---|---|---|---|---|---|']
    for r in rows:
        if r['case'] != name:
            continue
        if 'status' in r:
            lines.append(f"| {r['B']} | {r['tol']:g} | status {r['status']} | | | | | | | |")
        else:
            lines.append(f"| {r['B']} | {r['tol']:g} | {r['n_accept']} | {r['n_reject']} | {r['nfe']} | {r['us']:.1f} | {r['rows_per_s']:.3g} | "
                         f"{r['tflops']:.2f} | {r['hbm_gbs']:.2f} | {r['on_simplex']} |")
    lines.append('')
text = '\n'.join(lines)
print(text)
if args.out:
    open(args.out, 'w').write(text + '\n')
    json.dump(rows, open(os.path.splitext(args.out)[0] + '.json', 'w'))
