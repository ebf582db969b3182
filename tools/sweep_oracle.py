"""CPU side of the C5 sweep: the oracle's (restated torchdiffeq 0.2.2 dopri5 around the reference-pinned RHS) accepted /
rejected / NFE counts for every (dynamics, B <= 256, tol) cell of tools/sweep.py, plus the counts the SAME oracle gives
when the static features are perturbed by 1e-6 relative noise (4 draws): the step controller of a stiff solve is chaotic
(dt_{n+1} = dt_n * 0.9 * ratio^-1/5 with ratios that are round-off at loose tolerances), so two correct fp32
implementations -- or one implementation on two CPUs -- land on different counts; the band says by how much.

    python tools/sweep_oracle.py --out profiles/r02_sweep_oracle.json      (CPU only, ~10 min)
"""
import argparse
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tests import fixtures as fx                     # noqa: E402
from oracle import layers as ol, dynamics as od, dopri5 as odp   # noqa: E402

CASES = [
    ('README', dict(alpha_1=100., alpha_2=20., scale_nominal=False), 1.0),
    ('stress A', dict(alpha_1=100., alpha_2=50., scale_nominal=True), 5.0),
    ('stress B', dict(alpha_1=1000., alpha_2=200., scale_nominal=True), 5.0),
]
BATCHES = [64, 128, 256]
TOLS = [1e-2, 1e-3, 1e-4, 1e-5]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--out', default='profiles/r02_sweep_oracle.json')
    ap.add_argument('--draws', type=int, default=4)
    args = ap.parse_args()
    torch.set_num_threads(os.cpu_count())
    sd = fx.synth_state_dict()
    x = fx.synth_images(256, seed=1234)
    s_all = ol.backbone_forward(x, sd, fx.PM)
    Q = od.dyn_Q(sd, fx.DYN)
    rows = []
    t0 = time.time()
    for name, kw, t_max in CASES:
        cfg = fx.dyn_cfg(**kw)
        for B in BATCHES:
            h0 = torch.full((B, 10), 0.1)
            for tol in TOLS:
                def solve(s):
                    _, st = odp.odeint_dopri5(lambda t, hh: od.eval_dot(hh[0], s, sd, cfg, Q, fx.DYN), (h0,),
                                              torch.tensor([0., t_max]), tol, tol, return_stats=True)
                    return [st.n_accept, st.n_reject, st.nfe]
                base = solve(s_all[:B])
                g = torch.Generator().manual_seed(B * 7 + int(-torch.log10(torch.tensor(tol)).item()))
                noisy = [solve(s_all[:B] * (1 + 1e-6 * torch.randn(B, 10, generator=g))) for _ in range(args.draws)]
                rows.append(dict(case=name, B=B, tol=tol, oracle=base, oracle_noise_1e6=noisy))
                print(name, B, tol, base, noisy, f'{time.time() - t0:.0f}s', flush=True)
    json.dump(dict(how='oracle/dopri5.py on CPU torch fp32; s from oracle backbone of fixtures.synth_images(256, 1234); '
                       'noise = s * (1 + 1e-6 randn), %d draws' % args.draws, threads=torch.get_num_threads(), rows=rows),
              open(args.out, 'w'), indent=1)


if __name__ == '__main__':
    main()
