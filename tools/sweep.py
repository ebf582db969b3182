"""BASELINE config 5: dopri5 batch / tolerance sweep on the CIFAR dynamics (batch 64..8192, rtol = atol 1e-2..1e-5),
over BOTH solver launches -- `fiode_dopri5_simplex` (one cooperative launch per solve) and
`fiode_dopri5_simplex_batched` (the cluster-per-solve launch the bench uses; B <= 64 rows x cluster size, clusters of 4, 8 and 16) -- reporting
accepted / rejected steps, NFE, device time per solve, rows/s, achieved FLOP/s (NFE x 37 888 x B) and the algorithmic
HBM GB/s (120 B per row).  For B <= 256 the oracle's counts (tools/sweep_oracle.py, run on the CPU in the
build container: the restated torchdiffeq 0.2.2 dopri5 around the reference-pinned RHS) are printed beside every GPU
row, with a verdict column: `equal`, `noise` (inside the band the oracle itself spans when the static features carry
1e-6 relative noise, widened by one step: the step controller is chaotic at these tolerances), or `different`.
Because the README hyper-parameters at random init never reject a step, two stress dynamics that do exercise the
reject path are swept too.

    python tools/sweep.py [--out profiles/rNN_sweep.md] [--oracle profiles/r02_sweep_oracle.json]
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from fiode_b200 import ops  # noqa: E402

CASES = [
    ('README', 'README (alpha_1=100, alpha_2=20, t_max=1)', dict(alpha_1=100., alpha_2=20., scale_nominal=False), 1.0),
    ('stress A', 'stress A (alpha_2=50, scale_nominal, t_max=5)', dict(alpha_1=100., alpha_2=50., scale_nominal=True), 5.0),
    ('stress B', 'stress B (alpha_1=1000, alpha_2=200, scale_nominal, t_max=5)', dict(alpha_1=1000., alpha_2=200., scale_nominal=True), 5.0),
]
BATCHES = [64, 128, 256, 512, 1024, 2048, 4096, 8192]
TOLS = [1e-2, 1e-3, 1e-4, 1e-5]


def verdict(gpu, row):
    """equal / noise / different of a GPU (accepted, rejected, nfe) triple against the oracle row."""
    if row is None:
        return ''
    if list(gpu) == list(row['oracle']):
        return 'equal'
    alls = [row['oracle']] + row['oracle_noise_1e6']
    for k in range(3):
        lo, hi = min(a[k] for a in alls), max(a[k] for a in alls)
        slack = (1, 1, 6)[k] + (hi - lo)          # the band the oracle spans under 1e-6 noise, doubled, plus one step
        if not (lo - slack <= gpu[k] <= hi + slack):
            return 'different'
    return 'noise'


def sweep(iters=5, oracle_path=None, batches=BATCHES, tols=TOLS, dev=None):
    dev = dev or torch.device('cuda', 0)
    # the SAME weights and images as tools/sweep_oracle.py (tests/fixtures.py: reference-layout state_dict + seed-1234 images):
    # the step controller is sensitive to the static features, so counts are only comparable on identical inputs
    from tests import fixtures as fx
    ivp = fx.build_b200_model(fx.synth_state_dict(), device=dev)
    dyn = ivp.dyn_fun
    pm = ivp.init_coordinates.param_map
    orc = {}
    if oracle_path and os.path.exists(oracle_path):
        for r in json.load(open(oracle_path))['rows']:
            orc[(r['case'], r['B'], r['tol'])] = r
    nmax = max(batches)
    x = torch.cat([fx.synth_images(256, seed=1234), torch.rand(max(nmax - 256, 0), 3, 32, 32, generator=torch.Generator().manual_seed(4321))])[:nmax].to(dev)
    s_all = torch.cat([pm(x[i:i + 4096]) for i in range(0, nmax, 4096)])
    h_all = torch.full((nmax, 10), 0.1, device=dev)
    rows = []
    saved = {k: getattr(dyn, k) for k in ('alpha_1', 'alpha_2', 'scale_nominal')}
    for key, name, cfg, t_max in CASES:
        for k, v in cfg.items():
            setattr(dyn, k, v)
        dw = dyn.dyn_weights()
        t = torch.tensor([0., t_max], dtype=torch.float64, device=dev)
        for B in batches:
            for tol in tols:
                launches = [('cooperative', lambda: ops.dopri5_simplex(dw, h_all[:B], s_all[:B], t, tol, tol), 1)]
                for C in (4, 8, 16):
                    if B % C == 0 and B // C <= 64 and B // C >= 8:
                        G = max(1, ops.dopri5_max_concurrent_solves(B, C))
                        G = min(G, nmax // B)
                        launches.append((f'cluster{C} x{G}', (lambda C=C, G=G: ops.dopri5_simplex_batched(
                            dw, h_all[:G * B], s_all[:B].repeat(G, 1), G, t, tol, tol, cluster_size=C)), G))
                for lname, run, G in launches:
                    r = run()
                    st = r.stats() if G == 1 and lname == 'cooperative' else r.stats()[0]
                    if st.status:
                        rows.append(dict(case=key, title=name, launch=lname, B=B, tol=tol, status=st.status))
                        continue
                    ms = bench.timed(run, iters) / iters
                    out = r.out[-1][:B]
                    ok = bool(((out.sum(-1) - 1).abs() < 1e-3).all())
                    o = orc.get((key, B, tol))
                    rows.append(dict(case=key, title=name, launch=lname, B=B, tol=tol, n_accept=st.n_accept, n_reject=st.n_reject, nfe=st.nfe,
                                     us=ms * 1e3 / G, us_launch=ms * 1e3, solves_per_launch=G, rows_per_s=G * B * 1e3 / ms,
                                     tflops=G * st.nfe * bench.RHS_FLOP * B / (ms * 1e-3) / 1e12, hbm_gbs=120.0 * G * B / (ms * 1e-3) / 1e9,
                                     on_simplex=ok, oracle=o['oracle'] if o else None,
                                     oracle_noise=o['oracle_noise_1e6'] if o else None,
                                     verdict=verdict((st.n_accept, st.n_reject, st.nfe), o)))
    for k, v in saved.items():
        setattr(dyn, k, v)
    return rows


def summary(rows):
    cmp_rows = [r for r in rows if r.get('verdict')]
    cnt = {v: sum(1 for r in cmp_rows if r['verdict'] == v) for v in ('equal', 'noise', 'different')}
    worst = max(cmp_rows, key=lambda r: abs(r['nfe'] - r['oracle'][2]) / max(r['oracle'][2], 1), default=None)
    best = max((r for r in rows if 'rows_per_s' in r), key=lambda r: r['rows_per_s'], default=None)
    return dict(cells=len(rows), compared_with_oracle=len(cmp_rows), **cnt, failed=sum(1 for r in rows if 'status' in r),
                worst=None if worst is None else dict(case=worst['case'], launch=worst['launch'], B=worst['B'], tol=worst['tol'],
                                                      gpu=[worst['n_accept'], worst['n_reject'], worst['nfe']], oracle=worst['oracle']),
                peak_rows_per_s=None if best is None else dict(value=best['rows_per_s'], B=best['B'], tol=best['tol'], launch=best['launch'],
                                                              case=best['case'], tflops=best['tflops'], hbm_gbs=best['hbm_gbs']))


def render(rows, iters):
    lines = ['# dopri5 batch / tolerance sweep (BASELINE config 5), one B200',
             '', 'Launches: `cooperative` = `fiode_dopri5_simplex` (one cooperative launch per solve); `clusterC xG` = '
             '`fiode_dopri5_simplex_batched`, G independent solves in one launch, a cluster of C CTAs each (what `bench.py` times); '
             'us / solve = launch time / G.  Device time = CUDA events around %d back-to-back launches, after a warm-up.  FLOP = NFE x 37 888 x B; '
             'GB/s = 120 B per row (h0 in, s in, y out) over the solve time: the solver is latency-bound, not HBM-bound.  '
             'Oracle column (B <= 256): accepted / rejected / NFE of `oracle/dopri5.py` on the same inputs, and the min-max of the same oracle '
             'under 1e-6 relative noise in the static features (`tools/sweep_oracle.py`).' % iters, '',
             'Summary: `%s`' % json.dumps(summary(rows)), '']
    for key, name, _, _ in CASES:
        lines += [f'## {name}', '', '| B | tol | launch | accepted | rejected | NFE | oracle acc / rej / NFE | oracle under 1e-6 noise | verdict | us / solve | rows/s | TFLOP/s | GB/s | on simplex |',
                  '|---|---|---|---|---|---|---|---|---|---|---|---|---|---|']
        for r in rows:
            if r['case'] != key:
                continue
            if 'status' in r:
                lines.append(f"| {r['B']} | {r['tol']:g} | {r['launch']} | status {r['status']} | | | | | | | | | | |")
                continue
            if r['oracle']:
                o = ' / '.join(str(v) for v in r['oracle'])
                ns = r['oracle_noise']
                band = ', '.join('%d-%d' % (min(a[k] for a in ns), max(a[k] for a in ns)) for k in range(3))
            else:
                o = band = ''
            lines.append(f"| {r['B']} | {r['tol']:g} | {r['launch']} | {r['n_accept']} | {r['n_reject']} | {r['nfe']} | {o} | {band} | {r['verdict']} | "
                         f"{r['us']:.1f} | {r['rows_per_s']:.3g} | {r['tflops']:.2f} | {r['hbm_gbs']:.2f} | {r['on_simplex']} |")
        lines.append('')
    return '\n'.join(lines)


if __name__ == '__main__':
    ap = argparse.ArgumentParser()
    ap.add_argument('--out', default='')
    ap.add_argument('--oracle', default=os.path.join(ROOT, 'profiles', 'r02_sweep_oracle.json'))
    ap.add_argument('--iters', type=int, default=5)
    args = ap.parse_args()
    rows = sweep(args.iters, args.oracle)
    text = render(rows, args.iters)
    print(text)
    if args.out:
        open(args.out, 'w').write(text + '\n')
        json.dump(dict(summary=summary(rows), rows=rows), open(os.path.splitext(args.out)[0] + '.json', 'w'))
