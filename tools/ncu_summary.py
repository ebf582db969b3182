"""Turn ncu outputs into the markdown summaries kept under profiles/.

    python tools/ncu_summary.py launches gpurun_out/<tag>/launches.csv  > profiles/rNN_launches_summary.md
    python tools/ncu_summary.py full     gpurun_out/<tag>/full.ncu-rep  > profiles/rNN_ncu_full_summary.md
    python tools/ncu_summary.py traffic  gpurun_out/<tag>/full.ncu-rep  > profiles/ncu_traffic.json   (what bench.py's roofline.traffic reads)
"""
import csv
import io
import re
import subprocess
import sys
from collections import OrderedDict

METRICS = [
    'gpu__time_duration.sum', 'launch__grid_size', 'launch__block_size', 'launch__cluster_size', 'launch__registers_per_thread',
    'launch__shared_mem_per_block_dynamic', 'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem',
    'sm__warps_active.avg.pct_of_peak_sustained_active', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
    'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active', 'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active',
    'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active',
    'sm__inst_executed_pipe_tensor.sum', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
    'dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
    'lts__t_sector_hit_rate.pct', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
    'smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio',
    'smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio',
    'smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio',
    'smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio',
    'smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio',
    'smsp__average_warps_issue_stalled_wait_per_issue_active.ratio',
]


def short(name, n=90):
    name = name.replace('fiode::', '')
    return name if len(name) <= n else name[:n] + '...'


def launches(path):
    rows = []
    with open(path, newline='') as f:
        lines = [l for l in f if l.startswith('"')]
    rd = csv.DictReader(io.StringIO(''.join(lines)))
    for r in rd:
        if r.get('Metric Name') != 'gpu__time_duration.sum':
            continue
        v = float(r['Metric Value'].replace(',', ''))
        unit = r.get('Metric Unit', 'ns')
        v_us = v / 1e3 if unit in ('ns', 'nsecond') else v if unit in ('us', 'usecond') else v * 1e3 if unit in ('ms', 'msecond') else v
        rows.append((r['Kernel Name'], v_us))
    agg = OrderedDict()
    for k, v in rows:
        a = agg.setdefault(k, [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(a[1] for a in agg.values())
    print('| kernel | launches | total us | share | avg us |\n|---|---|---|---|---|')
    for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f'| `{short(k)}` | {n} | {t:.1f} | {100 * t / tot:.1f}% | {t / n:.1f} |')
    print(f'\n{len(rows)} launches, {tot / 1e3:.2f} ms in total.')


def full(path):
    out = subprocess.run(['ncu', '-i', path, '--page', 'raw', '--csv'], capture_output=True, text=True, check=True).stdout
    lines = [l for l in out.splitlines(True) if l.startswith('"')]
    rd = list(csv.reader(io.StringIO(''.join(lines))))
    head, units, data = rd[0], rd[1], rd[2:]
    col = {h: i for i, h in enumerate(head)}
    kn = col['Kernel Name']
    seen = OrderedDict()
    for r in data:
        seen.setdefault(r[kn], []).append(r)
    for k, rs in seen.items():
        r = rs[-1]
        print(f'## `{short(k, 140)}`  ({len(rs)} launch{"es" if len(rs) > 1 else ""} captured; last one shown)\n')
        print('| metric | value | unit |\n|---|---|---|')
        for m in METRICS:
            if m in col:
                print(f'| {m} | {r[col[m]]} | {units[col[m]]} |')
        print()


def traffic(path):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of every fiode kernel in the capture (tools/prof_kernels.py:
    one 4096-image backbone pass + the 16-solve dopri5 launch + one certify call), as JSON keyed by the kernel name (a list per kernel, in launch order)."""
    import json
    import os
    out = subprocess.run(['ncu', '-i', path, '--page', 'raw', '--csv'], capture_output=True, text=True, check=True).stdout
    lines = [l for l in out.splitlines(True) if l.startswith('"')]
    rd = list(csv.reader(io.StringIO(''.join(lines))))
    head, units, data = rd[0], rd[1], rd[2:]
    col = {h: i for i, h in enumerate(head)}
    scale = {'byte': 1.0, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}
    tab = OrderedDict()
    for r in data:
        name = r[col['Kernel Name']]
        if name.startswith('void at::') or name.startswith('at::'):
            continue
        by = sum(float(r[col[m]].replace(',', '')) * scale[units[col[m]]] for m in ('dram__bytes_read.sum', 'dram__bytes_write.sum'))
        t = float(r[col['gpu__time_duration.sum']].replace(',', ''))
        tu = units[col['gpu__time_duration.sum']]
        t_us = t * {'ns': 1e-3, 'nsecond': 1e-3, 'us': 1.0, 'usecond': 1.0, 'ms': 1e3, 'msecond': 1e3}.get(tu, 1.0)
        key = re.sub(r'\(.*$', '', name.replace('fiode::', '').replace('void ', ''))
        tab.setdefault(key, []).append(dict(dram_bytes=by, us_under_ncu=t_us, grid=int(r[col['launch__grid_size']].replace(',', ''))))
    print(json.dumps(dict(source=os.path.basename(os.path.dirname(os.path.abspath(path))) + '/' + os.path.basename(path),
                          capture='tools/prof_kernels.py all: certify grid_20 x 296 images; one step of 16 batch-256 solves (4096 images)',
                          backbone_images=4096, kernels=tab), indent=1))


if __name__ == '__main__':
    {'launches': launches, 'full': full, 'traffic': traffic}[sys.argv[1]](sys.argv[2])
