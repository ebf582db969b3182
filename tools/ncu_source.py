"""Per-instruction view of an `ncu --set full --import-source on` capture: stall samples by code region and the top instructions.
    python tools/ncu_source.py <capture.ncu-rep> [kernel-regex]"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'sass'] + (['-k', 'regex:' + sys.argv[2]] if len(sys.argv) > 2 else []),
                     capture_output=True, text=True).stdout
lines = out.splitlines()
start = next(i for i, l in enumerate(lines) if l.startswith('"Address"'))
rd = list(csv.reader(io.StringIO('\n'.join(lines[start:]))))
head = rd[0]
col = {h: i for i, h in enumerate(head)}
rows = []
for r in rd[1:]:
    if r and r[0] == 'Address':      # several launches in the capture: the first one is shown
        break
    if len(r) == len(head):
        rows.append(r)
S = col['# Samples']
tot = sum(int(r[S] or 0) for r in rows)
print('instructions', len(rows), 'samples', tot)
stalls = [h for h in head if h.startswith('stall_') and 'Not Issued' not in h]
agg = {h: sum(int(r[col[h]] or 0) for r in rows) for h in stalls}
print('stall totals:', ', '.join('%s %.1f%%' % (h[6:], 100 * v / max(tot, 1)) for h, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v > 0.01 * tot))
wf, wfi = col.get('L1 Wavefronts Shared'), col.get('L1 Wavefronts Shared Ideal')
if wf is not None:
    print('shared wavefronts %d ideal %d' % (sum(int(r[wf] or 0) for r in rows), sum(int(r[wfi] or 0) for r in rows)))
print('\n# top instructions by samples')
for r in sorted(rows, key=lambda r: -int(r[S] or 0))[:40]:
    top = sorted(((int(r[col[h]] or 0), h[6:]) for h in stalls), reverse=True)[:2]
    extra = ''
    if wf is not None and r[wf] not in ('', '0'):
        extra = ' wf %s/%s' % (r[wf], r[wfi])
    print('%6.2f%%  %s  %-70s %s%s' % (100 * int(r[S] or 0) / max(tot, 1), r[col['Address']][-5:], r[col['Source']][:70], ' '.join('%s:%d' % (n, v) for v, n in top), extra))
# cumulative profile in 20 chunks of the instruction stream
print('\n# samples by position (5% chunks of the instruction stream)')
n = len(rows)
for k in range(20):
    ch = rows[k * n // 20:(k + 1) * n // 20]
    s = sum(int(r[S] or 0) for r in ch)
    print('%3d%%-%3d%%  %5.1f%%  first: %s' % (5 * k, 5 * k + 5, 100 * s / max(tot, 1), ch[0][col['Source']][:60]))
