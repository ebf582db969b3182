#!/bin/bash
# A/B of the mixed-spectrum layout of model.4 (FIODE_YF4=1: [phi][b][co]; default [b][phi][co])
OUT=gpurun_out/${1:-yf4}; mkdir -p $OUT
for S in 1 2; do
  FIODE_YF4=$S timeout 300 python -m pytest tests/test_gpu_parity.py -q -x -k "backbone or ivp_matches" > $OUT/pytest_$S.log 2>&1; echo "YF4=$S pytest rc=$? $(tail -1 $OUT/pytest_$S.log)"
  FIODE_YF4=$S timeout 200 python bench.py --steps 60 --warmup 5 --no-cpu-baseline --no-certify --no-fi > $OUT/t_$S.json 2> $OUT/t_$S.err
  python - <<PY
import json
try:
    d = json.loads([l for l in open('$OUT/t_$S.json') if l.startswith('{')][-1])
    print('YF4', $S, 'value', round(d['value']), 'ms', round(d['ms_per_step'], 3), [(s['stage'][:12], round(s['ms'], 4)) for s in d['roofline']['parts']['stages'] if 'mixing' in s['stage'] or 'irfft2 16' in s['stage']])
except Exception as e:
    print('YF4', $S, 'failed', e)
PY
done
