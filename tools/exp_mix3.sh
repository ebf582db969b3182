#!/bin/bash
# A/B of the fused conv0 kernel (FIODE_MIX3_T=160: 160 threads, two FFTs per row task; 288: one round per phase)
OUT=gpurun_out/${1:-mix3}; mkdir -p $OUT
for S in 160 288; do
  FIODE_MIX3_T=$S timeout 300 python -m pytest tests/test_gpu_parity.py -q -x -k "backbone or ivp_matches or cayley_conv or u8" > $OUT/pytest_$S.log 2>&1; echo "T=$S pytest rc=$? $(tail -1 $OUT/pytest_$S.log)"
  FIODE_MIX3_T=$S timeout 200 python bench.py --steps 60 --warmup 5 --no-cpu-baseline --no-certify --no-fi > $OUT/t_$S.json 2> $OUT/t_$S.err
  python - <<PY
import json
try:
    d = json.loads([l for l in open('$OUT/t_$S.json') if l.startswith('{')][-1])
    print('T', $S, 'value', round(d['value']), 'ms', round(d['ms_per_step'], 3), [(s['stage'][:12], round(s['ms'], 4)) for s in d['roofline']['parts']['stages'] if 'mix 3' in s['stage']])
except Exception as e:
    print('T', $S, 'failed', e)
PY
done
