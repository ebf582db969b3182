#!/bin/bash
# A/B of the row-major GEMM epilogue (FIODE_GEMM_ST=0 thread-per-row float4 stores, 1 transposed whole-row stores)
OUT=gpurun_out/${1:-gemmst}; mkdir -p $OUT
for S in 0 1; do
  FIODE_GEMM_ST=$S timeout 300 python -m pytest tests/test_gpu_parity.py -q -x -k "tensor_core or backbone or ivp_matches" > $OUT/pytest_st$S.log 2>&1; echo "st=$S pytest rc=$?"
  FIODE_GEMM_ST=$S timeout 200 python bench.py --steps 60 --warmup 5 --no-cpu-baseline --no-certify --no-fi > $OUT/st_$S.json 2> $OUT/st_$S.err
  python - <<PY
import json
try:
    d = json.loads([l for l in open('$OUT/st_$S.json') if l.startswith('{')][-1])
    print('st', $S, 'value', round(d['value']), 'ms', round(d['ms_per_step'], 3), [(s['stage'][:12], round(s['ms'], 4)) for s in d['roofline']['parts']['stages'] if 'mixing' in s['stage'] or 'linear' in s['stage']])
except Exception as e:
    print('st', $S, 'failed', e)
PY
done
