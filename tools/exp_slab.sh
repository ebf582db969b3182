#!/bin/bash
# backbone slab-size sweep (FIODE_SLAB images per slab) on the headline step; prints value / ms per step
OUT=gpurun_out/${1:-slab}; mkdir -p $OUT
for S in 512 1024 2048 4096 8192; do
  FIODE_SLAB=$S timeout 200 python bench.py --steps 60 --warmup 5 --no-cpu-baseline --no-certify --no-fi > $OUT/slab_$S.json 2> $OUT/slab_$S.err
  python - <<PY
import json
try:
    d = json.loads([l for l in open('$OUT/slab_$S.json') if l.startswith('{')][-1])
    print('slab', $S, 'value', round(d['value']), 'ms', round(d['ms_per_step'], 3), 'e2e', round(d['e2e']['value']), 'bb_ms', round(d['roofline']['parts']['backbone_ms'], 3))
except Exception as e:
    print('slab', $S, 'failed', e)
PY
done
