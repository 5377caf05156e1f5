#!/bin/bash
# A/B of the two kNN kernels: parity tests, then the C2 kNN stage and the map-scale kNN with each.
python -m pytest tests/test_gpu_matcher.py tests/test_gpu_sharded.py -x -q 2>&1 | tail -5
for v in 0 1; do
  echo "== ORBFE_KNN_MMA=$v"
  ORBFE_KNN_MMA=$v python bench.py --no-cpu-baseline --latency-iters 0 --stereo-pairs 0 --search-frames 0 --c4-frames 0 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print(d['value'], d['e2e']['value'], {k: round(v['ms_per_step'], 3) for k, v in d['stages'].items()}); m = d.get('knn_map') or {}; print('map', m.get('ms_per_step'), m.get('gpairs_per_s'), m.get('parity_sample'))
"
done
