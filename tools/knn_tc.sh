#!/bin/bash
# the tcgen05 kNN kernel (ORBFE_KNN_TC=1): parity tests, then the C2 kNN stage and the map-scale kNN
export ORBFE_KNN_TC=1
timeout 300 python -m pytest tests/test_gpu_matcher.py tests/test_gpu_extractor.py -x -q -k "knn or tensor or large or c2 or ties" 2>&1 | tail -3
timeout 600 python bench.py --no-cpu-baseline --latency-iters 0 --stereo-pairs 0 --search-frames 0 --c4-frames 0 ${BENCH_EXTRA} 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print(d['value'], d['e2e']['value'], {k: round(v['ms_per_step'], 3) for k, v in d['stages'].items()}); m = d.get('knn_map') or {}; print('map', m.get('ms_per_step'), m.get('gpairs_per_s'), m.get('parity_sample'))
"
for so in variants/*.so; do
  [ -f "$so" ] || continue
  echo "== $so"
  ORBFE_SO=$PWD/$so timeout 600 python bench.py --no-cpu-baseline --latency-iters 0 --stereo-pairs 0 --search-frames 0 --c4-frames 0 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print(d['value'], d['e2e']['value'], {k: round(v['ms_per_step'], 3) for k, v in d['stages'].items()}); m = d.get('knn_map') or {}; print('map', m.get('ms_per_step'), m.get('gpairs_per_s'), m.get('parity_sample'))
"
done
