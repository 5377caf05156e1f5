#!/bin/bash
# GPU-box check used during kernel work: parity tests, then the C2 bench without the slow legs (both FAST kernels).
python -m pytest tests -m gpu -x -q 2>&1 | tail -15
B="python bench.py --no-cpu-baseline --map-rows 0 --latency-iters 0 --stereo-pairs 0 --search-frames 0 --c4-frames 0"
show() { python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print(d['value'], d['e2e']['value'], {k: v['ms_per_step'] for k, v in d['stages'].items()})
"; }
$B 2>&1 | show
for v in "$@"; do echo "== $v"; env $v $B 2>&1 | show; done
