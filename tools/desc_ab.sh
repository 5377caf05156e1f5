#!/bin/bash
B="python bench.py --no-cpu-baseline --map-rows 0 --latency-iters 0 --stereo-pairs 0 --search-frames 0 --c4-frames 0 --no-nf1000 --no-two-handles"
show() { python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print(d['value'], d['e2e']['value'], {k: v['ms_per_step'] for k, v in d['stages'].items()})
"; }
for v in "$@"; do echo "== $v"; env $v timeout 300 $B 2>&1 | show; done
