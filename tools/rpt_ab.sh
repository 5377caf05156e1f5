B="python bench.py --no-cpu-baseline --map-rows 0 --latency-iters 0 --stereo-pairs 0 --search-frames 0 --c4-frames 0"
for r in 16 12 8 6 4; do
ORBFE_RESIZE_RPT=$r $B 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('rpt $r', round(d['value']), round(d['e2e']['value']), d['stages']['pyramid']['ms_per_step'])
"
done
