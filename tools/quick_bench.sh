#!/bin/bash
# usage: tools/quick_bench.sh <workload> [env assignments...]  -> prints ms/frame, Mrays/s, e2e, frac
wl=$1; shift
env "$@" python bench.py ${LAYOUT:+--layout $LAYOUT} --workload $wl --steps 10 --warmup 3 --no-cpu-baseline 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']
print('$wl $*', 'ms=%.4f Mrays/s=%.1f e2e=%.1f frac=%.3f' % (d['ms_per_step'], d['value'], d['e2e']['value'], r['frac']))"
