#!/bin/bash
# usage: tools/k2_matrix.sh <workload> <rows> "ENV1=.. ENV2=.." "..."   (each arg = one configuration)
w=$1; rows=$2; shift 2
for cfg in "$@"; do
  printf "%-60s " "$cfg"
  env $cfg timeout 90 python bench.py --workload $w --rows $rows $EXTRA --steps 5 --no-cpu-baseline --no-e2e 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        j=json.loads(l); r=j['roofline']; print('%.3f ms  issued_frac %.3f  frac %.3f %s  variant %s' % (r['kernel_ms'], r.get('issued_frac',0), r['frac'], r['bound'], j['config']['search_variant']))
    elif 'rror' in l: print(l.strip())
"
done
