#!/bin/bash
# tools/ab_variants.sh "<variant dirs>" "<scenes>" [spp] -- same-box comparison of library builds kept under ab/ (see tools/build_variant.sh); "." = this tree
variants=$1; scenes=${2:-"teapot"}; spp=${3:-64}
for round in $(seq 1 ${ROUNDS:-2}); do
  for s in $scenes; do
    for which in $variants; do
      d=$which; [ "$which" != "." ] && d=ab/$which
      timeout 300 python $d/tools/quick_bench.py scenes/$s.yml $spp 1920 1080 ${PREC:-32} 2>&1 | tail -2 | python -c "
import sys,json
for ln in sys.stdin:
    try: d=json.loads(ln)
    except ValueError: print(ln.strip()[:200]); continue
    keep={k:d[k] for k in ('Msamples_s','ms_extend','ms_shade','ms_shadow','ms_total') if k in d}
    print('$which'.ljust(8), '$s'.ljust(12), 'timed' if d.get('timed') else 'plain', keep)
"
    done
  done
done
