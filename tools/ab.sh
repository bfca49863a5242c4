#!/bin/bash
# tools/ab.sh "<scenes>" [spp] -- on the GPU box: this tree's library against the round-1 build kept under ab/r01 (same box, alternating runs)
scenes=${1:-"teapot dragon environment"}; spp=${2:-64}
for round in 1 2; do
  for s in $scenes; do
    for which in ab/r01 .; do
      timeout 300 python $which/tools/quick_bench.py scenes/$s.yml $spp 2>&1 | tail -2 | python -c "
import sys,json
for ln in sys.stdin:
    try: d=json.loads(ln)
    except ValueError: print(ln.strip()[:200]); continue
    keep={k:d[k] for k in ('Msamples_s','Mrays_s','ms_extend','ms_shade','ms_shadow','ms_accumulate','ms_total') if k in d}
    print('$which'.ljust(7), '$s'.ljust(12), 'timed' if d.get('timed') else 'plain', keep)
"
    done
  done
done
