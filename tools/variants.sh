#!/bin/bash
# development probe: extend-kernel variants (0 plain batches, 3 two-wide refill, 16 four-wide speculative) on some scenes.
# usage: variants.sh "<scenes>" v1 v2 ...   (env PATHS_SHADE_VARIANT / PATHS_PIPELINE pass through)
scenes=$1; shift
for v in "$@"; do
  for s in $scenes; do
    PATHS_EXTEND_VARIANT=$v timeout 90 python tools/quick_bench.py scenes/$s.yml 16 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('variant', $v, {k:d[k] for k in d if k in ('ms_extend','ms_shade','ms_shadow','ms_total','Msamples_s','Mrays_s','scene')})"
  done
done
