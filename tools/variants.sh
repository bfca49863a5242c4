#!/bin/bash
# development probe: extend-kernel variants on two scenes
for v in "$@"; do
  for s in teapot dragon; do
    PATHS_EXTEND_VARIANT=$v python tools/quick_bench.py scenes/$s.yml 16 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('variant', $v, {k:d[k] for k in d if k in ('ms_extend','ms_shade','ms_shadow','Msamples_s','Mrays_s','scene')})"
  done
done
