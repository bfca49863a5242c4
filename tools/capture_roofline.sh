#!/bin/bash
# tools/capture_roofline.sh [scene ...] -- regenerates profiles/traffic.json and profiles/issue.json for the CURRENT code from one ncu pass
# per scene.  Run from the container (it goes through gpurun and stamps the files with `git rev-parse HEAD` and a hash of the kernel
# sources; bench.py prints the ncu-derived fields only when that hash matches the code it runs).
#   tools/capture_roofline.sh teapot dragon
set -e
cd "$(dirname "$0")/.."
HEAD=$(git rev-parse --short=12 HEAD)
if ! git diff --quiet -- paths_b200/csrc; then HEAD="$HEAD+dirty"; fi
SCENES="${*:-teapot dragon}"
TAG=${TAG:-r02}
CMD=""
for s in $SCENES; do
  M="gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,smsp__thread_inst_executed_per_inst_executed.ratio"
  CMD="$CMD python tools/roofline_probe.py $s > gpurun_out/${TAG}_probe_$s.json 2> gpurun_out/${TAG}_probe_$s.err && "
  CMD="$CMD W=\$(python -c \"import json;print(json.loads(open('gpurun_out/${TAG}_probe_$s.json').read().strip().splitlines()[-1])['warmup_launches'])\") && "
  CMD="$CMD N=\$(python -c \"import json;print(json.loads(open('gpurun_out/${TAG}_probe_$s.json').read().strip().splitlines()[-1])['launches_per_pass'])\") && "
  CMD="$CMD ncu --metrics $M --clock-control none -k regex:k_ -s \$W -c \$N --csv --log-file gpurun_out/${TAG}_roofline_$s.csv python tools/roofline_probe.py $s > gpurun_out/${TAG}_probe_ncu_$s.log 2>&1 ; "
done
tools/gpu_retry.sh 900 "$CMD ls -la gpurun_out | tail -20"
for s in $SCENES; do
  python tools/ncu_roofline.py $s gpurun_out/${TAG}_roofline_$s.csv gpurun_out/${TAG}_probe_$s.json "$HEAD"
  cp gpurun_out/${TAG}_roofline_$s.csv profiles/${TAG}_roofline_$s.csv
done
