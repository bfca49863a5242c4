#!/bin/bash
# tools/run_multi.sh N TAG -- on an N-GPU box: the multi-GPU tests, then the bench lines DESIGN.md quotes at N GPUs (weak and strong
# scaling on teapot, the tile-sharded configs), written to gpurun_out/TAG_*.  Run through `tools/gpu_retry.sh --gpus N`.
N=${1:-8}; TAG=${2:-r02zz}
cd "$(dirname "$0")/.."
python -m pytest tests/test_gpu_multi.py -q -m gpu -p no:cacheprovider 2>&1 | tail -8 > gpurun_out/${TAG}_multi_n$N.log; cat gpurun_out/${TAG}_multi_n$N.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N --warmup 3 --cpu-budget 0 --also none"
$TR --steps 10 --scene teapot > gpurun_out/${TAG}_bench_n${N}_teapot.json 2> gpurun_out/${TAG}_n${N}_t.err; tail -1 gpurun_out/${TAG}_n${N}_t.err
$TR --steps 8 --scene teapot --scaling strong --spp-frame 256 --no-strong > gpurun_out/${TAG}_bench_n${N}_teapot_strong.json 2> gpurun_out/${TAG}_n${N}_ts.err; tail -1 gpurun_out/${TAG}_n${N}_ts.err
$TR --steps 8 --scene environment --shard tiles --no-strong > gpurun_out/${TAG}_bench_n${N}_environment_tiles.json 2> gpurun_out/${TAG}_n${N}_env.err; tail -1 gpurun_out/${TAG}_n${N}_env.err
$TR --steps 5 --scene doom --shard tiles --no-strong > gpurun_out/${TAG}_bench_n${N}_doom_tiles.json 2> gpurun_out/${TAG}_n${N}_doom.err; tail -1 gpurun_out/${TAG}_n${N}_doom.err
for f in gpurun_out/${TAG}_bench_n${N}_*.json; do echo $f; cut -c1-160 $f; done
