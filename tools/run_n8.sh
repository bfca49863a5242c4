python -m pytest tests/test_gpu_multi.py -q -m gpu 2>&1 | tail -8 > gpurun_out/r02n_multi_n8.log; cat gpurun_out/r02n_multi_n8.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 8 --warmup 3 --cpu-budget 0 --also none"
$TR --steps 8 --scene environment --shard tiles > gpurun_out/r02n_bench_n8_environment_tiles.json 2> gpurun_out/r02n_n8_env.err; tail -1 gpurun_out/r02n_n8_env.err
$TR --steps 5 --scene doom --shard tiles > gpurun_out/r02n_bench_n8_doom_tiles.json 2> gpurun_out/r02n_n8_doom.err; tail -1 gpurun_out/r02n_n8_doom.err
$TR --steps 8 --scene teapot --scaling strong --spp-frame 256 > gpurun_out/r02n_bench_n8_teapot_strong.json 2> gpurun_out/r02n_n8_ts.err; tail -1 gpurun_out/r02n_n8_ts.err
$TR --steps 5 --scene dragon --scaling strong --spp-frame 256 > gpurun_out/r02n_bench_n8_dragon_strong.json 2> gpurun_out/r02n_n8_ds.err; tail -1 gpurun_out/r02n_n8_ds.err
$TR --steps 10 --scene teapot > gpurun_out/r02n_bench_n8_teapot.json 2> gpurun_out/r02n_n8_t.err; tail -1 gpurun_out/r02n_n8_t.err
for f in gpurun_out/r02n_bench_n8_*.json; do echo $f; cut -c1-200 $f; done
