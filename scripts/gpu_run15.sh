set -x
cd $GRAFT_REPO_ROOT
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
BAMCL_SHARD_TIMING=1 $TR --nproc-per-node 8 --master-port 29522 bench.py --gpus 8 --steps 4 --warmup 2 > gpurun_out/r02o_bench_n8_timing.json 2> gpurun_out/r02o_bench_n8_timing.err
BAMCL_SHARD_TIMING=1 $TR --nproc-per-node 8 --master-port 29524 bench.py --gpus 8 --steps 2 --warmup 2 --particles 8000000 > gpurun_out/r02o_bench_n8_64M_timing.json 2> gpurun_out/r02o_bench_n8_64M_timing.err
echo done
