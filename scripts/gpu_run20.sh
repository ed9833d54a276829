set -x
cd $GRAFT_REPO_ROOT
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
BAMCL_CC_REPLICATED=1 $TR --nproc-per-node 8 --master-port 29551 bench.py --gpus 8 --steps 20 --warmup 3 > gpurun_out/r02x_bench_n8_ccrep.json 2> gpurun_out/r02x_bench_n8_ccrep.err
BAMCL_CC_REPLICATED=1 $TR --nproc-per-node 8 --master-port 29552 bench.py --gpus 8 --steps 5 --warmup 3 --particles 8000000 > gpurun_out/r02x_bench_n8_64M_ccrep.json 2> gpurun_out/r02x_bench_n8_64M_ccrep.err
BAMCL_CC_REPLICATED=1 $TR --nproc-per-node 4 --master-port 29553 bench.py --gpus 4 --steps 20 --warmup 3 > gpurun_out/r02x_bench_n4_ccrep.json 2> gpurun_out/r02x_bench_n4_ccrep.err
echo done
