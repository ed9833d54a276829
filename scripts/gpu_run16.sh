set -x
cd $GRAFT_REPO_ROOT
(time python -m pytest tests/test_gpu_sharded.py tests/test_gpu_nccl.py tests/test_gpu_map3d.py -m gpu -x -q 2>&1 | tail -15) > gpurun_out/r02p_pytest.log 2>&1
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r02p_bench_n2.json 2> gpurun_out/r02p_bench_n2.err
BAMCL_SHARD_TIMING=1 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 3 --warmup 2 > gpurun_out/r02p_bench_n2_timing.json 2> gpurun_out/r02p_bench_n2_timing.err
echo done
