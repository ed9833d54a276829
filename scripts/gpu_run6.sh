set -x
cd $GRAFT_REPO_ROOT
nvidia-smi -L | wc -l
(time python -m pytest tests/test_gpu_nccl.py -m gpu -x -q 2>&1 | tail -30) > gpurun_out/r02f_nccl8_pytest.log 2>&1
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
$TR --nproc-per-node 8 --master-port 29521 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/r02f_bench_n8.json 2> gpurun_out/r02f_bench_n8.err
BAMCL_SHARD_TIMING=1 $TR --nproc-per-node 8 --master-port 29522 bench.py --gpus 8 --steps 3 --warmup 2 > gpurun_out/r02f_bench_n8_timing.json 2> gpurun_out/r02f_bench_n8_timing.err
$TR --nproc-per-node 8 --master-port 29523 bench.py --gpus 8 --steps 4 --warmup 3 --particles 8000000 > gpurun_out/r02f_bench_n8_64M.json 2> gpurun_out/r02f_bench_n8_64M.err
BAMCL_SHARD_TIMING=1 $TR --nproc-per-node 8 --master-port 29524 bench.py --gpus 8 --steps 2 --warmup 2 --particles 8000000 > gpurun_out/r02f_bench_n8_64M_timing.json 2> gpurun_out/r02f_bench_n8_64M_timing.err
$TR --nproc-per-node 4 --master-port 29525 bench.py --gpus 4 --steps 10 --warmup 3 > gpurun_out/r02f_bench_n4.json 2> gpurun_out/r02f_bench_n4.err
python scripts/gather_peak.py > gpurun_out/r02f_gather_peak.log 2>&1
echo done
