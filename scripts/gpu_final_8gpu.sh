set -x
cd $GRAFT_REPO_ROOT
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
python bench.py --gpus 1 --steps 20 --warmup 3 --legs none > gpurun_out/r02y_bench_n1.json 2> gpurun_out/r02y_bench_n1.err
$TR --nproc-per-node 2 --master-port 29541 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/r02y_bench_n2.json 2> gpurun_out/r02y_bench_n2.err
$TR --nproc-per-node 4 --master-port 29542 bench.py --gpus 4 --steps 20 --warmup 3 > gpurun_out/r02y_bench_n4.json 2> gpurun_out/r02y_bench_n4.err
$TR --nproc-per-node 8 --master-port 29543 bench.py --gpus 8 --steps 20 --warmup 3 > gpurun_out/r02y_bench_n8.json 2> gpurun_out/r02y_bench_n8.err
$TR --nproc-per-node 8 --master-port 29544 bench.py --gpus 8 --steps 5 --warmup 3 --particles 8000000 > gpurun_out/r02y_bench_n8_64M.json 2> gpurun_out/r02y_bench_n8_64M.err
BAMCL_SHARD_TIMING=1 $TR --nproc-per-node 8 --master-port 29545 bench.py --gpus 8 --steps 3 --warmup 2 > gpurun_out/r02y_timing_n8.json 2> gpurun_out/r02y_timing_n8.err
BAMCL_SHARD_TIMING=1 $TR --nproc-per-node 8 --master-port 29546 bench.py --gpus 8 --steps 2 --warmup 2 --particles 8000000 > gpurun_out/r02y_timing_n8_64M.json 2> gpurun_out/r02y_timing_n8_64M.err
python scripts/write_bench_input.py /tmp/bench_in_8M.bin 8000000
$TR --no-python --nproc-per-node 8 --master-port 29547 badger_amcl_b200/host/test/bench_dist /tmp/bench_in_8M.bin 10 3 > gpurun_out/r02y_cpp_dist_n8.json 2> gpurun_out/r02y_cpp_dist_n8.err
(time python -m pytest tests/test_gpu_nccl.py -m gpu -x -q 2>&1 | tail -5) > gpurun_out/r02y_nccl8_pytest.log 2>&1
echo done
