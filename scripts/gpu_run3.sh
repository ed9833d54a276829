set -x
cd $GRAFT_REPO_ROOT
nvidia-smi -L
(time python -m pytest tests/test_gpu_nccl.py -m gpu -x -q 2>&1 | tail -40) > gpurun_out/r02c_nccl_pytest.log 2>&1
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r02c_bench_n2.json 2> gpurun_out/r02c_bench_n2.err
python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r02c_bench_n1.json 2> gpurun_out/r02c_bench_n1.err
python bench.py --steps 5 --warmup 3 --no-cpu-baseline --particles 2000000 > gpurun_out/r02c_bench_n1_2M.json 2> gpurun_out/r02c_bench_n1_2M.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r02c_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/r02c_ncu.log 2>&1
echo done
