# last multi-GPU sanity of the round on 2 GPUs: real-NCCL parity tests + the 2-rank bench line (both arms' launch as the driver does)
set -x
cd $GRAFT_REPO_ROOT
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
(time python -m pytest tests/test_gpu_nccl.py -m gpu -x -q 2>&1 | tail -5) > gpurun_out/final2_nccl_pytest.log 2>&1
$TR --nproc-per-node 2 --master-port 29551 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/final2_bench_n2.json 2> gpurun_out/final2_bench_n2.err
echo done
