set -x
cd $GRAFT_REPO_ROOT
(time python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -5) > gpurun_out/r02u_smoke.log 2>&1
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29561 bench.py --gpus 2 --steps 5 --warmup 3 --legs c5 > gpurun_out/r02u_bench_n2_c5.json 2> gpurun_out/r02u_bench_n2_c5.err
python bench.py --steps 5 --warmup 3 > gpurun_out/r02u_bench_n1.json 2> gpurun_out/r02u_bench_n1.err
echo done
