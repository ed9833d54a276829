set -x
cd $GRAFT_REPO_ROOT
python scripts/write_bench_input.py /tmp/bench_in_2M.bin 2000000
python -m torch.distributed.run --no-python --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 badger_amcl_b200/host/test/bench_dist /tmp/bench_in_2M.bin 10 3 > gpurun_out/r02r_cpp_dist_n2.json 2> gpurun_out/r02r_cpp_dist_n2.err
python scripts/time_configs.py c2 > gpurun_out/r02r_c2.log 2>&1
echo done
