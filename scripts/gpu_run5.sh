set -x
cd $GRAFT_REPO_ROOT
(time python -m pytest tests/test_gpu_resample.py tests/test_gpu_sharded.py tests/test_golden.py tests/test_global_init.py -m gpu -x -q 2>&1 | tail -30) > gpurun_out/r02e_pytest.log 2>&1
python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r02e_bench_n1.json 2> gpurun_out/r02e_bench_n1.err
python bench.py --steps 5 --warmup 3 --no-cpu-baseline --particles 8000000 > gpurun_out/r02e_bench_n1_8M.json 2> gpurun_out/r02e_bench_n1_8M.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r02e_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/r02e_ncu.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r02e_launches_8M.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --particles 8000000 > gpurun_out/r02e_ncu8.log 2>&1
python scripts/gather_peak.py > gpurun_out/r02e_gather_peak.log 2>&1
echo done
