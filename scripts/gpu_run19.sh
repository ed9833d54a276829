set -x
cd $GRAFT_REPO_ROOT
(time python -m pytest tests/test_gpu_resample.py tests/test_gpu_sharded.py tests/test_golden.py -m gpu -x -q 2>&1 | tail -6) > gpurun_out/r02s_pytest.log 2>&1
python bench.py --steps 10 --warmup 3 --no-cpu-baseline --legs none > gpurun_out/r02s_bench_init.json 2>/dev/null
BAMCL_CC_NO_INIT=1 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --legs none > gpurun_out/r02s_bench_noinit.json 2>/dev/null
python bench.py --steps 5 --warmup 3 --no-cpu-baseline --legs none --particles 8000000 > gpurun_out/r02s_bench_init_8M.json 2>/dev/null
BAMCL_CC_NO_INIT=1 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --legs none --particles 8000000 > gpurun_out/r02s_bench_noinit_8M.json 2>/dev/null
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r02s_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --legs none > gpurun_out/r02s_ncu.log 2>&1
echo done
