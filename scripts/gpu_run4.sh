set -x
cd $GRAFT_REPO_ROOT
(time python -m pytest tests -m gpu -x -q 2>&1 | tail -40) > gpurun_out/r02d_pytest.log 2>&1
python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r02d_bench_n1.json 2> gpurun_out/r02d_bench_n1.err
python bench.py --steps 5 --warmup 3 --no-cpu-baseline --particles 8000000 > gpurun_out/r02d_bench_n1_8M.json 2> gpurun_out/r02d_bench_n1_8M.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r02d_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/r02d_ncu.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r02d_launches_8M.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --particles 8000000 > gpurun_out/r02d_ncu8.log 2>&1
echo done
