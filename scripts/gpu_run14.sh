set -x
cd $GRAFT_REPO_ROOT
(time python -m pytest tests/test_gpu_map2d.py tests/test_gpu_sharded.py tests/test_gpu_score.py tests/test_gpu_resample.py -m gpu -x -q 2>&1 | tail -15) > gpurun_out/r02n_pytest.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r02n_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --legs none > gpurun_out/r02n_ncu.log 2>&1
echo done
