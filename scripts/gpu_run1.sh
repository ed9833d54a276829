set -x
cd $GRAFT_REPO_ROOT
nvidia-smi -L | head -2
(time python -m pytest tests -m gpu -x -q 2>&1 | tail -5) > gpurun_out/r02a_pytest.log 2>&1
python bench.py --steps 10 --warmup 3 > gpurun_out/r02a_bench.json 2> gpurun_out/r02a_bench.err
(timeout 700 compute-sanitizer --tool memcheck --error-exitcode 9 python -m pytest tests/test_golden.py -m gpu -x -q 2>&1 | tail -40) > gpurun_out/r02a_memcheck.log 2>&1
(timeout 700 compute-sanitizer --tool racecheck --error-exitcode 9 python -m pytest tests/test_golden.py -m gpu -x -q 2>&1 | tail -60) > gpurun_out/r02a_racecheck.log 2>&1
python scripts/time_configs.py c2 c4 > gpurun_out/r02a_configs.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_score_beam -c 1 -f -o gpurun_out/r02a_beam python scripts/time_configs.py c2 > gpurun_out/r02a_ncu_beam.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_score_cloud -c 1 -f -o gpurun_out/r02a_cloud python scripts/time_configs.py c4 > gpurun_out/r02a_ncu_cloud.log 2>&1
echo done
