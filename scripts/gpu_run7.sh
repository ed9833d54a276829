set -x
cd $GRAFT_REPO_ROOT
(time python -m pytest tests/test_gpu_map3d.py tests/test_golden.py -m gpu -x -q 2>&1 | tail -30) > gpurun_out/r02g_pytest.log 2>&1
python scripts/time_configs.py c4 > gpurun_out/r02g_configs.log 2>&1
python scripts/gather_peak.py > gpurun_out/r02g_gather_peak.log 2>&1
echo done
