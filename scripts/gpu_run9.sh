set -x
cd $GRAFT_REPO_ROOT
(time python -m pytest tests/test_gpu_score.py tests/test_golden.py tests/test_host_cpp.py -m gpu -x -q 2>&1 | tail -30) > gpurun_out/r02i_pytest.log 2>&1
python scripts/time_configs.py c2 > gpurun_out/r02i_c2.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_score_cloud -c 1 -f -o gpurun_out/r02i_cloud python scripts/time_configs.py c4 > gpurun_out/r02i_ncu_cloud.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_score_beam -c 1 -f -o gpurun_out/r02i_beam python scripts/time_configs.py c2 > gpurun_out/r02i_ncu_beam.log 2>&1
echo done
