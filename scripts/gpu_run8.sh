set -x
cd $GRAFT_REPO_ROOT
(time python -m pytest tests/test_gpu_map3d.py -m gpu -x -q 2>&1 | tail -30) > gpurun_out/r02h_pytest.log 2>&1
python scripts/time_configs.py c4 > gpurun_out/r02h_c4_default.log 2>&1
BAMCL_NO_SCORE_ORDER=1 python scripts/time_configs.py c4 > gpurun_out/r02h_c4_noorder.log 2>&1
BAMCL_CLOUD_BITS=1 python scripts/time_configs.py c4 > gpurun_out/r02h_c4_bits.log 2>&1
echo done
