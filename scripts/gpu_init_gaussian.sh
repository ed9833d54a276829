# initWithGaussian on the device + the refactored motion update
set -x
cd $GRAFT_REPO_ROOT
(time python -m pytest tests/test_node_chain.py tests/test_host_cpp.py tests/test_motion.py tests/test_gpu_sharded.py -m gpu -x -q 2>&1 | tail -30) > gpurun_out/init_gaussian_pytest.log 2>&1
echo done
