# the reference's node sources relinked against the host mirror + CUDA library (oracle/_ref/libamcl_node_gpu.so), the
# host-drawn multinomial mode
set -x
cd $GRAFT_REPO_ROOT
(time python -m pytest tests/test_node_chain.py tests/test_gpu_resample.py tests/test_host_cpp.py -m gpu -x -q 2>&1 | tail -40) > gpurun_out/relinked_node_pytest.log 2>&1
echo done
