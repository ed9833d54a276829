# GPU tests added with the node pinning: the node-sequence replays, ingestion, global localisation (2-D and 3-D)
set -x
cd $GRAFT_REPO_ROOT
(time python -m pytest tests/test_node_chain.py tests/test_ingest.py tests/test_global_init.py tests/test_readout.py tests/test_gpu_resample.py -m gpu -q 2>&1 | tail -40) > gpurun_out/node_chain_pytest.log 2>&1
echo done
