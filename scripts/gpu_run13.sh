set -x
cd $GRAFT_REPO_ROOT
(time python -m pytest tests -m gpu -x -q 2>&1 | tail -15) > gpurun_out/r02m_pytest.log 2>&1
(time python bench.py --steps 10 --warmup 3 > gpurun_out/r02m_bench.json 2> gpurun_out/r02m_bench.err) > gpurun_out/r02m_bench.time 2>&1
echo done
