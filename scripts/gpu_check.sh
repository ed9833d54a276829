# full single-GPU check on the GPU box: smoke, GPU tests, the bench line with all legs
set -x
cd $GRAFT_REPO_ROOT
(time python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3) > gpurun_out/check_smoke.log 2>&1
(time python -m pytest tests -m gpu -x -q 2>&1 | tail -8) > gpurun_out/check_pytest.log 2>&1
(time python bench.py --steps 20 --warmup 3 --legs none > gpurun_out/check_bench.json 2> gpurun_out/check_bench.err) > gpurun_out/check_bench.time 2>&1
echo done
