set -x
cd $GRAFT_REPO_ROOT
(time python bench.py --steps 10 --warmup 3 > gpurun_out/r02l_bench.json 2> gpurun_out/r02l_bench.err) > gpurun_out/r02l_bench.time 2>&1
(time python scripts/parity_report.py > gpurun_out/r02l_parity.log 2>&1) > gpurun_out/r02l_parity.time 2>&1
echo done
