# last check of the round: smoke() and the default bench line after the pose-source refactor / no-alloc entry point
set -x
cd $GRAFT_REPO_ROOT
(time python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3) > gpurun_out/final_smoke.log 2>&1
(time python bench.py --steps 20 --warmup 3 > gpurun_out/final_bench.json 2> gpurun_out/final_bench.err) > gpurun_out/final_bench.time 2>&1
(time python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/final_bench_ref.json 2> gpurun_out/final_bench_ref.err) > gpurun_out/final_bench_ref.time 2>&1
echo done
