set -x
cd $GRAFT_REPO_ROOT
(time python -m pytest tests -m gpu -x -q 2>&1 | tail -8) > gpurun_out/r02z_pytest.log 2>&1
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02z_bench_ref.json 2> gpurun_out/r02z_bench_ref.err
python bench.py --steps 20 --warmup 3 > gpurun_out/r02z_bench_n1.json 2> gpurun_out/r02z_bench_n1.err
python bench.py --steps 5 --warmup 3 --no-cpu-baseline --legs none --particles 8000000 > gpurun_out/r02z_bench_n1_8M.json 2> gpurun_out/r02z_bench_n1_8M.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/r02z_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --legs none > gpurun_out/r02z_ncu.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_score_lf|k_cc_union_dense|k_cdf_resolve|k_fast_bucket_warp|k_labels_stats" -c 10 -f -o gpurun_out/r02z_top python bench.py --steps 1 --warmup 1 --no-cpu-baseline --legs none > gpurun_out/r02z_ncu_full.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_score_cloud -c 1 -f -o gpurun_out/r02z_cloud python scripts/time_configs.py c4 > gpurun_out/r02z_ncu_cloud.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_score_beam -c 1 -f -o gpurun_out/r02z_beam python scripts/time_configs.py c2 > gpurun_out/r02z_ncu_beam.log 2>&1
echo done
