set -x
cd $GRAFT_REPO_ROOT
(time python -m pytest tests/test_gpu_resample.py tests/test_golden.py tests/test_global_init.py tests/test_readout.py -m gpu -x -q 2>&1 | tail -6) > gpurun_out/r02v_pytest.log 2>&1
B="python bench.py --steps 10 --warmup 3 --no-cpu-baseline --legs none"
show() { python -c "
import json,sys; d=json.load(open('$1')); c=d['config']; print('$2', 'up %.3f rs %.3f'%(c['update_ms'], c['resample_ms']), c['kld_leaf_count'], c['clusters'], c['result_hash'])"; }
$B > gpurun_out/t_a.json 2>/dev/null; show gpurun_out/t_a.json tiles
BAMCL_CC_NO_TILES=1 $B > gpurun_out/t_b.json 2>/dev/null; show gpurun_out/t_b.json no_tiles
$B --particles 8000000 --steps 4 > gpurun_out/t_c.json 2>/dev/null; show gpurun_out/t_c.json tiles_8M
BAMCL_CC_NO_TILES=1 $B --particles 8000000 --steps 4 > gpurun_out/t_d.json 2>/dev/null; show gpurun_out/t_d.json no_tiles_8M
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r02v_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --legs none > gpurun_out/r02v_ncu.log 2>&1
echo done
