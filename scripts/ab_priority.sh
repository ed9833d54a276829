# A/B of the side-stream priority of the resampler's clustering chain (development aid, run on the GPU box)
set -x
cd $GRAFT_REPO_ROOT
B="python bench.py --steps 10 --warmup 3 --no-cpu-baseline --legs none"
show() { python -c "
import json,sys; d=json.load(open('$1')); c=d['config']; print('$2', 'up %.3f rs %.3f e2e %.3f'%(c['update_ms'], c['resample_ms'], d['e2e']['ms_per_step']), c['kld_leaf_count'], c['result_hash'])"; }
for p in 1000000 2000000 3000000 8000000; do
$B --particles $p --steps 6 > gpurun_out/t_a.json 2>/dev/null; show gpurun_out/t_a.json auto_$p
BAMCL_PRIO_MODE=low $B --particles $p --steps 6 > gpurun_out/t_a.json 2>/dev/null; show gpurun_out/t_a.json low_$p
BAMCL_PRIO_MODE=flat $B --particles $p --steps 6 > gpurun_out/t_a.json 2>/dev/null; show gpurun_out/t_a.json flat_$p
done
echo done
