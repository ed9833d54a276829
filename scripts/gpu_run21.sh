set -x
cd $GRAFT_REPO_ROOT
B="python bench.py --steps 10 --warmup 3 --no-cpu-baseline --legs none"
show() { python -c "
import json,sys; d=json.load(open('$1')); c=d['config']; print('$2', 'up %.3f rs %.3f'%(c['update_ms'], c['resample_ms']), c['kld_leaf_count'], c['clusters'], c['result_hash'])"; }
$B > gpurun_out/t_a.json 2>/dev/null; show gpurun_out/t_a.json default
BAMCL_CC_PER_NODE=1 $B > gpurun_out/t_b.json 2>/dev/null; show gpurun_out/t_b.json per_node
BAMCL_CC_PER_NODE=1 $B --particles 8000000 --steps 4 > gpurun_out/t_c.json 2>/dev/null; show gpurun_out/t_c.json per_node_8M
$B --particles 8000000 --steps 4 > gpurun_out/t_d.json 2>/dev/null; show gpurun_out/t_d.json default_8M
BAMCL_EXTRA_ONLY=kdtree BAMCL_EXTRA_NVCC="-DTOP2_DIV=16" python -c "from badger_amcl_b200 import build; build.build(force=True)" 2>&1 | grep -i error
$B > gpurun_out/t_e.json 2>/dev/null; show gpurun_out/t_e.json div16
$B --particles 8000000 --steps 4 > gpurun_out/t_f.json 2>/dev/null; show gpurun_out/t_f.json div16_8M
BAMCL_EXTRA_ONLY=kdtree BAMCL_EXTRA_NVCC="-DTOP2_DIV=64" python -c "from badger_amcl_b200 import build; build.build(force=True)" 2>&1 | grep -i error
$B > gpurun_out/t_g.json 2>/dev/null; show gpurun_out/t_g.json div64
$B --particles 8000000 --steps 4 > gpurun_out/t_h.json 2>/dev/null; show gpurun_out/t_h.json div64_8M
echo done
