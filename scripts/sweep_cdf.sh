#!/bin/bash
# Rebuild with different CDF tile shapes and report bench timings (development aid, run on the GPU box).
cd "$(dirname "$0")/.."
run() {
  echo "=== cfg: $1"
  BAMCL_EXTRA_NVCC="$1" python -c "from badger_amcl_b200 import build; build.build(force=True)" 2>&1 | grep -i error
  python bench.py --steps 10 --no-cpu-baseline | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('step %.3f update %.3f resample %.3f e2e %.3f' % (d['ms_per_step'], d['config']['update_ms'], d['config']['resample_ms'], d['e2e']['ms_per_step']))"
}
run ""
run "-DCDF_BLOCK_THREADS=512"
run "-DCDF_BLOCK_THREADS=1024"
run "-DCDF_BLOCK_THREADS=512 -DCDF_ITEMS_PER_THREAD=4"
run "-DCDF_BLOCK_THREADS=1024 -DCDF_ITEMS_PER_THREAD=4"
run ""
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
