#!/bin/bash
# Rebuild score_beam.cu with different k_score_beam knobs and time config 2 (development aid, run on the GPU box).
set -u
cd "$(dirname "$0")/.."
run() {
  echo "=== cfg: $1"
  BAMCL_EXTRA_ONLY=score_beam BAMCL_EXTRA_NVCC="$1" python -c "from badger_amcl_b200 import build; build.build(force=True)" 2>&1 | grep -i error
  python scripts/time_configs.py c2 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read())['c2_beam']; print('%.2f ms' % d['ms'])"
}
run "-DBEAM_CHUNK=8 -DBEAM_REFILL_MIN=8"
run "-DBEAM_CHUNK=4 -DBEAM_REFILL_MIN=8"
run "-DBEAM_CHUNK=16 -DBEAM_REFILL_MIN=8"
run "-DBEAM_CHUNK=4 -DBEAM_REFILL_MIN=16"
run "-DBEAM_CHUNK=8 -DBEAM_REFILL_MIN=16"
run "-DBEAM_CHUNK=8 -DBEAM_REFILL_MIN=4"
run "-DBEAM_CHUNK=2 -DBEAM_REFILL_MIN=12"
run "-DBEAM_CHUNK=8 -DBEAM_REFILL_MIN=8 -DBEAM_MINB=6"
run "-DBEAM_CHUNK=8 -DBEAM_REFILL_MIN=8 -DBEAM_MINB=8"
run "-DBEAM_CHUNK=8 -DBEAM_REFILL_MIN=8"
