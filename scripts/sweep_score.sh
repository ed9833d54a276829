#!/bin/bash
# Rebuild libbamcl_cuda.so with different k_score_lf knobs and time the update (development aid, run on the GPU box).
set -u
cd "$(dirname "$0")/.."
run() {
  echo "=== cfg: $1  env: ${2:-}"
  BAMCL_EXTRA_NVCC="$1" python -c "from badger_amcl_b200 import build; build.build(force=True)" 2>&1 | grep -i error
  env ${2:-X=1} python scripts/time_score.py 2>&1 | grep "update"
}
run ""
run "" "BAMCL_SCORE_CTAS=2"
run "-DSCORE_MINB=3" "BAMCL_SCORE_CTAS=3"
run "-DSCORE_PB=8 -DSCORE_MINB=3" "BAMCL_SCORE_CTAS=3"
run "-DSCORE_GB=8"
run "-DSCORE_BPL=8 -DSCORE_GB=2 -DSCORE_MINB=2"
run "-DSCORE_BPL=2 -DSCORE_GB=8"
run "-DSCORE_WARPS=16 -DSCORE_MINB=1" "BAMCL_SCORE_CTAS=1"
run "-DSCORE_WARPS=4" "BAMCL_SCORE_CTAS=8"
run ""   # leave the default build behind
