#!/bin/bash
# Rebuild score_cloud.cu with different k_score_cloud knobs and time config 4 (development aid, run on the GPU box).
set -u
cd "$(dirname "$0")/.."
run() {
  echo "=== cfg: $1  env: ${2:-}"
  BAMCL_EXTRA_ONLY=score_cloud BAMCL_EXTRA_NVCC="$1" python -c "from badger_amcl_b200 import build; build.build(force=True)" 2>&1 | grep -i error
  env ${2:-X=1} python scripts/time_configs.py c4 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read())['c4_cloud']; print('global %.2f ms tracking %.2f ms' % (d['global']['ms'], d['tracking']['ms']))"
}
run "-DCLOUD_PB=16 -DCLOUD_NPL=2 -DCLOUD_GB=8 -DCLOUD_MINB=4"
run "-DCLOUD_PB=16 -DCLOUD_NPL=2 -DCLOUD_GB=4 -DCLOUD_MINB=4"
run "-DCLOUD_PB=8 -DCLOUD_NPL=2 -DCLOUD_GB=8 -DCLOUD_MINB=4"
run "-DCLOUD_PB=8 -DCLOUD_NPL=2 -DCLOUD_GB=4 -DCLOUD_MINB=5"
run "-DCLOUD_PB=16 -DCLOUD_NPL=1 -DCLOUD_GB=16 -DCLOUD_MINB=4"
run "-DCLOUD_PB=16 -DCLOUD_NPL=2 -DCLOUD_GB=16 -DCLOUD_MINB=3"
run "-DCLOUD_PB=16 -DCLOUD_NPL=2 -DCLOUD_GB=8 -DCLOUD_MINB=3" "BAMCL_CLOUD_BITS=1"
run "-DCLOUD_PB=16 -DCLOUD_NPL=2 -DCLOUD_GB=8 -DCLOUD_MINB=3"
