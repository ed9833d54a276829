#!/bin/bash
# Rebuild with different refill thresholds of k_score_beam and time C2 (development aid, run on the GPU box).
cd "$(dirname "$0")/.."
for v in 4 16 24 32 8; do
  echo "=== REFILL_MIN=$v"
  BAMCL_EXTRA_NVCC="-DBEAM_REFILL_MIN=$v" python -c "from badger_amcl_b200 import build; build.build(force=True)" 2>&1 | grep -i error
  python scripts/time_configs.py c2 2>&1 | grep "^C2" | cut -c1-120
done
