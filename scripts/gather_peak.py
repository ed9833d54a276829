"""L2-resident random-gather peak: 2-byte random gathers over tables from 1 MB to 1 GB. One 32-B sector
moves per gather, so sectors/s = gathers/s. Output: JSON for profiles/."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from badger_amcl_b200 import cabi  # noqa: E402

ctx = cabi.Context(0)
out = {}
for mb in (1, 8, 32, 64, 96, 128, 256, 1024):
    r = ctx.gather_rate(mb << 20, 1 << 31)
    out["%dMB" % mb] = {"gathers_per_s": r, "sector_GBps": r * 32 / 1e9}
    print(mb, "MB", "%.3e gathers/s" % r, "%.0f GB/s of sectors" % (r * 32 / 1e9), flush=True)
print(json.dumps(out))
