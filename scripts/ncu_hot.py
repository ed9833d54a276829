"""Hot SASS instructions of one kernel from an .ncu-rep (source page): python scripts/ncu_hot.py rep kernel_regex [pct]"""
import csv
import subprocess
import sys

rep, kern = sys.argv[1], sys.argv[2]
pct = float(sys.argv[3]) if len(sys.argv) > 3 else 1.0
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kern],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"]
for k, h in enumerate(hi[:1]):
    hdr = rows[h]
    end = hi[k + 1] - 1 if k + 1 < len(hi) else len(rows)
    si, ie = hdr.index("# Samples"), hdr.index("Instructions Executed")
    body = [r for r in rows[h + 1:end] if len(r) > si and r[si].isdigit()]
    tot = sum(int(r[si]) for r in body)
    print("total samples", tot, "instructions", len(body))
    for n, r in enumerate(body):
        if int(r[si]) > tot * pct / 100:
            print("%4d %7s %10s  %s" % (n, r[si], r[ie], r[1][:100]))
