"""Pretty-print a tools/ncu_kernels.sh CSV: one block per launch."""
import csv
import sys
from collections import OrderedDict

rows = list(csv.DictReader(l for l in open(sys.argv[1]) if l.startswith('"')))
d = OrderedDict()
for r in rows:
    d.setdefault((r["ID"], r["Kernel Name"][:28], r["Grid Size"], r["Block Size"]), OrderedDict())[r["Metric Name"]] = r["Metric Value"]
short = lambda m: m.replace("smsp__average_warps_issue_stalled_", "stall_").replace("_per_issue_active.ratio", "").replace(
    ".avg.pct_of_peak_sustained_active", "%").replace(".sum", "")
for k, v in d.items():
    print(*k)
    print("   " + "  ".join(f"{short(m)}={x}" for m, x in v.items()))
