"""Per-source-line instruction counts and stall samples from `ncu --page source --print-source cuda,sass --csv`.
usage: python tools/ncu_lines.py <csv> [top_n]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
hdr = None
agg = {}
fpath = ""
for r in rows:
    if len(r) >= 2 and r[0] == "File Path":
        fpath = r[1].split("/")[-1]
    if r and r[0] == "Line No":
        hdr = r
        continue
    if hdr is None or len(r) < len(hdr) or not r[0].isdigit() or r[2] != "-":
        continue
    d = dict(zip(hdr, r))
    key = (fpath, int(r[0]), r[1].strip()[:90])
    inst = int(d.get("Instructions Executed", "0") or 0)
    samp = int(d.get("# Samples", "0") or 0)
    a = agg.setdefault(key, [0, 0])
    a[0] += inst
    a[1] += samp
tot_i = sum(a[0] for a in agg.values()) or 1
tot_s = sum(a[1] for a in agg.values()) or 1
print(f"total warp instructions {tot_i}, samples {tot_s}")
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
    print(f"{k[0]}:{k[1]:4d} inst {100 * a[0] / tot_i:5.1f}%  samples {100 * a[1] / tot_s:5.1f}%  {k[2]}")
