"""Summarise an `ncu --set full` report of the headline step into profiles/:
    ncu -i gpurun_out/prof_vN.ncu-rep --page raw --csv > /tmp/raw.csv; python tools/ncu_summary.py /tmp/raw.csv vN "<commit note>"
writes profiles/r1_ncu_full_vN_summary.txt (per-launch counters) and profiles/ncu_traffic.json (DRAM bytes per frame and
stage, which bench.py reports as roofline.traffic)."""
import csv
import json
import sys

raw, ver, note = sys.argv[1], sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else ""
rows = list(csv.reader(open(raw)))
hdr, units = rows[0], rows[1]
keys = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__occupancy_limit_shared_mem",
        "launch__occupancy_limit_registers", "smsp__inst_executed.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "lts__t_sector_hit_rate.pct"] + [f"smsp__average_warps_issue_stalled_{s}_per_issue_active.ratio" for s in
                                         ("long_scoreboard", "short_scoreboard", "barrier", "mio_throttle", "math_pipe_throttle", "wait", "not_selected")]
idx = {h: i for i, h in enumerate(hdr)}
out = ["# ncu --set full --clock-control none --import-source on, bench.py --steps 1 --warmup 3 --batch 64 --skip-hamming --no-cpu-baseline "
       "(64 frames 640x480 per step; the device entry splits a batch into two 32-frame sub-batches on two streams)",
       f"# source report: gpurun_out/prof_{ver}.ncu-rep (round 1, {note}); per-launch values; cold-cache, serialised: compare shares, not absolutes"]
scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
agg, seen = {}, set()
for r in rows[2:]:
    name = r[idx["Kernel Name"]]
    out.append(f"kernel: {name} grid {r[idx['launch__grid_size']]} block {r[idx['launch__block_size']]}")
    for k in keys:
        if k in idx:
            out.append(f"  {k:<78}{r[idx[k]]} {units[idx[k]]}")
    stage = ("pyramid" if "k_pyr" in name else "fast" if "k_fast" in name else
             "quadtree" if ("k_quadtree" in name or "k_cell_scan" in name or "k_qt_" in name) else
             "describe" if "k_describe" in name else "match" if "k_top2" in name else None)
    b = sum(float(r[idx[m]]) * scale[units[idx[m]]] for m in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
    key = (name[:30], r[idx["launch__grid_size"]])
    if key in seen:
        continue
    seen.add(key)
    agg.setdefault(stage, [0, []])
    agg[stage][0] += b
    agg[stage][1].append((name[:18], r[idx["launch__grid_size"]], round(b / 1e6, 2), r[idx["gpu__time_duration.sum"]]))
open(f"profiles/r1_ncu_full_{ver}_summary.txt", "w").write("\n".join(out) + "\n")
for k, v in agg.items():
    print(k, round(v[0] / 1e6, 2), v[1])
d = {"source": f"profiles/r1_ncu_full_{ver}_summary.txt (ncu --set full, 32-frame launches; dram__bytes_read.sum + dram__bytes_write.sum; "
               "writes of a 32-frame launch largely stay in the 126 MB L2 during the kernel)",
     "frames_per_launch": 32, "bytes_per_frame": {k: v[0] / 32 for k, v in agg.items() if k in ("pyramid", "fast", "quadtree", "describe")}}
json.dump(d, open("profiles/ncu_traffic.json", "w"), indent=1)
