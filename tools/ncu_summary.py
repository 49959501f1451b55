"""Summarise an `ncu --set full` report of the headline step into profiles/:
    ncu -i gpurun_out/<rep>.ncu-rep --page raw --csv > /tmp/raw.csv
    python tools/ncu_summary.py /tmp/raw.csv <tag> <frames_per_launch> "<capture command / commit note>"
writes profiles/<tag>_ncu_full_summary.txt (per-launch counters of every distinct kernel/grid) and
profiles/ncu_traffic.json (per frame and stage: DRAM bytes, warp instructions, ALU-pipe warp instructions), which bench.py
reads for roofline.traffic and the instruction-issue roofline of FAST."""
import csv
import json
import sys

raw, tag, fpl = sys.argv[1], sys.argv[2], int(sys.argv[3])
note = sys.argv[4] if len(sys.argv) > 4 else ""
# frames one launch of each stage covers: the extractor's kernels run per sub-batch (fpl frames), the frame-pair match once
# per step over both sub-batches
match_fpl = int(sys.argv[5]) if len(sys.argv) > 5 else 2 * fpl
frames_of = lambda stage: match_fpl if stage == "match" else fpl  # noqa: E731
rows = list(csv.reader(open(raw)))
hdr, units = rows[0], rows[1]
keys = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__occupancy_limit_shared_mem",
        "launch__occupancy_limit_registers", "smsp__inst_executed.sum", "sm__inst_executed_pipe_alu.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "lts__t_sector_hit_rate.pct", "sm__cycles_active.avg"] + [
    f"smsp__average_warps_issue_stalled_{s}_per_issue_active.ratio" for s in
    ("long_scoreboard", "short_scoreboard", "barrier", "mio_throttle", "math_pipe_throttle", "wait", "not_selected")]
idx = {h: i for i, h in enumerate(hdr)}
out = [f"# ncu --set full --clock-control none --import-source on; {note}",
       f"# {fpl} frames per launch; per-launch values of every distinct (kernel, grid); cold-cache, serialised: compare shares, not absolutes"]
scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def num(r, m):
    return float(r[idx[m]].replace(",", "")) if m in idx and r[idx[m]] not in ("", "n/a") else 0.0


agg, seen = {}, set()
for r in rows[2:]:
    name = r[idx["Kernel Name"]]
    key = (name[:40], r[idx["launch__grid_size"]])
    if key in seen:
        continue
    seen.add(key)
    out.append(f"kernel: {name} grid {r[idx['launch__grid_size']]} block {r[idx['launch__block_size']]}")
    for k in keys:
        if k in idx:
            out.append(f"  {k:<78}{r[idx[k]]} {units[idx[k]]}")
    stage = ("pyramid" if "k_pyr" in name else "fast" if "k_fast" in name else
             "quadtree" if ("k_quadtree" in name or "k_cell_scan" in name or "k_qt_" in name) else
             "describe" if "k_describe" in name else "match" if "k_top2" in name else "other")
    b = sum(num(r, m) * scale.get(units[idx[m]], 1) for m in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
    a = agg.setdefault(stage, {"bytes": 0.0, "inst": 0.0, "alu": 0.0, "us": 0.0, "kernels": []})
    a["bytes"] += b
    a["inst"] += num(r, "smsp__inst_executed.sum")
    us0 = num(r, "gpu__time_duration.sum")
    a["alu"] += num(r, "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active") * us0
    a["issue"] = a.get("issue", 0.0) + num(r, "smsp__issue_active.avg.pct_of_peak_sustained_active") * us0
    a["w"] = a.get("w", 0.0) + us0
    us = num(r, "gpu__time_duration.sum") * {"us": 1, "ms": 1e3, "ns": 1e-3, "s": 1e6}.get(units[idx["gpu__time_duration.sum"]], 1)
    a["us"] += us
    a["kernels"].append((name.split("(")[0][-28:], r[idx["launch__grid_size"]], round(b / 1e6, 2), round(us, 1)))
open(f"profiles/{tag}_ncu_full_summary.txt", "w").write("\n".join(out) + "\n")
for k, v in agg.items():
    print(k, f"{v['bytes'] / 1e6:.1f} MB, {v['inst'] / 1e6:.1f} M warp-inst, {v['us']:.0f} us", v["kernels"])
stages = ("pyramid", "fast", "quadtree", "describe", "match")
d = {"source": f"profiles/{tag}_ncu_full_summary.txt (ncu --set full, {fpl}-frame launches as in the timed bench; "
               "dram__bytes_read.sum + dram__bytes_write.sum, smsp__inst_executed.sum; pipe / issue percentages are time-weighted over the stage's kernels)",
     "frames_per_launch": fpl, "match_frames_per_launch": match_fpl,
     "bytes_per_frame": {k: agg[k]["bytes"] / frames_of(k) for k in stages if k in agg},
     "warp_inst_per_frame": {k: agg[k]["inst"] / frames_of(k) for k in stages if k in agg},
     "alu_pipe_pct_ncu": {k: agg[k]["alu"] / agg[k]["w"] for k in stages if k in agg},
     "issue_active_pct_ncu": {k: agg[k]["issue"] / agg[k]["w"] for k in stages if k in agg},
     "ncu_us_per_launch": {k: agg[k]["us"] for k in stages if k in agg}}
json.dump(d, open("profiles/ncu_traffic.json", "w"), indent=1)
