"""Turn the ncu captures in gpurun_out/ into the tracked summaries under profiles/ (run here, no GPU needed).

  launches csv  -> profiles/<tag>_ncu_launches_bench.csv (copy) + per-kernel share table (json)
  full .ncu-rep -> profiles/<tag>_ncu_full_summary.json (per kernel: duration, DRAM bytes, issue/pipe utilisation,
                   occupancy, top stall reasons) and profiles/ncu_traffic.json (DRAM bytes per launch, read by bench.py)
usage: summarize_profiles.py <tag> <launches.csv> <full.ncu-rep>
"""
import csv, json, os, re, shutil, subprocess, sys
from collections import defaultdict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag, launches, rep = sys.argv[1:4]
out_dir = os.path.join(ROOT, "profiles")

def short(name):
    m = re.search(r"(\w+)_kernel", name)
    if m:
        k = m.group(1)
        return {"blend_backward_streams": "blend_backward", "blend_forward_streams": "blend_forward",
                "scan_apply_sums": "scan_offsets"}.get(k, k)
    return re.sub(r"\(.*", "", name).split("::")[-1][:40]

# ---- launch list: share of the captured window per kernel
rows = [r for r in csv.reader(open(launches)) if len(r) > 14 and r[0].isdigit()]
# the work-counter kernels bench.py runs once after the timed region (roofline accounting) are not part of a step
rows = [r for r in rows if "stats_kernel" not in r[4]]
tot = defaultdict(lambda: [0, 0.0])
for r in rows:
    k = short(r[4]); tot[k][0] += 1; tot[k][1] += float(r[14]) / 1e6
total_ms = sum(v[1] for v in tot.values())
share = {k: {"launches": v[0], "ms": round(v[1], 4), "share": round(v[1] / total_ms, 4)} for k, v in
         sorted(tot.items(), key=lambda kv: -kv[1][1])}
shutil.copy(launches, os.path.join(out_dir, f"{tag}_ncu_launches_bench.csv"))
json.dump({"command": "ncu --metrics gpu__time_duration.sum --clock-control none -c 800 python bench.py --steps 2 --warmup 3 --no-extras "
                      "--no-cpu-baseline --no-e2e --no-kernel-timing (tools/gpu_ncu_list.sh)", "captured_launches": len(rows), "total_ms": round(total_ms, 4),
           "per_kernel": share}, open(os.path.join(out_dir, f"{tag}_ncu_launch_shares.json"), "w"), indent=1)

# ---- full capture
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
r = list(csv.reader(raw.splitlines()))
h, units, data = r[0], r[1], r[2:]
col = {k: i for i, k in enumerate(h)}
want = {"duration_us": "gpu__time_duration.sum", "dram_read_bytes": "dram__bytes_read.sum", "dram_write_bytes": "dram__bytes_write.sum",
        "dram_pct_of_peak": "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "issue_active_pct": "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "warps_active_pct": "sm__warps_active.avg.pct_of_peak_sustained_active",
        "pipe_fma_pct": "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "pipe_alu_pct": "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
        "warp_instructions": "smsp__inst_executed.sum", "registers": "launch__registers_per_thread",
        "global_red_instr": "smsp__inst_executed_op_global_red.sum"}
scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1, "ms": 1e3, "us": 1, "ns": 1e-3, "s": 1e6}
summary, traffic = {}, defaultdict(list)
for row in data:
    k = short(row[col["Kernel Name"]])
    if "Kernel Name" not in col:
        break
    tmpl = re.search(r"<\(?(?:bool\)|int\))?(\w+)>", row[col["Kernel Name"]])
    e = {}
    for name, metric in want.items():
        if metric in col and row[col[metric]] not in ("", "n/a"):
            v = float(row[col[metric]].replace(",", ""))
            u = units[col[metric]]
            e[name] = round(v * scale.get(u, 1), 3) if name.endswith("bytes") or name == "duration_us" else round(v, 3)
    stalls = {}
    for i, c in enumerate(h):
        if c.startswith("smsp__average_warps_issue_stalled_") and c.endswith("_per_issue_active.ratio") and "not_issued" not in c:
            try:
                stalls[c[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]] = round(float(row[i]), 2)
            except ValueError:
                pass
    e["stalls_per_issue"] = dict(sorted(stalls.items(), key=lambda kv: -kv[1])[:5])
    e["grid"] = row[col["Grid Size"]] if "Grid Size" in col else None
    if k == "onesweep":                      # same template for both sorts: the depth sort has as many CTAs as P needs
        seen_onesweep = len(summary.get("onesweep_depth", [])) + len(summary.get("onesweep_tile", []))
        k = "onesweep_depth" if seen_onesweep < 4 else "onesweep_tile"
    summary.setdefault(k, []).append(e)
    traffic[k].append(e.get("dram_read_bytes", 0) + e.get("dram_write_bytes", 0))
json.dump({"command": "ncu --set full --clock-control none --import-source on -k regex:... --launch-skip 375 -c 15 python bench.py --steps 1 "
                      "--warmup 3 ... (tools/gpu_ncu_full.sh: the 15 launches of one 3M-Gaussian 1080p view of the timed step)",
           "kernels": summary}, open(os.path.join(out_dir, f"{tag}_ncu_full_summary.json"), "w"), indent=1)
# bench.py names (per-launch DRAM bytes, mean over the captured launches of that kernel)
name_map = {k: k for k in ("blend_forward", "blend_backward", "preprocess_fwd", "preprocess_bwd", "emit_instances",
                           "onesweep_depth", "onesweep_tile", "radix_hist", "tile_table", "scan_offsets", "scan_tile_sums")}
tr = {name_map[k]: round(sum(v) / len(v)) for k, v in traffic.items() if k in name_map}
json.dump(tr, open(os.path.join(out_dir, "ncu_traffic.json"), "w"), indent=1)
print(json.dumps(share, indent=1)); print(json.dumps(tr))
