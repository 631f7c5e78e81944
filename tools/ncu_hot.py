"""Read an `ncu --page source --csv` dump (one kernel) and print the SASS instructions that carry the most executed
instructions / stall samples.   usage: ncu_hot.py dump.csv [min_pct]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
thr = float(sys.argv[2]) if len(sys.argv) > 2 else 0.4
h = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr, data = rows[h], [r for r in rows[h + 1:] if len(r) >= len(rows[h]) - 2 and r[0].startswith("0x")]
ia, ie, isamp, it = (hdr.index(k) for k in ("Source", "Instructions Executed", "# Samples", "Avg. Predicated-On Threads Executed"))
tot = sum(int(r[ie]) for r in data); ts = max(1, sum(int(r[isamp]) for r in data))
print("kernel:", rows[0][1][:100] if len(rows[0]) > 1 else "", "| warp-instr", tot, "samples", ts, "sass lines", len(data))
for n, r in enumerate(data):
    e = int(r[ie]); s = int(r[isamp])
    if e > tot * thr / 100 or s > ts * thr / 100:
        print(f"{n:4d} {e / tot * 100:5.2f}%i {s / ts * 100:5.2f}%s thr {float(r[it]):4.1f}  {r[ia].strip()[:100]}")
