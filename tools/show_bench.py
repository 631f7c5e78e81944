"""Prints the kernel table and headline numbers of bench JSON lines under gpurun_out/ (developer helper)."""
import glob, json, sys
pat = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/bench_*.json"
for f in sorted(glob.glob(pat)):
    try:
        d = json.load(open(f))
    except Exception as e:
        print(f, "ERR", e); continue
    print(f"== {f}: {d.get('value')} {d.get('unit')}  ms/view {d.get('fwd_bwd_ms_per_iter', d.get('ms_per_view'))}  e2e {(d.get('e2e') or {}).get('value')}  launches {d.get('gpu_launches')}")
    for k, v in (d.get("kernels") or {}).items():
        print(f"   {k:22s} {v.get('ms_per_view'):.4f}  x{v.get('launches_per_view')}  frac {v.get('frac_of_roofline')}")
