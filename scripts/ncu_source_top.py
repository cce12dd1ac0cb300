"""Summarise `ncu --page source --csv` output: per kernel, the SASS instructions with the most stall
samples and the most excessive shared-memory wavefronts.  usage: ncu_source_top.py file.csv [topN]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
secs = []; cur = None
for r in rows:
    if r and r[0] == "Kernel Name": cur = {"name": r[1], "rows": []}; secs.append(cur)
    elif r and r[0] == "Address": cur["hdr"] = r
    elif cur is not None and len(r) > 5: cur["rows"].append(r)
for s in secs:
    h = s["hdr"]; ix = {k: i for i, k in enumerate(h)}
    R = s["rows"]
    tot = sum(int(r[ix["# Samples"]]) for r in R)
    print("=" * 100); print(s["name"][:80], "instructions", len(R), "samples", tot)
    stall_cols = [k for k in h if k.startswith("stall_") and "Not Issued" not in k]
    agg = {k: sum(int(r[ix[k]]) for r in R) for k in stall_cols}
    print("stall totals:", {k: v for k, v in sorted(agg.items(), key=lambda x: -x[1]) if v > tot * 0.01})
    print("-- top by samples")
    for n, r in sorted(enumerate(R), key=lambda x: -int(x[1][ix["# Samples"]]))[:top]:
        st = {k[6:]: int(r[ix[k]]) for k in stall_cols if int(r[ix[k]]) > 0}
        st = dict(sorted(st.items(), key=lambda x: -x[1])[:3])
        print(f"{n:5d} {r[ix['Source']].strip()[:60]:60s} smp {r[ix['# Samples']]:>6s} exec {r[ix['Instructions Executed']]:>9s} {st}")
    print("-- top by excessive shared wavefronts")
    for n, r in sorted(enumerate(R), key=lambda x: -int(x[1][ix["L1 Wavefronts Shared Excessive"]]))[:8]:
        print(f"{n:5d} {r[ix['Source']].strip()[:60]:60s} excess {r[ix['L1 Wavefronts Shared Excessive']]:>9s} wavefronts {r[ix['L1 Wavefronts Shared']]:>9s} ideal {r[ix['L1 Wavefronts Shared Ideal']]:>9s}")
