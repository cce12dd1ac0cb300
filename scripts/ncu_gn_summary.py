"""Summarises an `ncu --set full` capture of the Gauss-Newton kernels (b)-(d) of one warm kitti_batch32 iteration
(gpurun_out/gn_kernels.ncu-rep, written by an `ncu --set full -k regex:k_` pass over scripts/timeline.py) into profiles/<tag>_ncu_gn_kernels_summary.txt:
per kernel the duration, DRAM bytes read + written, achieved DRAM GB/s (bytes / duration), the share of the measured
HBM copy bandwidth (MEASURED_PEAKS.json, fallback 6500 GB/s), L2 and SM throughput and the launch shape.
usage: python scripts/ncu_gn_summary.py r01i"""
import collections, csv, json, os, subprocess, sys
tag = sys.argv[1]
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rep = os.path.join(root, "gpurun_out", "gn_kernels.ncu-rep")
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
peak = 6500.0
try:
    mp = json.load(open(os.path.join(root, "MEASURED_PEAKS.json")))
    for k, v in mp.items():
        if "hbm" in k.lower() and isinstance(v, (int, float)):
            peak = float(v); break
        if isinstance(v, dict):
            for k2, v2 in v.items():
                if "hbm" in (k + k2).lower() and "burst" in (k + k2).lower() and isinstance(v2, (int, float)):
                    peak = float(v2)
except Exception:
    pass


def val(r, k):
    i = hdr.index(k)
    x = float(r[i].replace(",", ""))
    u = units[i]
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}
    return x * scale.get(u, 1)


agg = collections.OrderedDict()
for r in rows[2:]:
    if len(r) < len(hdr):
        continue
    name = r[hdr.index("Kernel Name")].split("(")[0].replace("void ", "")
    a = agg.setdefault(name, collections.defaultdict(float))
    a["n"] += 1
    a["us"] += val(r, "gpu__time_duration.sum")
    a["rd"] += val(r, "dram__bytes_read.sum")
    a["wr"] += val(r, "dram__bytes_write.sum")
    for k, short in (("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram_pct"),
                     ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "l2_pct"),
                     ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm_pct"),
                     ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ_pct")):
        if k in hdr:
            a[short] += val(r, k)
    a["grid"] = r[hdr.index("launch__grid_size")] if "launch__grid_size" in hdr else "?"
    a["block"] = r[hdr.index("launch__block_size")] if "launch__block_size" in hdr else "?"
    a["regs"] = r[hdr.index("launch__registers_per_thread")] if "launch__registers_per_thread" in hdr else "?"
out = os.path.join(root, "profiles", f"{tag}_ncu_gn_kernels_summary.txt")
with open(out, "w") as f:
    f.write("# ncu --set full --clock-control none, the Gauss-Newton kernels (b)-(d) of one warm kitti_batch32 iteration\n"
            "# (QSP_TIMELINE=0 python scripts/timeline.py, launches 120.. of the run: a warm step).  Per launch averages;\n"
            f"# GB/s = (dram__bytes_read.sum + dram__bytes_write.sum) / gpu__time_duration.sum; HBM peak {peak:.0f} GB/s (measured copy).\n"
            "# Under ncu every launch is cold-cache and serialised: the byte counts are the evidence, the times are upper bounds.\n")
    f.write(f"{'kernel':28s} {'n':>3s} {'us':>8s} {'rd KB':>9s} {'wr KB':>9s} {'GB/s':>8s} {'%HBM':>6s} {'dram%':>6s} {'L2%':>6s} "
            f"{'SM%':>6s} {'occ%':>6s} {'grid':>7s} {'block':>6s} {'regs':>5s}\n")
    for k, a in agg.items():
        n = a["n"]
        gbs = (a["rd"] + a["wr"]) / (a["us"] * 1e-6) / 1e9 if a["us"] else 0.0
        f.write(f"{k[:28]:28s} {int(n):3d} {a['us'] / n:8.1f} {a['rd'] / n / 1e3:9.1f} {a['wr'] / n / 1e3:9.1f} {gbs:8.1f} "
                f"{gbs / peak * 100:6.2f} {a['dram_pct'] / n:6.1f} {a['l2_pct'] / n:6.1f} {a['sm_pct'] / n:6.1f} "
                f"{a['occ_pct'] / n:6.1f} {a['grid']:>7s} {a['block']:>6s} {a['regs']:>5s}\n")
print(open(out).read())
