"""Copies the outputs of scripts/gpu_round2.sh from gpurun_out/ (scratch) into profiles/ under a round tag:
bench line, ncu launch list + per-kernel shares, `ncu --set full` summary of the decoder kernels, config timings,
in-situ timelines; refreshes profiles/ncu_traffic.json (DRAM bytes per launch, read by bench.py).
usage: python scripts/save_profiles.py r02a"""
import collections, csv, json, os, shutil, subprocess, sys
tag = sys.argv[1]
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
go, pr = os.path.join(root, "gpurun_out"), os.path.join(root, "profiles")
for src, dst in (("launches.csv", "launches_bench.csv"), ("bench.json", "bench.json"), ("configs.jsonl", "configs.jsonl"),
                 ("timeline.txt", "timeline_kitti_batch32.txt"), ("timeline_single.txt", "timeline_kitti_single.txt"),
                 ("bench_reference.json", "bench_reference.json"), ("single_ab.jsonl", "single_ab.jsonl"),
                 ("parity_decompose.md", "parity_decompose.md"), ("profile_decoder.log", "profile_decoder.txt"),
                 ("fullsize_parity.jsonl", "fullsize_parity.jsonl"), ("fullsize_parity_cases.jsonl", "fullsize_parity_cases.jsonl"),
                 ("bench_gpus2.json", "bench_gpus2.json"), ("bench_gpus4.json", "bench_gpus4.json"),
                 ("bench_gpus8.json", "bench_gpus8.json"), ("bench_gpus1_samebox.json", "bench_gpus1_samebox.json"),
                 ("pytest_gpu_tail.txt", "pytest_gpu_tail.txt"), ("smi.txt", "smi.txt"),
                 ("launches_kitti_single.csv", "launches_kitti_single.csv"), ("random_objects_errors.jsonl", "random_objects_errors.jsonl")):
    if os.path.exists(os.path.join(go, src)):
        shutil.copy(os.path.join(go, src), os.path.join(pr, f"{tag}_{dst}"))
bench = json.load(open(os.path.join(go, "bench.json")))
rows_it = bench.get("rows_per_iteration", {})
if not os.path.exists(os.path.join(go, "decoder_tc.ncu-rep")) or not os.path.exists(os.path.join(go, "launches.csv")):
    print("no ncu outputs in gpurun_out/: copied the plain files only")
    sys.exit(0)
raw = subprocess.run(["ncu", "-i", os.path.join(go, "decoder_tc.ncu-rep"), "--page", "raw", "--csv"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
keys = ["gpu__time_duration.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "dram__bytes_read.sum",
        "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__cycles_elapsed.avg", "sm__cycles_elapsed.avg.per_second",
        "sm__cycles_active.avg", "smsp__inst_executed.sum", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__cluster_size",
        "launch__shared_mem_per_block_dynamic"]
tr = {}
with open(os.path.join(pr, f"{tag}_ncu_decoder_summary.txt"), "w") as f:
    f.write(f"# ncu --set full --clock-control none --import-source on -k regex:k_decoder_ -c 4, scripts/profile_decoder.py "
            f"{rows_it.get('first_pass')} {rows_it.get('fwd_jac')}\n# k_decoder_tc<0,0,2> = TC_FAST forward (CTA pairs, M=256), "
            "k_decoder_p3<1> = three-product fp16-pair forward + reverse sweep\n# (CTA pairs, M=128).  Row counts of one "
            "kitti_batch32 GN iteration (first pass: all phases in one launch).  Launches 3-4 of 4 listed.\n")
    for r in rows[4:6]:
        name = r[hdr.index("Kernel Name")].split("(")[0]
        f.write(name + "\n")
        for k in keys:
            if k in hdr:
                i = hdr.index(k); f.write(f"   {k} = {r[i]} {units[i]}\n")
        def b(k):
            i = hdr.index(k); return float(r[i]) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[units[i]]
        tr[name] = int(b("dram__bytes_read.sum") + b("dram__bytes_write.sum"))
fast = [v for k, v in tr.items() if "k_decoder_tc<0" in k][0]
jac = [v for k, v in tr.items() if "k_decoder_p3<1>" in k][0]
json.dump({"source": f"profiles/{tag}_ncu_decoder_summary.txt (ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum per launch, "
                     "rows of one kitti_batch32 GN iteration)",
           "decoder_fwd": {"rows": rows_it.get("first_pass"), "dram_bytes": fast},
           "decoder_fwd_jac": {"rows": rows_it.get("fwd_jac"), "dram_bytes": jac}},
          open(os.path.join(pr, "ncu_traffic.json"), "w"), indent=1)
L = list(csv.reader(open(os.path.join(pr, f"{tag}_launches_bench.csv"))))
h = next(i for i, r in enumerate(L) if "Kernel Name" in r)
hd = L[h]; ki, vi, ui = hd.index("Kernel Name"), hd.index("Metric Value"), hd.index("Metric Unit")
agg = collections.OrderedDict()
for r in L[h + 1:]:
    if len(r) <= vi: continue
    name = r[ki].split("(")[0].replace("void ", ""); v = float(r[vi].replace(",", "")); u = r[ui]
    v = v / 1e3 if u == "ns" else v * 1e3 if u == "ms" else v
    a = agg.setdefault(name, [0, 0.0]); a[0] += 1; a[1] += v
tot = sum(a[1] for a in agg.values())
with open(os.path.join(pr, f"{tag}_step_breakdown.txt"), "w") as f:
    f.write("# per-kernel share of one bench command (python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-configs) from the ncu launch list\n"
            f"# profiles/{tag}_launches_bench.csv (gpu__time_duration.sum; cold-cache, serialised launches: shares, not absolutes)\n")
    f.write(f"{'kernel':44s} {'launches':>8s} {'total ms':>10s} {'avg us':>9s} {'share':>7s}\n")
    for k, a in sorted(agg.items(), key=lambda x: -x[1][1]):
        f.write(f"{k[:44]:44s} {a[0]:8d} {a[1] / 1e3:10.3f} {a[1] / a[0]:9.1f} {a[1] / tot * 100:6.1f}%\n")
    f.write(f"{'total':44s} {sum(a[0] for a in agg.values()):8d} {tot / 1e3:10.3f}\n")
print(open(os.path.join(pr, f"{tag}_step_breakdown.txt")).read())
print(open(os.path.join(pr, f"{tag}_ncu_decoder_summary.txt")).read()[:1500])
