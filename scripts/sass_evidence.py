"""Per-kernel SASS evidence of the product library (runs in the build container, no GPU): counts of the Blackwell
mnemonics per function of libqsp_recon.so - tcgen05.mma (UTCHMMA, .2CTA = cta_group::2), tcgen05.ld (LDTM), bulk
async copies (UBLKCP), tcgen05.commit / mbarrier (UTCBAR, SYNCS), programmatic dependent launch (PREEXIT / ACQBULK),
packed fp32 pairs (FADD2 / FFMA2 / FMUL2), mixed-precision FMA (FHFMA) - and of the legacy tensor path (HMMA: must be 0).
usage: python scripts/sass_evidence.py > profiles/r02_sass_grep.txt"""
import collections, os, re, subprocess, sys
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(root, "qsp_slam_b200", "libqsp_recon.so")
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
keys = ["UTCHMMA", "UTCHMMA.2CTA", "LDTM", "UBLKCP", "UTCBAR", "SYNCS", "PREEXIT", "ACQBULK", "FADD2", "FFMA2", "FMUL2", "FHFMA", "HMMA.", "BAR.SYNC"]
cur, cnt, total = None, collections.OrderedDict(), collections.Counter()
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
        cnt[cur] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and cur:
        op = m.group(1)
        cnt[cur]["instructions"] += 1
        for k in keys:
            if (k == "UTCHMMA" and op.startswith("UTCHMMA")) or (k != "UTCHMMA" and (op == k or op.startswith(k + ".") or (k.endswith(".") and op.startswith(k)) or (k == "UTCHMMA.2CTA" and ".2CTA" in op and op.startswith("UTCHMMA")))):
                cnt[cur][k] += 1
print("# cuobjdump -sass qsp_slam_b200/libqsp_recon.so (nvcc -gencode arch=compute_100a,code=sm_100a), counts per kernel")
print(f"{'kernel':46s} {'instr':>7s} " + " ".join(f"{k:>12s}" for k in keys))
for f, c in cnt.items():
    print(f"{f[:46]:46s} {c['instructions']:7d} " + " ".join(f"{c[k]:12d}" for k in keys))
    total.update(c)
print(f"{'total (' + str(len(cnt)) + ' kernels)':46s} {total['instructions']:7d} " + " ".join(f"{total[k]:12d}" for k in keys))
