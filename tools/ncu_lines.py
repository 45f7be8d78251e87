#!/usr/bin/env python3
"""Per-source-line hot spots from an .ncu-rep (needs -lineinfo + --import-source on).
   python tools/ncu_lines.py rep.ncu-rep [top]"""
import csv, subprocess, sys
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
cur = None; hdr = None; lines = []
for r in rows:
    if len(r) >= 2 and r[0] == "File Path": cur = r[1].split("/")[-1]; continue
    if len(r) >= 2 and r[0] == "Line No": hdr = r; continue
    if hdr and len(r) == len(hdr) and r[0] not in ("", "Line No"):
        d = dict(zip(range(len(hdr)), r))
        try:
            ln = int(r[0])
        except ValueError:
            continue
        def col(name):
            i = hdr.index(name); v = r[i]
            try: return float(v)
            except ValueError: return 0.0
        lines.append((col("Instructions Executed"), col("Thread Instructions Executed"), col("# Samples"), cur, ln, r[1].strip()[:110]))
tot_i = sum(l[0] for l in lines); tot_s = sum(l[2] for l in lines)
print("total warp instructions %.3g, samples %d" % (tot_i, tot_s))
print("by instructions:")
for l in sorted(lines, key=lambda x: -x[0])[:top]:
    print("%5.1f%% inst %5.1f%% smp  lanes %4.1f  %s:%d  %s" % (100*l[0]/max(tot_i,1), 100*l[2]/max(tot_s,1), l[1]/max(l[0],1), l[3], l[4], l[5]))
print("by stall samples:")
for l in sorted(lines, key=lambda x: -x[2])[:top]:
    print("%5.1f%% smp %5.1f%% inst  lanes %4.1f  %s:%d  %s" % (100*l[2]/max(tot_s,1), 100*l[0]/max(tot_i,1), l[1]/max(l[0],1), l[3], l[4], l[5]))
# ---- aggregate by file / line range (kernels_continuous.cu regions) ----
import collections
agg = collections.Counter(); aggt = collections.Counter()
def region(f, ln):
    if f == "kernels_continuous.cu":
        for name, a, b in REGIONS:
            if a <= ln <= b: return name
        return "kc:other"
    return f
REGIONS = []
try:
    src = open("/root/repo/hpp-core_b200/csrc/kernels_continuous.cu").read().split("\n")
    marks = [(i + 1, l.split("@region")[1].strip()) for i, l in enumerate(src) if "@region" in l]
    for j, (ln, name) in enumerate(marks):
        REGIONS.append((name, ln, marks[j + 1][0] - 1 if j + 1 < len(marks) else 10**9))
except Exception:
    pass
for l in lines:
    r = region(l[3], l[4]); agg[r] += l[0]; aggt[r] += l[1]
print("by region:")
for r, v in agg.most_common(20):
    print("%5.1f%% inst  lanes %4.1f  %s" % (100 * v / max(tot_i, 1), aggt[r] / max(v, 1), r))
