#!/usr/bin/env python
"""Per-source-line dynamic instruction counts and stall samples of one kernel from an ncu report captured with
--import-source on (run here, no GPU):  python scripts/ncu_lines.py <file.ncu-rep> <kernel substring> [top N]"""
import csv, io, subprocess, sys, collections
rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
cur_file = cur_fn = None
hdr = None
lines = collections.OrderedDict()
for row in csv.reader(io.StringIO(out)):
    if not row:
        continue
    if row[0] == "File Path":
        cur_file = row[1].split("/")[-1]; continue
    if row[0] == "Function Name":
        cur_fn = row[1]; continue
    if row[0] == "Line No":
        hdr = row; continue
    if hdr is None or cur_fn is None or kern not in cur_fn or row[0] == "":
        continue
    d = dict(zip(hdr, row))
    try:
        inst = int(d["Instructions Executed"]); samp = int(d["# Samples"]); thr = int(d["Thread Instructions Executed"])
    except (ValueError, KeyError):
        continue
    key = (cur_file, int(row[0]))
    a = lines.setdefault(key, [0, 0, 0, row[1]])
    a[0] += inst; a[1] += samp; a[2] += thr
tot_i = sum(a[0] for a in lines.values()); tot_s = sum(a[1] for a in lines.values())
print(f"kernel ~ {kern}: {tot_i} warp instructions, {tot_s} stall samples")
# by 20-line regions
reg = collections.Counter(); regs = collections.Counter()
for (f, ln), a in lines.items():
    reg[(f, ln // 10 * 10)] += a[0]; regs[(f, ln // 10 * 10)] += a[1]
print("--- regions (10 lines) by instructions")
for (f, ln), c in reg.most_common(top):
    print(f"{f}:{ln:4d}  inst {c:12d} {c/tot_i:6.3f}   samples {regs[(f, ln)]:8d} {regs[(f, ln)]/max(1,tot_s):6.3f}")
print("--- lines by stall samples")
for (f, ln), a in sorted(lines.items(), key=lambda x: -x[1][1])[:top]:
    print(f"{f}:{ln:4d} samples {a[1]:7d} {a[1]/max(1,tot_s):6.3f} inst {a[0]:10d} {a[0]/tot_i:6.3f} lanes {a[2]/max(1,a[0]):4.1f} | {a[3][:90]}")
