#!/usr/bin/env python
"""Summarise `ncu --page source --csv --print-source cuda,sass` output: stall reasons over the kernel, the hottest source
lines with their stall mix, shared-memory wavefronts per line. usage: ncu_source_summary.py <source.csv> [top]"""
import csv
import sys
from collections import defaultdict

path = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
rows = list(csv.reader(open(path, newline="")))
hdr_i = next(i for i, r in enumerate(rows) if r and r[0] == "Line No")
hdr = rows[hdr_i]
# the header repeats "Source": first = CUDA source, second = SASS
col = {}
for i, h in enumerate(hdr):
    col.setdefault(h, i)
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
lines, sass = [], []
cur = None
for r in rows[hdr_i + 1:]:
    if len(r) < len(hdr):
        continue
    if r[0] != "":
        cur = (r[0], r[1])
        lines.append(r)
    else:
        sass.append((cur, r))


def num(x):
    try:
        return float(x)
    except ValueError:
        return 0.0


tot = defaultdict(float)
for r in lines:
    for s in stalls:
        tot[s] += num(r[col[s]])
allsmp = sum(tot.values())
print("function:", rows[1][1] if len(rows) > 1 else "?")
print("warp-stall samples %d; by reason %%: %s" % (allsmp, {k.replace("stall_", ""): round(100 * v / allsmp, 1) for k, v in sorted(tot.items(), key=lambda kv: -kv[1]) if v > 0.01 * allsmp}))
inst = sum(num(r[col["Instructions Executed"]]) for r in lines)
wf = sum(num(r[col["L1 Wavefronts Shared"]]) for r in lines)
wfi = sum(num(r[col["L1 Wavefronts Shared Ideal"]]) for r in lines)
print("warp instructions executed %.4g; shared wavefronts %.4g (ideal %.4g)" % (inst, wf, wfi))
# opcode mix from SASS
ops = defaultdict(float)
for cur, r in sass:
    txt = r[3].strip().split()
    if not txt:
        continue
    op = txt[1] if txt[0].startswith("@") and len(txt) > 1 else txt[0]
    ops[op.split(".")[0]] += num(r[col["Instructions Executed"]])
print("opcode mix (warp instructions, %):", {k: round(100 * v / inst, 1) for k, v in sorted(ops.items(), key=lambda kv: -kv[1])[:16]})
print("--- hottest source lines (samples, %, instructions, shared wavefronts, top stalls)")
for r in sorted(lines, key=lambda r: -num(r[col["# Samples"]]))[:top]:
    smp = num(r[col["# Samples"]])
    mix = sorted(((num(r[col[s]]), s.replace("stall_", "")) for s in stalls), reverse=True)[:4]
    print("%7d %5.1f%% inst=%10d wf=%9d | %s:%s | %s" % (smp, 100 * smp / allsmp, num(r[col["Instructions Executed"]]), num(r[col["L1 Wavefronts Shared"]]),
                                                     r[0], r[1].strip()[:70], " ".join("%s=%d" % (n, v) for v, n in mix if v > 0)))
print("--- top shared-wavefront lines")
for r in sorted(lines, key=lambda r: -num(r[col["L1 Wavefronts Shared"]]))[:8]:
    w = num(r[col["L1 Wavefronts Shared"]])
    if w <= 0:
        break
    print("%10d %5.1f%% ideal=%10d | %s:%s" % (w, 100 * w / wf, num(r[col["L1 Wavefronts Shared Ideal"]]), r[0], r[1].strip()[:80]))
