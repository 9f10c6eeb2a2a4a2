#!/usr/bin/env python
"""Groups the SASS of the first kernel in an ncu report (--import-source on) into regions of equal execution count and
prints instructions, stall samples and shared wavefronts per region; with --top also the most-sampled instructions.
usage: python tools/ncu_regions.py report.ncu-rep [--kernel substr] [--top N] [--dump a:b]"""
import csv, io, subprocess, sys
from collections import Counter

rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
# several kernels in one report: blocks start with a ["Kernel Name", name] row followed by a header row; keep the first
# block, or the last one whose name contains the --kernel substring
starts = [k for k, r in enumerate(rows) if r and r[0] == "Kernel Name"]
pick = starts[0]
if "--kernel" in sys.argv:
    want = sys.argv[sys.argv.index("--kernel") + 1]
    pick = [k for k in starts if want in rows[k][1]][-1]
print(rows[pick][1][:120])
hdr = rows[pick + 1]
iS, iI, iSm, iW, iT = (hdr.index(x) for x in ("Source", "Instructions Executed", "# Samples", "L1 Wavefronts Shared", "Avg. Threads Executed"))
nxt = [k for k in starts if k > pick]
body = rows[pick + 2:(nxt[0] if nxt else len(rows))]
lines = [(int(r[iI]), int(r[iSm]), int(r[iW]), r[iS].strip(), r[iT]) for r in body if len(r) > iW]
tot = sum(l[0] for l in lines); ts = sum(l[1] for l in lines)
print(f"total warp instructions {tot/1e9:.3f} G, samples {ts}, shared wavefronts {sum(l[2] for l in lines)/1e9:.3f} G")
regs, cur = [], None
for i, l in enumerate(lines):
    if cur and abs(l[0] - cur["n"]) <= 0.02 * max(cur["n"], 1):
        cur["end"] = i; cur["instr"] += l[0]; cur["smp"] += l[1]; cur["wf"] += l[2]; cur["cnt"] += 1
    else:
        if cur: regs.append(cur)
        cur = dict(start=i, end=i, n=l[0], instr=l[0], smp=l[1], wf=l[2], cnt=1)
regs.append(cur)
for r in regs:
    if r["instr"] > 0.003 * tot or r["smp"] > 0.01 * ts:
        ops = Counter((lines[k][3].split()[1] if lines[k][3].startswith("@") else lines[k][3].split()[0]) for k in range(r["start"], r["end"] + 1))
        print(f"[{r['start']:4d}-{r['end']:4d}] x{r['n']/1e6:8.1f}M n_instr {r['cnt']:3d} total {r['instr']/1e9:6.3f}G ({100*r['instr']/tot:4.1f}%) smp {100*r['smp']/ts:4.1f}% wf {r['wf']/1e6:7.1f}M {dict(ops.most_common(7))}")
if "--top" in sys.argv:
    n = int(sys.argv[sys.argv.index("--top") + 1])
    for i in sorted(range(len(lines)), key=lambda i: -lines[i][1])[:n]:
        print(f"{i:5d} {100*lines[i][1]/ts:5.1f}% {lines[i][0]/1e6:8.1f}M thr {lines[i][4]:>4s} {lines[i][3]}")
if "--dump" in sys.argv:
    a, b = (int(x) for x in sys.argv[sys.argv.index("--dump") + 1].split(":"))
    for i in range(a, b):
        print(f"{i:5d} {100*lines[i][1]/ts:5.2f}% {lines[i][0]/1e6:8.1f}M thr {lines[i][4]:>4s} wf {lines[i][2]/1e6:7.1f}M {lines[i][3]}")
