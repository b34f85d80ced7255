#!/usr/bin/env python
"""Per-source-line roll-up of an .ncu-rep (needs -lineinfo and --import-source on): samples, instructions, shared wavefronts."""
import csv, io, subprocess, sys
rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass,cuda"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hi = next(i for i, x in enumerate(rows) if x and x[0] == "Line No")
h = rows[hi]
ci = {n: i for i, n in enumerate(h)}
S, I, W = h.index("# Samples"), h.index("Instructions Executed"), h.index("L1 Wavefronts Shared")
X = h.index("L1 Wavefronts Shared Excessive") if "L1 Wavefronts Shared Excessive" in h else None  # bank-conflict replays
fname = ""
agg = []
for x in rows[:hi]:
    if x and x[0] == "File Path": fname = x[1].split("/")[-1]
tot_s = tot_i = 0
for x in rows[hi + 1:]:
    if len(x) < len(h): 
        if x and x[0] == "File Path": fname = x[1].split("/")[-1]
        continue
    if x[0] and x[0] != "Line No":  # a source line (roll-up of its SASS)
        f = lambda v: float(v) if v not in ('', '-') else 0.0
        s, i, w = f(x[S - len(h)]), f(x[I - len(h)]), f(x[W - len(h)])  # from the right: source text may hold commas
        xs = f(x[X - len(h)]) if X is not None else 0.0
        agg.append((s, i, w, fname, x[0], x[1].strip(), xs))
        tot_s += s; tot_i += i
print(f"total samples {tot_s:.0f}  warp instructions {tot_i:.3e}")
for s, i, w, f, ln, code, xs in sorted(agg, key=lambda t: -t[0])[:top]:
    print(f"{100*s/tot_s:6.2f}% smp {100*i/tot_i:6.2f}% inst {w:12.0f} wf {xs:12.0f} excess  {f}:{ln:>4s}  {code[:90]}")
tot_x = sum(t[6] for t in agg)
if tot_x:
    print(f"-- shared-memory wavefronts in excess of the ideal (bank conflicts), by source line; total {tot_x:.0f}")
    for s, i, w, f, ln, code, xs in sorted(agg, key=lambda t: -t[6])[:20]:
        if xs: print(f"{100*xs/tot_x:6.2f}% of excess {xs:12.0f} of {w:12.0f} wf  {f}:{ln:>4s}  {code[:90]}")
