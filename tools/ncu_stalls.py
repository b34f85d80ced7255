#!/usr/bin/env python
"""Stall-reason breakdown, dynamic instruction mix and hottest SASS lines from an .ncu-rep (source page)."""
import csv, io, subprocess, sys
from collections import Counter
rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 15
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
r = list(csv.reader(io.StringIO(raw))); h = r[0]
for n in ("gpu__time_duration.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed",
          "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
          "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "smsp__warps_active.avg.per_cycle_active", "smsp__warps_eligible.avg.per_cycle_active",
          "smsp__inst_executed.sum", "launch__registers_per_thread", "l1tex__t_sector_hit_rate.pct"):
    if n in h: print(f"{n:75s} {r[2][h.index(n)]} {r[1][h.index(n)]}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hi = next(i for i, x in enumerate(rows) if x and x[0] == "Address")
h = rows[hi]; col = {n: i for i, n in enumerate(h)}
body = []
for x in rows[hi + 1:]:  # the first kernel of the report only (a second one repeats the header row)
    if len(x) == len(h) and x[0] == "Address":
        break
    if len(x) == len(h):
        body.append(x)
f = lambda x, n: float(x[col[n]] or 0)
tot = sum(f(x, "# Samples") for x in body)
print("static instructions", len(body), "samples", tot)
for n, v in sorted(((n, sum(f(x, n) for x in body)) for n in h if n.startswith("stall_") and "Not Issued" not in n), key=lambda t: -t[1])[:10]:
    print(f"  {n:26s} {100 * v / tot:6.2f}%")
c = Counter()
for x in body:
    s = x[col["Source"]].split()
    op = s[1] if s[0].startswith("@") else s[0]
    p = op.split(".")
    c[p[0] + ("." + p[1] if p[0] == "IMAD" and len(p) > 1 else "")] += f(x, "Instructions Executed")
T = sum(c.values())
print("dynamic mix:", ", ".join(f"{k} {100 * v / T:.1f}%" for k, v in c.most_common(18)))
print("hottest lines:")
for x in sorted(body, key=lambda x: -f(x, "# Samples"))[:top]:
    st = max(((n, f(x, n)) for n in h if n.startswith("stall_") and "Not Issued" not in n), key=lambda t: t[1])
    print(f"  {100 * f(x, '# Samples') / tot:5.2f}%  {x[col['Source']].strip()[:70]:70s} {st[0]}")
