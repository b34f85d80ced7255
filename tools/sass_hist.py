#!/usr/bin/env python
"""Static SASS opcode histogram of one kernel of libmkfhe_b200.so, split at BAR/loop boundaries (CPU only, cuobjdump).

    python tools/sass_hist.py <substring of the mangled kernel name> [lo_addr hi_addr]
"""
import collections, re, subprocess, sys
lib = "mkfhe_b200/libmkfhe_b200.so"
pat = sys.argv[1]
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
cur, lines = None, []
for l in out.splitlines():
    m = re.search(r"Function : (\S+)", l)
    if m:
        cur = m.group(1)
        continue
    if cur and pat in cur:
        m = re.match(r"\s+/\*([0-9a-f]+)\*/\s+(.*?);", l)
        if m:
            lines.append((int(m.group(1), 16), m.group(2).strip()))
lo = int(sys.argv[2], 16) if len(sys.argv) > 2 else 0
hi = int(sys.argv[3], 16) if len(sys.argv) > 3 else 1 << 40
h = collections.Counter()
for a, t in lines:
    if lo <= a < hi:
        t = re.sub(r"^@!?U?P\d+\s+", "", t)
        op = t.split()[0]
        parts = op.split(".")
        key = parts[0]
        if key == "IMAD" and len(parts) > 1 and parts[1] in ("HI", "WIDE", "MOV", "IADD", "SHL"):
            key += "." + parts[1]
        if key in ("LDS", "STS", "LDG") and "128" in op:
            key += ".128"
        h[key] += 1
tot = sum(h.values())
print(f"{tot} instructions in [{lo:#x}, {hi:#x})")
for k, v in h.most_common(40):
    print(f"  {k:12s} {v:6d}  {100.0 * v / tot:5.1f}%")
back = [(a, t) for a, t in lines if "BRA" in t and re.search(r"0x([0-9a-f]+)", t) and int(re.search(r"0x([0-9a-f]+)", t).group(1), 16) < a]
print("backward branches (loop ends):", ", ".join(f"{a:#x}->{re.search(r'0x[0-9a-f]+', t).group(0)}" for a, t in back[-12:]))
