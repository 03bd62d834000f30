#!/usr/bin/env python
"""Top stall-sample SASS instructions of a kernel in an .ncu-rep (source page).
usage: ncu_hot.py rep kernel-regex [top]"""
import csv, io, subprocess, sys
rep, rx = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + rx], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
h = rows[hi]
col = {c: i for i, c in enumerate(h)}
seen, data = set(), []
for r in rows[hi + 1:]:
    if len(r) < len(h) or r[0] in seen:
        continue
    seen.add(r[0])
    try:
        s = int(r[col["# Samples"]] or 0)
    except ValueError:
        continue
    stalls = {c[6:]: int(r[i] or 0) for c, i in col.items() if c.startswith("stall_") and "Not Issued" not in c and (r[i] or "0").isdigit() and int(r[i] or 0) > 0}
    data.append((s, r[0][-5:], r[col["Source"]].strip()[:70], r[col["Instructions Executed"]], stalls))
tot = sum(d[0] for d in data)
print("total samples", tot)
for s, a, src, ex, st in sorted(data, reverse=True)[:top]:
    print(f"{100.0 * s / tot:5.1f}% {a} {src:70s} x{ex} {dict(sorted(st.items(), key=lambda kv: -kv[1])[:3])}")
