"""Instruction-cache footprint of one kernel launch from `ncu --page source --csv --print-source sass`:
    python tools/icache_footprint.py <export.csv>
128-byte lines (8 instructions) by how often they were executed, relative to the most executed line."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
h = rows[hi]
ia, ie = h.index("Address"), h.index("Instructions Executed")
ins = [(int(r[ia], 16), int(r[ie])) for r in rows[hi + 1:] if len(r) > ie and r[ia].startswith("0x")]
base = ins[0][0]
lines = {}
for a, e in ins:
    lines.setdefault((a - base) // 128, []).append(e)
mx = max(max(v) for v in lines.values())
tot = sum(e for _, e in ins)
print(f"{len(ins)} instructions, {len(lines)} lines of 128 B = {len(lines) * 128 / 1024:.1f} KB; {tot / 1e9:.2f} G warp instructions")
for thr in (0, 1e-5, 1e-4, 1e-3, 1e-2, 0.05, 0.1, 0.3):
    hot = [k for k, v in lines.items() if max(v) > thr * mx]
    share = sum(sum(lines[k]) for k in hot) / tot
    print(f"  lines executed more than {thr:7.5f} x max: {len(hot):4d} = {len(hot) * 128 / 1024:5.1f} KB, {100 * share:6.2f} % of executed instructions")
