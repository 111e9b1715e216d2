"""SASS instruction counts of one kernel by source function: python tools/sass_by_function.py <nvdisasm -g -c listing of one kernel>"""
import collections
import os
import re
import sys

CSRC = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "raytracingweekend_b200", "csrc")


def funcs(f):
    p = os.path.join(CSRC, f)
    if not os.path.exists(p):
        return []
    out = []
    for i, l in enumerate(open(p), 1):
        m = re.match(r"^(?:template.*>\s*)?(?:static\s+)?(?:__device__|__global__|__host__|RTW_DEV|RTW_UNIFIED_FN)[^;{]*?\b(\w+)\s*\(", l)
        if m:
            out.append((i, m.group(1)))
    return out


cur, byline, n, subs = None, collections.Counter(), 0, []
for l in open(sys.argv[1]):
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4}\*/", l):
        n += 1
        byline[cur] += 1
    m = re.match(r"^(\$[\w\.\$]+):", l)
    if m:
        subs.append((n, m.group(1)[-70:]))
print("instructions", n)
for a, b in subs:
    print("  sub-function at", a, b)
files = collections.Counter()
for (f, _), c in byline.items():
    files[f] += c
for f, c in files.most_common():
    fs, agg = funcs(f), collections.Counter()
    for (ff, ln), k in byline.items():
        if ff != f:
            continue
        name = "?"
        for i, nm in fs:
            if i <= ln:
                name = f"{nm}@{i}"
        agg[name] += k
    print(f, c, agg.most_common(12))
if len(sys.argv) > 2:  # per-line listing of one file
    for (f, ln), c in sorted(byline.items()):
        if f == sys.argv[2]:
            print(ln, c)
