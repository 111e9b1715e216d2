"""Summarise an `ncu --page source --csv --print-source cuda,sass` export per source line and per SASS opcode.

    ncu -i X.ncu-rep --page source --csv --print-source cuda,sass > /tmp/x.csv
    python tools/ncu_hotspots.py /tmp/x.csv [top_n]

Prints, for every source file in the export, the lines ranked by executed warp instructions with their share of
warp-stall samples and average active threads, then an opcode histogram weighted by executed instructions.
"""
import csv
import sys
from collections import defaultdict


def main():
    path = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
    rows = list(csv.reader(open(path)))
    cur_file = None
    header = None
    lines = {}  # (file, line) -> [text, inst, thread_inst, samples]
    ops = defaultdict(lambda: [0, 0, 0])
    cur = None
    for r in rows:
        if not r:
            continue
        if r[0] in ("File Name", "File Path"):
            cur_file = r[1].split("/")[-1]
            continue
        if r[0] == "Function Name":
            cur_file = cur_file or "kernels.cu"
            continue
        if r[0] == "Line No":
            header = r
            ix = {name: i for i, name in enumerate(header)}
            i_inst = header.index("Instructions Executed")
            i_tinst = header.index("Thread Instructions Executed")
            i_samp = header.index("# Samples")
            continue
        if header is None or len(r) < len(header):
            continue
        if r[0] != "":  # source line row
            cur = (cur_file, int(r[0]))
            lines[cur] = [r[1].strip(), 0, 0, 0]
            continue
        # sass row belonging to cur
        try:
            inst, tinst, samp = int(r[i_inst]), int(r[i_tinst]), int(r[i_samp])
        except ValueError:
            continue
        if cur:
            lines[cur][1] += inst
            lines[cur][2] += tinst
            lines[cur][3] += samp
        op = r[3].strip().split()
        if op and op[0].startswith("@"):
            op = op[1:]
        name = op[0].split(".")[0] if op else "?"
        ops[name][0] += inst
        ops[name][1] += tinst
        ops[name][2] += samp
    tot_inst = sum(v[1] for v in lines.values()) or 1
    tot_samp = sum(v[3] for v in lines.values()) or 1
    print(f"total warp instructions {tot_inst:,}  samples {tot_samp:,}")
    by_file = defaultdict(lambda: [0, 0])
    for (f, _), v in lines.items():
        by_file[f][0] += v[1]
        by_file[f][1] += v[3]
    for f, (i, s) in sorted(by_file.items(), key=lambda kv: -kv[1][0]):
        print(f"  {f:20s} inst {100 * i / tot_inst:5.1f}%  samples {100 * s / tot_samp:5.1f}%")
    print(f"\n{'file:line':26s} {'inst%':>6s} {'samp%':>6s} {'thr':>5s}  source")
    for (f, ln), v in sorted(lines.items(), key=lambda kv: -kv[1][3])[:top]:
        thr = v[2] / v[1] if v[1] else 0
        print(f"{f + ':' + str(ln):26s} {100 * v[1] / tot_inst:6.2f} {100 * v[3] / tot_samp:6.2f} {thr:5.1f}  {v[0][:110]}")
    # where the divergence is: warp instructions (issue slots) by how many lanes were active on their source line
    buckets = [[0, 0, 0] for _ in range(4)]
    for v in lines.values():
        if v[1]:
            b = min(3, int(v[2] / v[1] / 8.0))
            buckets[b][0] += v[1]
            buckets[b][1] += v[2]
            buckets[b][2] += v[3]
    print(f"\n{'active lanes':14s} {'inst%':>6s} {'samp%':>6s} {'thread-inst%':>13s}")
    tot_t = sum(b[1] for b in buckets) or 1
    for k, b in enumerate(buckets):
        print(f"{8 * k:3d}-{8 * k + 8:<10d} {100 * b[0] / tot_inst:6.2f} {100 * b[2] / tot_samp:6.2f} {100 * b[1] / tot_t:13.2f}")
    print(f"\n{'opcode':12s} {'inst%':>6s} {'samp%':>6s} {'thr':>5s}")
    for name, (i, t, s) in sorted(ops.items(), key=lambda kv: -kv[1][0])[:30]:
        print(f"{name:12s} {100 * i / tot_inst:6.2f} {100 * s / tot_samp:6.2f} {t / i if i else 0:5.1f}")


if __name__ == "__main__":
    main()
