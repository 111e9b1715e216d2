"""Basic-block view of an `ncu --page source --csv --print-source cuda,sass` export.

    python tools/ncu_blocks.py /tmp/x.csv [min_share_pct]

SASS instructions are de-duplicated by address (the cuda view repeats an instruction under every inlined frame), then
consecutive instructions with the same execution count are merged into blocks.  For each block: number of
instructions, executions per kernel-wide "iteration" (normalised by the most executed block), average active lanes,
share of all issued warp instructions, share of stall samples, and the source lines it came from.
"""
import csv
import sys
from collections import OrderedDict, Counter


def main():
    path = sys.argv[1]
    min_share = float(sys.argv[2]) if len(sys.argv) > 2 else 0.5
    rows = list(csv.reader(open(path)))
    header = None
    cur_file, cur_line = None, None
    sass = OrderedDict()  # addr -> [text, inst, tinst, samples, Counter(lines)]
    for r in rows:
        if not r:
            continue
        if r[0] in ("File Name", "File Path"):
            cur_file = r[1].split("/")[-1]
            continue
        if r[0] == "Line No":
            header = r
            i_addr, i_inst = header.index("Address"), header.index("Instructions Executed")
            i_t, i_s = header.index("Thread Instructions Executed"), header.index("# Samples")
            continue
        if header is None or len(r) <= i_t:
            continue
        if r[0] != "":
            cur_line = f"{cur_file}:{r[0]}"
            continue
        try:
            addr, inst, t, s = int(r[i_addr], 16), int(r[i_inst]), int(r[i_t]), int(r[i_s])
        except ValueError:
            continue
        e = sass.setdefault(addr, [r[3].strip(), inst, t, s, Counter()])
        e[4][cur_line] += 1
    addrs = sorted(sass)
    blocks = []
    for a in addrs:
        txt, inst, t, s, lines = sass[a]
        if blocks and blocks[-1]["exec"] == inst and blocks[-1]["thr"] == t and blocks[-1]["end"] + 16 == a:
            b = blocks[-1]
            b["n"] += 1
            b["samples"] += s
            b["lines"].update(lines)
            b["end"] = a
        else:
            blocks.append({"start": a, "end": a, "n": 1, "exec": inst, "thr": t, "samples": s, "lines": Counter(lines)})
    tot = sum(b["n"] * b["exec"] for b in blocks) or 1
    tot_s = sum(b["samples"] for b in blocks) or 1
    top_exec = max(b["exec"] for b in blocks if b["n"] >= 4) or 1
    base = addrs[0]
    print(f"SASS instructions {len(addrs)}, issued warp instructions {tot:,}, normalising executions by {top_exec:,}")
    print(f"{'pc':>6s} {'n':>4s} {'exec/it':>8s} {'lanes':>6s} {'issue%':>7s} {'samp%':>6s}  lines")
    for b in blocks:
        share = 100.0 * b["n"] * b["exec"] / tot
        if share < min_share:
            continue
        lanes = b["thr"] / b["exec"] if b["exec"] else 0
        ls = ", ".join(k for k, _ in b["lines"].most_common(4))
        print(f"{b['start'] - base:6x} {b['n']:4d} {b['exec'] / top_exec:8.3f} {lanes:6.1f} {share:7.2f} {100.0 * b['samples'] / tot_s:6.2f}  {ls}")


if __name__ == "__main__":
    main()
