"""SASS listing of a pc range from an `ncu --page source --csv --print-source cuda,sass` export, with execution counts
and stall samples:  python tools/ncu_sass.py x.csv 0x7380 0x7620   (offsets as printed by tools/ncu_blocks.py)"""
import csv
import sys


def main():
    path, lo, hi = sys.argv[1], int(sys.argv[2], 16), int(sys.argv[3], 16)
    header, sass = None, {}
    for r in csv.reader(open(path)):
        if not r:
            continue
        if r[0] == "Line No":
            header = r
            i_addr, i_inst = header.index("Address"), header.index("Instructions Executed")
            i_t, i_s = header.index("Thread Instructions Executed"), header.index("# Samples")
            continue
        if header is None or len(r) <= i_t or r[0] != "":
            continue
        try:
            sass.setdefault(int(r[i_addr], 16), (r[3].strip(), int(r[i_inst]), int(r[i_t]), int(r[i_s])))
        except ValueError:
            continue
    base = min(sass)
    for a in sorted(sass):
        if lo <= a - base < hi:
            txt, inst, t, s = sass[a]
            print(f"{a - base:6x} {inst:12d} {t / inst if inst else 0:5.1f} {s:7d}  {txt}")


if __name__ == "__main__":
    main()
