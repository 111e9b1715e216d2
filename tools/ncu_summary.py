"""Fold the counters bench.py quotes (DRAM bytes per launch, issue slots busy, FP32 share of executed instructions, lanes per
instruction) out of an `ncu --page details` text export into profiles/ncu_summary.json, keyed by bench config:

    ncu -i cap.ncu-rep --page details > profiles/r2_<what>_ncu_details.txt
    ncu -i cap.ncu-rep --page raw --csv > profiles/r2_<what>_ncu_raw.csv
    python tools/ncu_summary.py cfg4 profiles/r2_<what>_ncu_details.txt [profiles/r2_<what>_ncu_raw.csv] ["command line of the capture"]

bench.py reads roofline.traffic / roofline.issue_slots_busy_ncu from that file, so they follow the committed capture
instead of being literals in the source."""
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "profiles", "ncu_summary.json")
UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def field(txt, name, units=None):
    m = re.search(r"^\s*" + re.escape(name) + r"\s+(?:(\S+)\s+)?([0-9][0-9.,]*)\s*$", txt, flags=re.M)
    if not m:
        return None
    v = float(m.group(2).replace(",", ""))
    return v * UNIT.get(m.group(1) or "", 1.0) if units else v


def raw_csv(path):
    """`ncu --page raw --csv` reduced to metric,unit,value rows (tools/ncu_raw_select.sh) or ncu's own wide CSV."""
    import csv
    rows = list(csv.reader(open(path)))
    out = {}
    if rows and rows[0][:3] == ["metric", "unit", "value"]:
        for r in rows[1:]:
            if len(r) >= 3 and r[2]:
                out[r[0]] = float(r[2]) * UNIT.get(r[1], 1.0)
    elif len(rows) >= 3:  # header, units, one row per launch: take the last launch
        for name, unit, val in zip(rows[0], rows[1], rows[-1]):
            try:
                out[name] = float(val.replace(",", "")) * UNIT.get(unit, 1.0)
            except ValueError:
                pass
    return out


def main():
    cfg, path = sys.argv[1], sys.argv[2]
    txt = open(path).read()
    s = {"source": os.path.relpath(os.path.abspath(path), ROOT)}
    rest = sys.argv[3:]
    rd, wr = field(txt, "dram__bytes_read.sum", True), field(txt, "dram__bytes_write.sum", True)
    if rest and rest[0].endswith(".csv"):  # raw page of the same capture
        raw = raw_csv(rest[0])
        rd, wr = raw.get("dram__bytes_read.sum", rd), raw.get("dram__bytes_write.sum", wr)
        s["raw_source"] = os.path.relpath(os.path.abspath(rest[0]), ROOT)
        rest = rest[1:]
    if rest:
        s["command"] = rest[0]
    if rd is not None and wr is not None:
        s["dram_bytes_per_launch"] = int(rd + wr)
    busy = field(txt, "Issue Slots Busy")
    if busy is not None:
        s["issue_slots_busy"] = busy / 100.0
    for key, name in (("duration_ms", "Duration"), ("registers", "Registers Per Thread"), ("achieved_occupancy_pct", "Achieved Occupancy"),
                      ("executed_ipc_active", "Executed Ipc Active"), ("avg_active_threads_per_warp", "Avg. Active Threads Per Warp"),
                      ("l1_hit_rate_pct", "L1/TEX Hit Rate"), ("l2_hit_rate_pct", "L2 Hit Rate")):
        v = field(txt, name)
        if v is not None:
            s[key] = v
    m = re.search(r"([0-9.]+)-way bank conflict", txt)
    k = re.search(r"^\s*(render_\w+<[^>]*>)", txt, flags=re.M)
    if k:
        s["kernel"] = k.group(1)
    allsum = {}
    if os.path.exists(OUT):
        allsum = json.load(open(OUT))
    allsum[cfg] = s
    json.dump(allsum, open(OUT, "w"), indent=1, sort_keys=True)
    print(json.dumps(s, indent=1))


if __name__ == "__main__":
    main()
