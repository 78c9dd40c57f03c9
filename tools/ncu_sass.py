#!/usr/bin/env python
"""Compact per-instruction listing from `ncu --page source --csv --print-source sass` (optionally .gz):
  python tools/ncu_sass.py FILE.csv[.gz] KERNEL_SUBSTRING [occurrence] [min_share_pct]
Prints offset, executed warp instructions, average active threads, stall samples and the SASS text."""
import csv
import gzip
import sys


def sections(path):
    op = gzip.open if path.endswith(".gz") else open
    with op(path, "rt", newline="") as f:
        cur = None
        for row in csv.reader(f):
            if row and row[0] == "Kernel Name":
                cur = {"name": row[1], "hdr": None, "rows": []}
                yield cur
            elif cur is not None and cur["hdr"] is None:
                cur["hdr"] = row
            elif cur is not None:
                cur["rows"].append(row)


def main():
    path, sub = sys.argv[1], sys.argv[2]
    occ = int(sys.argv[3]) if len(sys.argv) > 3 else 0
    min_share = float(sys.argv[4]) if len(sys.argv) > 4 else 0.0
    secs = [s for s in sections(path) if sub in s["name"]]
    # ncu prints each kernel twice (two views); keep those with rows
    secs = [s for s in secs if s["rows"]]
    s = secs[occ]
    h = {n: i for i, n in enumerate(s["hdr"])}
    tot_inst = sum(int(r[h["Instructions Executed"]]) for r in s["rows"])
    tot_thr = sum(int(r[h["Thread Instructions Executed"]]) for r in s["rows"])
    tot_smp = sum(int(r[h["# Samples"]]) for r in s["rows"])
    print(f"# {s['name']}  occurrence {occ} of {len(secs)}: {tot_inst} warp instr, {tot_thr} thread instr, "
          f"avg active {tot_thr / max(1, tot_inst):.2f}, {tot_smp} samples")
    base = int(s["rows"][0][h["Address"]], 16)
    for r in s["rows"]:
        n = int(r[h["Instructions Executed"]])
        share = 100.0 * n / max(1, tot_inst)
        smp = int(r[h["# Samples"]])
        if share < min_share and 100.0 * smp / max(1, tot_smp) < min_share:
            continue
        off = int(r[h["Address"]], 16) - base
        print(f"{off:6x} {n:12d} {share:5.2f}% thr {float(r[h['Avg. Threads Executed']]):5.1f} smp {smp:7d} "
              f"{100.0 * smp / max(1, tot_smp):5.2f}%  {r[h['Source']].strip()}")


if __name__ == "__main__":
    main()
