#!/usr/bin/env python
"""Executed warp instructions, active lanes and stall samples per SOURCE LINE of one profiled launch: joins
`ncu --page source --csv --print-source sass` (optionally .gz) with `nvdisasm -gi -c` of the same cubin (build with
-lineinfo; `cuobjdump -xelf all lib.so` extracts it) on the instruction offset.
  python tools/ncu_lines.py SASS.csv[.gz] DISASM.txt KERNEL_SUBSTRING [occurrence] [min_share_pct]
Three tables: by the chain of inlined functions (which device function the instructions belong to), by the line of the
kernel body (outermost frame of the inline chain) and by the innermost line inside this repository's sources."""
import bisect
import os
import collections
import re
import sys

sys.path.insert(0, __import__("os").path.dirname(__file__))
from ncu_sass import sections  # noqa: E402


def disasm_lines(path, mangled_sub):
    """offset -> list of (file, line) frames, innermost first, for the function whose label contains mangled_sub"""
    out, cur, frames, pending = {}, None, [], []
    for ln in open(path, errors="replace"):
        m = re.match(r"\s*\.section\s+\.text\.(\S+?),", ln)
        if m:
            cur = m.group(1)
            frames, pending = [], []
            continue
        if cur is None or mangled_sub not in cur:
            continue
        m = re.match(r'\s*//## File "([^"]+)", line (\d+)(?: inlined at "([^"]+)", line (\d+))?', ln)
        if m:
            pending.append((m.group(1), int(m.group(2))))
            continue
        m = re.match(r"\s*/\*([0-9a-f]{4,6})\*/\s+(.*)", ln)
        if m:
            if pending:
                frames, pending = pending, []
            out[int(m.group(1), 16)] = frames
    return out


def function_table(csrc):
    """file name -> sorted [(first line, function name)] of the top-level functions of the .cuh files"""
    out = {}
    for f in os.listdir(csrc):
        fs = []
        for i, l in enumerate(open(os.path.join(csrc, f), errors="replace"), 1):
            m = re.match(r"^(k_\w+)\(", l) or re.match(r"^(?:__device__|static|inline)[^;]*?\b(\w+)\s*\(", l)
            if m:
                fs.append((i, m.group(1)))
        out[f] = fs
    return out


def function_of(ftab, f, line):
    f = f.split("/")[-1]
    fs = ftab.get(f)
    if not fs:
        return f
    k = bisect.bisect_right([x[0] for x in fs], line) - 1
    return fs[k][1] if k >= 0 else f


def main():
    path, dis, sub = sys.argv[1], sys.argv[2], sys.argv[3]
    occ = int(sys.argv[4]) if len(sys.argv) > 4 else 0
    min_share = float(sys.argv[5]) if len(sys.argv) > 5 else 0.3
    secs = [s for s in sections(path) if sub in s["name"]]  # (the generator fills the rows after it yields)
    secs = [s for s in secs if s["rows"]]
    s = secs[occ]
    mang = re.sub(r"[^A-Za-z0-9_]", "", sub.split("<")[0].split("::")[-1])
    tmpl = sys.argv[6] if len(sys.argv) > 6 else mang
    loc = disasm_lines(dis, tmpl)
    h = {n: i for i, n in enumerate(s["hdr"])}
    base = int(s["rows"][0][h["Address"]], 16)
    ftab = function_table(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "path-tracing_b200", "csrc"))
    chains = collections.defaultdict(lambda: [0, 0, 0])
    outer, inner = collections.defaultdict(lambda: [0, 0, 0]), collections.defaultdict(lambda: [0, 0, 0])
    tot = [0, 0, 0]
    for r in s["rows"]:
        off = int(r[h["Address"]], 16) - base
        n, thr, smp = int(r[h["Instructions Executed"]]), int(r[h["Thread Instructions Executed"]]), int(r[h["# Samples"]])
        fr = loc.get(off) or [("?", 0)]
        own = [f for f in fr if "/csrc/" in f[0]] or fr
        o, i = own[-1], own[0]
        chain = " > ".join(function_of(ftab, f, l) for f, l in reversed(own))
        for d, k in ((outer, o), (inner, i), (chains, chain)):
            d[k][0] += n; d[k][1] += thr; d[k][2] += smp
        tot[0] += n; tot[1] += thr; tot[2] += smp
    print(f"# {s['name'][:110]}  occurrence {occ} of {len(secs)}: {tot[0]} warp instr, avg active {tot[1] / max(1, tot[0]):.2f}, {tot[2]} samples")
    print("## by chain of inlined functions")
    for k, (n, thr, smp) in sorted(chains.items(), key=lambda kv: -kv[1][0]):
        if 100.0 * n / tot[0] >= min_share:
            print(f"  instr {100.0 * n / tot[0]:6.2f}%  lanes {thr / max(1, n):5.1f}  stall samples {100.0 * smp / max(1, tot[2]):6.2f}%  {k}")
    for title, d in (("by line of the kernel body (outermost frame)", outer), ("by innermost line in csrc/", inner)):
        print(f"## {title}")
        for (f, l), (n, thr, smp) in sorted(d.items(), key=lambda kv: -kv[1][0]):
            if 100.0 * n / tot[0] < min_share and 100.0 * smp / max(1, tot[2]) < min_share:
                continue
            print(f"  {f.split('/')[-1]:18s}:{l:4d}  instr {100.0 * n / tot[0]:6.2f}%  lanes {thr / max(1, n):5.1f}  stall samples {100.0 * smp / max(1, tot[2]):6.2f}%")


if __name__ == "__main__":
    main()
