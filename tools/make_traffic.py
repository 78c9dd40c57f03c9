"""profiles/traffic.json from ncu raw-page CSV exports of production-shaped launches (tools/gpu_session_r02.sh):
  python tools/make_traffic.py torus1m:gpurun_out/full_torus1m_raw.csv cornell:gpurun_out/full_cornell_raw.csv ...
Per workload and kernel class the MEDIAN launch by duration is taken (the first traversal launch walks coherent primary
rays and is not typical)."""
import csv
import json
import os
import statistics
import sys

HERE = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PATHS = 1920 * 1080 * 64  # path slots of the profiled step (tools/prof_one.py: 1080p, 64 samples per step)


def launches(path):
    rows = list(csv.reader(open(path, newline="")))
    hdr = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    names = rows[hdr]
    col = {n: i for i, n in enumerate(names)}
    out = []
    for r in rows[hdr + 2:]:
        if len(r) < len(names):
            continue
        f = lambda k: float(r[col[k]].replace(",", "")) if k in col and r[col[k]] not in ("", "n/a") else None
        unit = lambda k: rows[hdr + 1][col[k]] if k in col else ""
        scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}
        rd = f("dram__bytes_read.sum") * scale.get(unit("dram__bytes_read.sum"), 1.0)
        wr = f("dram__bytes_write.sum") * scale.get(unit("dram__bytes_write.sum"), 1.0)
        dur = f("gpu__time_duration.sum") * {"ms": 1.0, "us": 1e-3, "ns": 1e-6, "s": 1e3}.get(unit("gpu__time_duration.sum"), 1.0)
        out.append({"kernel": r[col["Kernel Name"]].split("<")[0].split("(")[0].replace("void ", "").replace("ptd::", "").strip(), "id": r[col["ID"]],
                    "ms": dur, "dram": rd + wr, "issue": f("smsp__issue_active.avg.pct_of_peak_sustained_active"),
                    "lanes": f("smsp__thread_inst_executed_per_inst_executed.ratio"), "l2": f("lts__t_sector_hit_rate.pct"),
                    "l1": f("l1tex__t_sector_hit_rate.pct"), "alu": f("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"),
                    "fma": f("sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active")})
    return out


def main():
    p = os.path.join(HERE, "profiles", "traffic.json")
    data = json.load(open(p))
    for arg in sys.argv[1:]:
        wl, path = arg.split(":", 1)
        by = {}
        for l in launches(path):
            by.setdefault(l["kernel"], []).append(l)
        data[wl] = {}
        for k, ls in by.items():
            ls.sort(key=lambda l: l["ms"])
            m = ls[len(ls) // 2]
            data[wl][k] = {"dram_bytes_per_launch": m["dram"], "paths_per_launch": PATHS, "launch_ms_under_ncu": round(m["ms"], 3),
                           "issue_active_pct": m["issue"], "lanes": m["lanes"], "l2_hit_pct": m["l2"], "l1_hit_pct": m["l1"],
                           "alu_pipe_pct": m["alu"], "fma_pipe_pct": m["fma"],
                           "source": f"profiles/r02_final_ncu_{wl}.txt, launch id {m['id']} (median of {len(ls)} profiled {k} launches, 1920x1080x64 slots)"}
    json.dump(data, open(p, "w"), indent=1)
    print(json.dumps(data, indent=1)[:3000])


if __name__ == "__main__":
    main()
