"""Markdown rows of DESIGN.md's "Measured" table from the bench lines of a session:
  python tools/measured_table.py gpurun_out      (reads bench_<workload>.json)"""
import json
import os
import sys


def main():
    d = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out"
    names = [("cornell", "C2 Cornell, 34 triangles"), ("torus1m", "C3 torus, 1 000 014 triangles"),
             ("composite4k", "C4 composite (290 triangles), 3840×2160, 64 spp"),
             ("glossy", "C5 without roulette / refraction (reference semantics), depth 32, 64 spp"),
             ("tor", "C1 `Tor.obj` 512², 20 spp, adaptive"),
             ("glossy_rr", "C5 as BASELINE names it: glossy + refractive, depth 32, roulette from bounce 3, 256 spp")]
    print("| workload | Grays/s (device-resident) | frame | Grays/s e2e (`pt_render`, host buffers, unseen scene) | same arrays again | reference CPU, all host threads |")
    print("|---|---|---|---|---|---|")
    for w, label in names:
        p = os.path.join(d, f"bench_{w}.json")
        if not os.path.exists(p):
            continue
        try:
            j = json.loads(open(p).read().strip().splitlines()[-1])
        except Exception:
            continue
        cpu = j.get("cpu_baseline") or {}
        c = j.get("e2e_cached") or {}
        print(f"| {label} | {j['value'] / 1e3:.2f} | {j['ms_per_step']:.1f} ms | {j['e2e']['value'] / 1e3:.2f} ({j['frame_ms_e2e']:.1f} ms) | "
              f"{c.get('value', 0) / 1e3:.2f} | {cpu.get('value', 0) / 1e3:.2g} ({cpu.get('cores')} threads, {cpu.get('kind')}) |")


if __name__ == "__main__":
    main()
