"""The small case run under compute-sanitizer (one --tool per gpurun call: memcheck, racecheck, initcheck):
Tor.obj 64x64x16 with the adaptive rule on, both pipelines, alive lists on and off, 8-wide BVH and brute force, a
Russian-roulette render, a row band, a resumed render, the watertight intersector, pt_intersect, the post stage, and
renders on a strided grid of the per-slot kernels.
  compute-sanitizer --tool memcheck python tools/sanitize_case.py"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(HERE, "path-tracing_b200"))
import ptb200  # noqa: E402
import scenes  # noqa: E402


def main():
    lib = ptb200.Library()
    tor = ptb200.HostLoader().load_obj(os.path.join(HERE, "assets", "models"), "Tor.obj")
    mesh = scenes.torus_room(24, 12)
    n = 0
    for sc in (tor, mesh):
        for mode in (0, 1):
            for alive in (1, 2):
                for accel in (ptb200.PT_ACCEL_LBVH, ptb200.PT_ACCEL_BRUTE):
                    cfg = ptb200.config(64, 64, 16, 8, 1e-4, 0.001, seed=42, mode=mode, alive=alive, accel=accel)
                    sm, s2, cnt = lib.render_accum(sc, cfg)
                    n += int(cnt.sum())
    sm, s2, cnt, wt = lib.render_accum_w(scenes.glass_mix(), ptb200.config(48, 32, 8, 16, 1e-4, -1.0, seed=3, rr_depth=2))
    band = lib.render_accum(tor, ptb200.config(64, 64, 8, 8, 1e-4, 0.001, rows=(10, 30)))
    st = lib.render_accum(tor, ptb200.config(64, 64, 6, 8, 1e-4, 0.001))
    st = lib.render_accum(tor, ptb200.config(64, 64, 12, 8, 1e-4, 0.001, sample_begin=6, resume=1), st)
    wtm = lib.render_accum(mesh, ptb200.config(64, 64, 8, 8, 1e-4, -1.0, intersector=ptb200.PT_ISECT_WATERTIGHT))
    ids, t = lib.intersect(mesh, scenes.random_rays(20000, seed=1))
    img, stats = lib.render_image(tor, ptb200.config(64, 64, 8, 8, 1e-4, 0.001), lib.post_config(64, 64, gauss=1))
    img, stats = lib.postprocess(sm, s2, cnt, lib.post_config(48, 32, median=1), weight=wt)
    big = scenes.torus_room(250, 200)  # 100k triangles: a tree of several levels
    lib.render_accum(big, ptb200.config(160, 90, 16, 8, 1e-4, -1.0))
    lib.render_accum(big, ptb200.config(160, 90, 40, 8, 1e-4, 0.001, q_shift=3, alive=1))
    os.environ["PT_WF_GRID_CAP"] = "64"  # the per-slot kernels stride over the step's blocks (many passes per block)
    lib.render_accum(big, ptb200.config(160, 90, 16, 8, 1e-4, -1.0))
    lib.render_accum(big, ptb200.config(160, 90, 40, 8, 1e-4, 0.001, q_shift=3, alive=1))
    lib.render_accum_w(scenes.glass_mix(), ptb200.config(48, 32, 8, 16, 1e-4, -1.0, seed=3, rr_depth=2))
    del os.environ["PT_WF_GRID_CAP"]
    lib.lib.pt_release_workspace()
    print(f"sanitize case done: {n} contributing paths, {int((ids >= 0).sum())} hits")
    try:
        print("bounds checks:", lib.debug_checks())
    except ptb200.PtError as e:
        print("bounds checks: not a checked build:", e)


if __name__ == "__main__":
    main()
