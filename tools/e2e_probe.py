import sys, time, os
sys.path.insert(0, "path-tracing_b200")
import numpy as np, ptb200, scenes
lib = ptb200.Library()
for name, sc in (("cornell", scenes.cornell()), ("torus1m", scenes.torus_room(1000, 500))):
    for spp in (8, 64):
        cfg = ptb200.config(1920, 1080, spp, 8, 1e-4, -1.0, seed=42)
        for i in range(3):
            t0 = time.perf_counter(); rgb = lib.render(sc, cfg); t1 = time.perf_counter()
            print(name, spp, "pt_render call", i, round((t1 - t0) * 1e3, 1), "ms", flush=True)
        t0 = time.perf_counter(); ds = lib.scene_create(sc, cfg); t1 = time.perf_counter()
        st = ds.stats(); print("  scene_create", round((t1 - t0) * 1e3, 1), "ms  upload", round(st["upload_ms"], 2), "build", round(st["build_ms"], 2))
        for i in range(2):
            t0 = time.perf_counter(); ds.render_device(cfg); st = ds.stats(); t1 = time.perf_counter()
            print("  render_device", i, round((t1 - t0) * 1e3, 1), "ms kernel", round(st["kernel_ms"], 1))
        ds.close()
