"""pt_scene_create of the 1M-triangle scene, twice (pool warm the second time): run under
  ncu --metrics gpu__time_duration.sum --clock-control none --csv python tools/prof_build.py
for the duration of every build kernel."""
import os
import sys

HERE = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(HERE, "path-tracing_b200"))
import ptb200  # noqa: E402
import scenes  # noqa: E402

lib = ptb200.Library()
sc = scenes.torus_room(1000, 500)
cfg = ptb200.config(1920, 1080, 8, 8, 1e-4, -1.0)
for i in range(2):
    ds = lib.scene_create(sc, cfg)
    st = ds.stats()
    print(f"create {i}: build {st['build_ms']:.2f} ms, upload {st['upload_ms']:.2f} ms, {st['n_nodes']} wide nodes", flush=True)
    ds.close()
