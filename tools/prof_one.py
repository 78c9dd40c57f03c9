"""One production-shaped wavefront step for profiling: C3 (1,000,014 triangles) — or another scenes.py scene — at 1080p,
64 samples per pixel (one full step of 64 samples per pixel, depth 8).  Run plain first, then under ncu:
  python tools/prof_one.py [torus1m|cornell|glossy] [spp]"""
import os
import sys

HERE = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(HERE, "path-tracing_b200"))
import ptb200  # noqa: E402
import scenes  # noqa: E402


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "torus1m"
    spp = int(sys.argv[2]) if len(sys.argv) > 2 else 64
    sc = {"torus1m": lambda: scenes.torus_room(1000, 500), "cornell": scenes.cornell, "glossy": scenes.glossy_mix}[name]()
    lib = ptb200.Library()
    cfg = ptb200.config(1920, 1080, spp, 8, 1e-4, -1.0, seed=42)
    ds = lib.scene_create(sc, cfg)
    ds.render_device(cfg)
    st = ds.stats()
    print(f"{name}: {st['segments']} segments in {st['kernel_ms']:.2f} ms = {st['segments'] / st['kernel_ms'] / 1e3:.0f} Mrays/s, "
          f"{st['launches']} launches")
    ds.close()


if __name__ == "__main__":
    main()
