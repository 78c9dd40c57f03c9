"""BVH quality counters on the 1M-triangle scene (C3) — one instrumented frame at 1080p, 8 spp:
wide nodes / scan tests / exact tests per ray that walks the tree, tree size and build time.
  python tools/bvh_stats.py [variant-name ...]     (variants built by tools/ab_variants.sh; default: the in-tree library)"""
import os
import sys

HERE = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(HERE, "path-tracing_b200"))
import ptb200  # noqa: E402
import scenes  # noqa: E402


def main():
    big = scenes.torus_room(1000, 500)
    for name in sys.argv[1:] or [""]:
        if name:
            ptb200.RENDER_SO = os.path.join(HERE, "path-tracing_b200", "lib", "variants", f"libptrender_{name}.so")
        lib = ptb200.Library()
        cfg = ptb200.config(1920, 1080, 8, 8, 1e-4, -1.0, seed=42)
        ds = lib.scene_create(big, cfg, counters=True)
        ds.render_device(cfg)
        st = ds.stats()
        ds.close()
        q = max(1, st["rays_queued"])
        leaf_tris = st["tris_tested"] - st["n_large"] * st["segments_traced"]
        line = (f"{name or 'default':8s} wide nodes {st['n_nodes']} ({st['n_nodes'] * 96 / 1e6:.1f} MB), levels {st['n_levels']}, "
                f"build {st['build_ms']:.2f} ms, upload {st['upload_ms']:.2f} ms | per queued ray: nodes {st['nodes_visited'] / q:.2f}, "
                f"scan tests {leaf_tris / q:.2f}, exact tests {st['exact_tests'] / q:.2f} | queued {q / st['segments_traced']:.4f} of "
                f"{st['segments_traced']} segments | node histogram (log2 buckets) {st['node_hist'][:8]}")
        print(line, flush=True)
        with open(os.path.join(HERE, "gpurun_out", "bvh_stats.txt"), "a") as f:
            f.write(line + "\n")


if __name__ == "__main__":
    os.makedirs(os.path.join(HERE, "gpurun_out"), exist_ok=True)
    main()
