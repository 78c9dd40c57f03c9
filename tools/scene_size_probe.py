"""Does the traversal kernel's cost per unit of work depend on whether nodes + scan records + exact records fit L2?
The torus room at several tessellations (0.25 M .. 2 M triangles: 35 .. 280 MB of nodes and records), 1080p, 64 spp:
an instrumented pass for the work per queued ray, a timed pass for the time per k_wf_traverse launch.
  python tools/scene_size_probe.py"""
import os
import sys

HERE = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(HERE, "path-tracing_b200"))
import ptb200  # noqa: E402
import scenes  # noqa: E402


def main():
    lib = ptb200.Library()
    lines = []
    for nu, nv in ((500, 250), (700, 350), (1000, 500), (1400, 700)):
        sc = scenes.torus_room(nu, nv)
        n = int(sc["verts"].shape[0])
        ccfg = ptb200.config(1920, 1080, 8, 8, 1e-4, -1.0, seed=42)
        ds = lib.scene_create(sc, ccfg, counters=True)
        ds.render_device(ccfg)
        c = ds.stats()
        ds.close()
        q = max(1, c["rays_queued"])
        nodes, scans, exact = c["nodes_visited"] / q, (c["tris_tested"] - c["n_large"] * c["segments_traced"]) / q, c["exact_tests"] / q
        cfg = ptb200.config(1920, 1080, 64, 8, 1e-4, -1.0, seed=42)
        dt = lib.scene_create(sc, cfg, timed=True)
        for _ in range(2):
            dt.render_device(cfg)
            t = dt.stats()
        dt.close()
        lib.lib.pt_release_workspace()
        queued = q / c["segments_traced"] * t["segments_traced"]
        trav = t["class_ms"][1]
        mb = (c["n_nodes"] * 96 + 2 * 64 * n) / 1e6
        line = (f"{n:8d} triangles, {c['n_nodes']:7d} wide nodes, {mb:6.1f} MB nodes + scan + exact records | per queued ray: nodes {nodes:5.2f} "
                f"scans {scans:5.2f} exact {exact:4.2f} | queued {queued / 1e6:6.1f} M rays/step, traverse {trav:6.2f} ms/step = "
                f"{trav * 1e6 / queued:5.2f} ns/ray = {trav * 1e6 / queued / (nodes + scans + exact):5.3f} ns per node-or-record | "
                f"bounce {t['class_ms'][2]:6.2f} ms")
        print(line, flush=True)
        lines.append(line)
    os.makedirs(os.path.join(HERE, "gpurun_out"), exist_ok=True)
    open(os.path.join(HERE, "gpurun_out", "scene_size_probe.txt"), "a").write("\n".join(lines) + "\n")


if __name__ == "__main__":
    main()
