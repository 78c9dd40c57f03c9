"""PT_ISECT_WATERTIGHT against the reference's intersector: hit-ID mismatch rates on every golden ray set (answers of
the unmodified reference, tests/golden/*.npz) and on the edge / vertex / grazing stress rays of the parity tests
(answers of the device's reference-exact mode), 4096-spp RMSE against the reference's converged images, and C3
throughput in both modes.  Writes gpurun_out/watertight_report.txt (copied to profiles/ and cited in DESIGN.md).
  python tools/watertight_report.py"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for d in ("path-tracing_b200", "tests", "oracle"):
    sys.path.insert(0, os.path.join(HERE, d))
import ptb200  # noqa: E402
import scenes  # noqa: E402
from test_gpu_parity import _stress_rays  # noqa: E402
from test_oracle import tor_scene  # noqa: E402

G = np.load(os.path.join(HERE, "tests", "golden", "golden.npz"))
GB = np.load(os.path.join(HERE, "tests", "golden", "golden_big.npz"))
CV = np.load(os.path.join(HERE, "tests", "golden", "converged.npz"))
WT = ptb200.PT_ISECT_WATERTIGHT


def main():
    lib = ptb200.Library()
    tor = ptb200.HostLoader().load_obj(os.path.join(HERE, "assets", "models"), "Tor.obj")
    out = ["# PT_ISECT_WATERTIGHT (Woop-style test of the geometric triangle) against the reference's plane + area-sum test",
           "# hit-ID mismatch = rays whose nearest-hit triangle differs; dt = max relative difference of t where the IDs agree",
           f"{'ray set':34s} {'rays':>8s} {'mismatch':>9s} {'rate':>8s} {'of which hit<->miss':>20s} {'max dt (same id)':>17s}"]
    sets = [("tor_cam (Tor.obj, vn planes)", tor_scene(), G["tor_cam_rays"], G["tor_cam_ids"], G["tor_cam_t"]),
            ("tor_rnd (Tor.obj, vn planes)", tor_scene(), G["tor_rnd_rays"], G["tor_rnd_ids"], G["tor_rnd_t"]),
            ("cornell", scenes.cornell(), G["cornell_rays"], G["cornell_ids"], G["cornell_t"]),
            ("glossy", scenes.glossy_mix(), G["glossy_rays"], G["glossy_ids"], G["glossy_t"]),
            ("torus40 (1,614 triangles)", scenes.torus_room(40, 20), G["torus40_rays"], G["torus40_ids"], G["torus40_t"]),
            ("torus1m (C3, 1,000,014 triangles)", scenes.torus_room(1000, 500), GB["torus1m_rays"], GB["torus1m_ids"], GB["torus1m_t"]),
            ("composite (C4, 290 triangles)", scenes.composite(tor), GB["composite_rays"], GB["composite_ids"], GB["composite_t"])]
    for name, sc in (("cornell", scenes.cornell()), ("torus40", scenes.torus_room(40, 20)), ("composite", scenes.composite(tor))):
        rays = _stress_rays(sc, 400000, 11)
        ids, t = lib.intersect(sc, rays, ptb200.config())
        sets.append((f"stress rays: {name} (edges, vertices, grazing)", sc, rays, ids, t))
    for name, sc, rays, ids, t in sets:
        wi, wt = lib.intersect(sc, rays, ptb200.config(intersector=WT))
        bi, bt = lib.intersect(sc, rays, ptb200.config(intersector=WT, accel=ptb200.PT_ACCEL_BRUTE))
        assert np.array_equal(wi, bi) and np.array_equal(wt.view(np.uint32), bt.view(np.uint32)), "BVH != brute force in watertight mode"
        bad = wi != ids
        same = ~bad & (ids >= 0)
        dt = float(np.max(np.abs(wt[same] - t[same]) / np.maximum(np.abs(t[same]), 1e-9))) if same.any() else 0.0
        out.append(f"{name:34s} {len(ids):8d} {int(bad.sum()):9d} {100 * bad.mean():7.3f}% {int(((wi < 0) != (ids < 0)).sum()):20d} {dt:17.2e}")
    out.append("")
    out.append("# converged images, 4096 spp (north_star bar: RMSE 1e-3 against the reference's image)")
    for name, sc in (("tor", tor_scene()), ("cornell", scenes.cornell())):
        W, H, spp, depth, seed = map(int, CV[name + "_cfg"])
        sm, cnt = CV[name + "_sum"], CV[name + "_count"]
        ref = np.where(cnt[..., None] > 0, sm / np.maximum(cnt, 1)[..., None].astype(np.float32), np.float32(0))
        for label, isect in (("reference-exact", ptb200.PT_ISECT_REFERENCE), ("watertight", WT)):
            rgb = lib.render(sc, ptb200.config(W, H, spp, depth, 1e-4, -1.0, seed=seed, intersector=isect))
            rmse = float(np.sqrt(np.mean((rgb.astype(np.float64) - ref) ** 2)))
            out.append(f"{name:8s} {W}x{H} {label:16s} RMSE {rmse:.3e}   mean radiance {rgb.mean():.4f} (reference {ref.mean():.4f})")
    out.append("")
    out.append("# C3 throughput, 1080p 128 spp depth 8, 3 timed frames")
    big = scenes.torus_room(1000, 500)
    for label, isect in (("reference-exact", ptb200.PT_ISECT_REFERENCE), ("watertight", WT)):
        cfg = ptb200.config(1920, 1080, 128, 8, 1e-4, -1.0, seed=42, intersector=isect)
        ds = lib.scene_create(big, cfg, timed=True)
        ms, cls = [], np.zeros(4)
        for i in range(4):
            ds.render_device(cfg)
            st = ds.stats()
            if i:
                ms.append(st["kernel_ms"]); cls += np.array(st["class_ms"][:4])
        out.append(f"{label:16s} frame {np.mean(ms):7.2f} ms  {st['segments'] / np.mean(ms) / 1e3:7.0f} Mrays/s  "
                   + "  ".join(f"{c} {v / 3:.2f}" for c, v in zip(("begin", "traverse", "bounce", "commit"), cls)))
        ds.close()
    text = "\n".join(out)
    print(text)
    os.makedirs(os.path.join(HERE, "gpurun_out"), exist_ok=True)
    open(os.path.join(HERE, "gpurun_out", "watertight_report.txt"), "w").write(text + "\n")


if __name__ == "__main__":
    main()
