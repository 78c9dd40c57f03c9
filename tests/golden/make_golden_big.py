"""Generates tests/golden/golden_big.npz: pins of BASELINE configs[2] (C3, the 1,000,014-triangle torus room) and
configs[3] (C4, Tor.obj + the two Cornell boxes) on the UNMODIFIED reference (oracle/_ref/libptref.so), plus the raw
`vn` of every Tor.obj face for the device's Triangle-constructor check.  Run in the build container (where
/root/reference exists; about two minutes on 8 cores: the reference is brute force, scene.cpp:114-120):
    python tests/golden/make_golden_big.py
Only rays / small frames are stored (the scenes are rebuilt by path-tracing_b200/scenes.py on the GPU box)."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, os.path.join(ROOT, "path-tracing_b200"))
from oracle_py import Reference  # noqa: E402
import scenes  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
MODELS = os.path.join(ROOT, "assets", "models")


def render_case(ref, name, W, H, spp, depth, error, seed=42, s_begin=0, s_step=1):
    ref.set_config(W, H, spp, depth, 1e-4, error)
    sm, s2, cnt, segs, _ = ref.render_accum(seed=seed, s_begin=s_begin, s_step=s_step)
    return {f"{name}_sum": sm, f"{name}_sum2": s2, f"{name}_count": cnt, f"{name}_segments": np.int64(segs),
            f"{name}_cfg": np.array([W, H, spp, depth, seed, s_begin, s_step], np.int64), f"{name}_error": np.float32(error)}


def torus_surface_rays(n, seed):
    """Rays aimed at the torus itself from inside the room (a uniform ray set mostly ends on the walls)."""
    rng = np.random.default_rng(seed)
    tt = scenes.torus_tris(200, 100).reshape(-1, 3, 3)
    pick = rng.integers(0, tt.shape[0], n)
    w = rng.dirichlet([1.0, 1.0, 1.0], n)
    target = np.einsum("nk,nkc->nc", w, tt[pick])
    org = rng.uniform([-9.5, -9.5, -20.5], [9.5, 9.5, 9.5], (n, 3))
    return np.concatenate([org, target - org], 1).astype(np.float32)


def obj_face_vn(path):
    """(n_faces, 3) float32: the `vn` the reference's loader applies to each face (scene.cpp:102-104: the normal index
    of the face's FIRST vertex), and a has_vn byte per face."""
    vns, out, has = [], [], []
    for line in open(path):
        tok = line.split()
        if not tok:
            continue
        if tok[0] == "vn":
            vns.append([np.float32(x) for x in tok[1:4]])
        elif tok[0] == "f":
            parts = tok[1].split("/")
            if len(parts) >= 3 and parts[2]:
                out.append(vns[int(parts[2]) - 1])
                has.append(1)
            else:
                out.append([0, 0, 0])
                has.append(0)
    return np.asarray(out, np.float32), np.asarray(has, np.uint8)


def main():
    ref = Reference()
    out = {}
    t0 = time.time()
    # ---- C3: torus_room(1000, 500), 1,000,014 triangles -------------------------------------------------------
    sc = scenes.torus_room(1000, 500)
    assert sc["verts"].shape[0] == 1000014
    ref.set_config(24, 14, 2, 8, 1e-4, -1.0)
    ref.from_arrays(sc["verts"], sc["mat_ids"], sc["mats"])
    rays = np.concatenate([scenes.camera_rays(48, 48), scenes.random_rays(2048, seed=31), torus_surface_rays(8192, 32)])
    ids, t = ref.intersect(rays)
    out["torus1m_rays"], out["torus1m_ids"], out["torus1m_t"] = rays, ids, t
    print(f"torus1m rays: {len(rays)} rays, {(ids >= 0).sum()} hits, {(ids < 1000000).sum() - (ids < 0).sum()} on the torus, "
          f"{time.time() - t0:.0f} s", flush=True)
    out.update(render_case(ref, "torus1m", 24, 14, 2, 8, -1.0))
    print(f"torus1m render: {int(out['torus1m_segments'])} segments, {time.time() - t0:.0f} s", flush=True)
    # ---- C4: composite (Tor.obj + the two Cornell boxes, 290 triangles) -----------------------------------------
    import ptb200
    tor = ptb200.HostLoader().load_obj(MODELS, "Tor.obj")
    sc = scenes.composite(tor)
    assert sc["verts"].shape[0] == 290
    ref.set_config(32, 32, 16, 8, 1e-4, -1.0)
    ref.from_arrays(sc["verts"], sc["mat_ids"], sc["mats"])
    rays = np.concatenate([scenes.camera_rays(48, 48), scenes.random_rays(4096, seed=33)])
    ids, t = ref.intersect(rays)
    out["composite_rays"], out["composite_ids"], out["composite_t"] = rays, ids, t
    out.update(render_case(ref, "composite", 32, 32, 16, 8, -1.0))
    out.update(render_case(ref, "composite_part", 32, 32, 16, 8, -1.0, s_begin=1, s_step=4))
    print(f"composite: {time.time() - t0:.0f} s", flush=True)
    # ---- Tor.obj's `vn` planes: raw vn per face + what the reference's loader made of them ------------------------
    vn, has = obj_face_vn(os.path.join(MODELS, "Tor.obj"))
    ref.set_config(64, 64, 1, 8)
    n = ref.load_obj(MODELS, "Tor.obj")
    planes, verts, squares = ref.dump()
    assert n == len(vn) == 270
    out.update(torvn_vn=vn, torvn_has=has, torvn_verts=verts, torvn_planes=planes, torvn_squares=squares)
    np.savez_compressed(os.path.join(OUT, "golden_big.npz"), **out)
    print("wrote golden_big.npz:", {k: getattr(v, "shape", None) for k, v in out.items()})


if __name__ == "__main__":
    main()
