"""Generates tests/golden/*.npz from the UNMODIFIED reference (oracle/_ref/libptref.so = /root/reference's scene.cpp,
triangles.h, material.h, ray.h compiled with the glm/RNG shims; see oracle/Makefile).  Run in the build container
(where /root/reference exists):  python tests/golden/make_golden.py
The fixtures travel to the GPU box; the reference does not."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, os.path.join(ROOT, "path-tracing_b200"))
from oracle_py import Oracle, Reference  # noqa: E402
import scenes  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
MODELS = os.path.join(ROOT, "assets", "models")


def render_case(ref, name, W, H, spp, depth, error, seed=42, s_begin=0, s_step=1):
    ref.set_config(W, H, spp, depth, 1e-4, error)
    sm, s2, cnt, segs, _ = ref.render_accum(seed=seed, s_begin=s_begin, s_step=s_step)
    return {f"{name}_sum": sm, f"{name}_sum2": s2, f"{name}_count": cnt, f"{name}_segments": np.int64(segs),
            f"{name}_cfg": np.array([W, H, spp, depth, seed, s_begin, s_step], np.int64), f"{name}_error": np.float32(error)}


def render_case_sky(ref, name, W, H, spp, depth, bmp):
    sm, s2, cnt, segs, _ = ref.render_accum(seed=42)
    return {f"{name}_sum": sm, f"{name}_sum2": s2, f"{name}_count": cnt, f"{name}_segments": np.int64(segs),
            f"{name}_cfg": np.array([W, H, spp, depth, 42, 0, 1], np.int64), f"{name}_error": np.float32(-1.0)}


def main():
    ref, orc = Reference(), Oracle()
    out = {}
    # ---- Tor.obj through the reference's own loader ----
    ref.set_config(64, 64, 1, 8)
    n = ref.load_obj(MODELS, "Tor.obj")
    planes, verts, squares = ref.dump()
    out.update(tor_n=np.int64(n), tor_planes=planes, tor_verts=verts, tor_squares=squares)
    cam = scenes.camera_rays(64, 64)
    rnd = scenes.random_rays(8192, seed=1)
    for nm, rays in (("cam", cam), ("rnd", rnd)):
        ids, t = ref.intersect(rays)
        out[f"tor_{nm}_rays"], out[f"tor_{nm}_ids"], out[f"tor_{nm}_t"] = rays, ids, t
    prim = np.array([[0, 0, 0], [5, 7, 1], [63, 63, 3], [31, 2, 19], [17, 40, 255]], np.int32)
    ref.set_config(64, 64, 256, 8)
    out["prim_xys"] = prim
    out["prim_rays"] = np.array([np.concatenate(ref.primary_ray(42, *map(int, p))) for p in prim], np.float32)
    out.update(render_case(ref, "tor_flat", 32, 32, 16, 8, -1.0))
    out.update(render_case(ref, "tor_adapt", 32, 32, 24, 8, 0.001))
    out.update(render_case(ref, "tor_part", 32, 32, 16, 8, -1.0, s_begin=1, s_step=2))
    # ---- synthetic scenes through the reference's Triangle/Factory constructors ----
    for nm, sc, depth in (("cornell", scenes.cornell(), 8), ("glossy", scenes.glossy_mix(), 32)):
        ref.set_config(32, 32, 16, depth, 1e-4, -1.0)
        ref.from_arrays(sc["verts"], sc["mat_ids"], sc["mats"])
        p, v, s = ref.dump()
        out[f"{nm}_planes"], out[f"{nm}_squares"] = p, s
        rays = scenes.random_rays(4096, seed=2)
        ids, t = ref.intersect(rays)
        out[f"{nm}_rays"], out[f"{nm}_ids"], out[f"{nm}_t"] = rays, ids, t
        out.update(render_case(ref, nm, 32, 32, 16, depth, -1.0))
    # small torus mesh with vn-free planes: BVH vs brute-force reference answer
    sc = scenes.torus_room(40, 20)
    ref.set_config(32, 32, 4, 8, 1e-4, -1.0)
    ref.from_arrays(sc["verts"], sc["mat_ids"], sc["mats"])
    rays = np.concatenate([scenes.random_rays(4096, seed=3), scenes.camera_rays(48, 48)])
    ids, t = ref.intersect(rays)
    out["torus40_rays"], out["torus40_ids"], out["torus40_t"] = rays, ids, t
    out.update(render_case(ref, "torus40", 32, 32, 4, 8, -1.0))
    # skybox (scene.cpp:125-156): open room, synthetic equirect image written as BMP for the reference's own loader
    sky = scenes.sky_image()
    bmp = os.path.join(OUT, "_sky_tmp.bmp")
    scenes.write_bmp(bmp, sky)
    sc = scenes.open_scene()
    ref.set_config(32, 32, 8, 8, 1e-4, -1.0, bmp)
    ref.from_arrays(sc["verts"], sc["mat_ids"], sc["mats"])
    out["sky_bgr"] = sky
    out.update(render_case_sky(ref, "sky", 32, 32, 8, 8, bmp))
    os.unlink(bmp)
    # after the loop (main.cpp:162-201): the accumulators of two render cases above -> float image (the oracle's
    # restatement of the inline resolve, host libm powf) -> the REFERENCE's own GaussBlur / MedianFilter (main.cpp
    # compiled into oracle/_ref) -> 8 bits.  Also a synthetic image with values over the whole 0..255 range.
    for nm in ("tor_adapt", "cornell"):
        color, img, st = orc.resolve_display(out[f"{nm}_sum"], out[f"{nm}_sum2"], out[f"{nm}_count"], 1 / 2.2)
        out[f"post_{nm}_rgb8"], out[f"post_{nm}_stats"] = img, np.array(st, np.float32)
        for r in (1, 2):
            out[f"post_{nm}_gauss{r}"] = orc.to_rgb8(ref.gauss_blur(color, r), out[f"{nm}_count"])
        for w in (1, 2, 3):
            out[f"post_{nm}_median{w}"] = orc.to_rgb8(ref.median_filter(color, w), out[f"{nm}_count"])
    rng = np.random.default_rng(12)
    img = (rng.random((20, 28, 3)) * 255).astype(np.float32)
    out["post_img"] = img
    out["post_img_gauss2"], out["post_img_gauss3"] = ref.gauss_blur(img, 2), ref.gauss_blur(img, 3)
    out["post_img_median1"], out["post_img_median2"] = ref.median_filter(img, 1), ref.median_filter(img, 2)
    # Philox4x32-10 known answers (Random123 kat_vectors) through the oracle's implementation
    out["philox_kat"] = np.array([orc.philox([0, 0, 0, 0], [0, 0]), orc.philox([0xffffffff] * 4, [0xffffffff] * 2),
                                  orc.philox([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344], [0xa4093822, 0x299f31d0])],
                                 np.uint32)
    np.savez_compressed(os.path.join(OUT, "golden.npz"), **out)
    print("wrote golden.npz:", {k: getattr(v, "shape", None) for k, v in out.items()})


if __name__ == "__main__":
    main()
