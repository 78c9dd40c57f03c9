"""Generates tests/golden/converged.npz: 4096-spp images of the UNMODIFIED reference (oracle/_ref/libptref.so, see
make_golden.py) under the Philox stream contract — the north_star's third check ("converged images at 4096 spp agree
within RMSE 1e-3").  Small frames: the reference is brute force (Tor.obj: 24x24 px x 4096 spp x ~8 segments x 270
triangles).  Run in the build container:  python tests/golden/make_converged.py"""
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
SPP = 4096


def case(ref, name, W, H, depth, seed):
    ref.set_config(W, H, SPP, depth, 1e-4, -1.0)
    t0 = time.time()
    sm, s2, cnt, segs, _ = ref.render_accum(seed=seed)
    print(f"{name}: {W}x{H} x {SPP} spp, {segs} segments, {time.time() - t0:.1f} s")
    return {f"{name}_sum": sm, f"{name}_count": cnt, f"{name}_segments": np.int64(segs),
            f"{name}_cfg": np.array([W, H, SPP, depth, seed], np.int64)}


def main():
    ref = Reference()
    out = {}
    ref.set_config(24, 24, SPP, 8)
    ref.load_obj(MODELS, "Tor.obj")
    out.update(case(ref, "tor", 24, 24, 8, 42))
    sc = scenes.cornell()
    ref.set_config(32, 18, SPP, 8, 1e-4, -1.0)
    ref.from_arrays(sc["verts"], sc["mat_ids"], sc["mats"])
    out.update(case(ref, "cornell", 32, 18, 8, 7))
    np.savez_compressed(os.path.join(OUT, "converged.npz"), **out)


if __name__ == "__main__":
    main()
