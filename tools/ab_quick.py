"""Quick A/B of compile-time variants built by `tools/ab_variants.sh build`, in ONE process (a GPU-box call of well
under a minute):  python tools/ab_quick.py base top128 top320
For each variant: accumulators of two small renders must equal the first variant's bit for bit (the first one is the
default build, which the goldens pin), then the 1M-triangle scene (C3) at 1080p, 128 spp: 1 warm-up + 3 timed
frames with per-class event timing."""
import os
import sys
import time

HERE = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(HERE, "path-tracing_b200"))
import numpy as np  # noqa: E402
import ptb200  # noqa: E402
import scenes  # noqa: E402

CLASSES = ("begin", "traverse", "bounce", "commit")


def run_variant(name, big, small, state):
    ptb200.RENDER_SO = os.path.join(HERE, "path-tracing_b200", "lib", "variants", f"libptrender_{name}.so")
    lib = ptb200.Library()
    outs = [lib.render_accum(small, ptb200.config(64, 64, 16, 8, 1e-4, -1.0, seed=42)),
            lib.render_accum(big, ptb200.config(320, 180, 8, 8, 1e-4, -1.0, seed=42)),
            lib.render_accum(big, ptb200.config(96, 54, 24, 8, 1e-4, 0.001, seed=7))]
    flat = [np.ascontiguousarray(a).view(np.uint32).ravel() for o in outs for a in o]
    if state["ref"] is None:
        state["ref"], parity = flat, "reference"
    else:
        parity = "bit-equal" if all(np.array_equal(a, b) for a, b in zip(flat, state["ref"])) else "DIFFERS"
    cfg = ptb200.config(1920, 1080, 128, 8, 1e-4, -1.0, seed=42)
    ds = lib.scene_create(big, cfg, timed=True)
    ms, cls = [], np.zeros(4)
    for i in range(4):
        ds.render_device(cfg)
        st = ds.stats()
        if i:
            ms.append(st["kernel_ms"])
            cls += np.array(st["class_ms"][:4])
    segs = st["segments"]
    ds.close()
    # C2 (Cornell, every triangle in shared memory): the per-slot kernels alone
    cfg2 = ptb200.config(1920, 1080, 128, 8, 1e-4, -1.0, seed=42)
    ds2 = lib.scene_create(state["cornell"], cfg2)
    ms2 = []
    for i in range(3):
        ds2.render_device(cfg2)
        st2 = ds2.stats()
        if i:
            ms2.append(st2["kernel_ms"])
    ds2.close()
    # C5 as BASELINE names it (glass_mix, depth 32, roulette from bounce 3): alive lists, most launches nearly empty; and the
    # reference's default frame (Tor.obj-like small frame: launch-bound)
    extra = ""
    for tag, sc3, cfg3 in (("glossy_rr", state["glass"], ptb200.config(1920, 1080, 128, 32, 1e-4, -1.0, seed=42, rr_depth=3)),
                           ("small512", small, ptb200.config(512, 512, 20, 8, 1e-4, 0.001, seed=42))):
        ds3 = lib.scene_create(sc3, cfg3)
        ms3 = []
        for i in range(3):
            ds3.render_device(cfg3)
            st3 = ds3.stats()
            if i:
                ms3.append(st3["kernel_ms"])
        ds3.close()
        extra += f"  | {tag} {np.mean(ms3):.2f} ms"
    lib.lib.pt_release_workspace()
    return (f"{name:8s} parity {parity:9s} frame {np.mean(ms):7.2f} ms (min {min(ms):.2f})  "
            f"{segs / np.mean(ms) / 1e3:7.0f} Mrays/s  " + "  ".join(f"{c} {v / 3:.2f}" for c, v in zip(CLASSES, cls)) +
            f"  | cornell {st2['segments'] / np.mean(ms2) / 1e3:7.0f} Mrays/s" + extra)


def main():
    names = sys.argv[1:] or ["base"]
    t00 = time.time()
    big, small = scenes.torus_room(1000, 500), scenes.torus_room(40, 20)
    print(f"scenes built in {time.time() - t00:.1f} s", flush=True)
    state = {"ref": None, "cornell": scenes.cornell(), "glass": scenes.glass_mix()}
    lines = []
    for name in names:
        try:
            line = run_variant(name, big, small, state)
        except Exception as e:  # a broken variant must not take the rest of the session with it
            line = f"{name:8s} FAILED: {e}"
        print(line, flush=True)
        lines.append(line)
    os.makedirs(os.path.join(HERE, "gpurun_out"), exist_ok=True)
    with open(os.path.join(HERE, "gpurun_out", "ab_quick.txt"), "a") as f:
        f.write("\n".join(lines) + "\n")
    print(f"total {time.time() - t00:.1f} s")


if __name__ == "__main__":
    main()
