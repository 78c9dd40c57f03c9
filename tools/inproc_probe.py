"""pt_render with PtConfig.n_gpus = 1, 2, 4, 8 inside ONE process (host buffers in, image out): ms per call.
Run on a multi-GPU box: gpurun --gpus 2 -- 'python tools/inproc_probe.py'"""
import json, sys, time
sys.path.insert(0, "path-tracing_b200")
import ptb200, scenes
lib = ptb200.Library()
have = lib.device_count()
out = {}
for name, sc in (("cornell", scenes.cornell()), ("torus1m", scenes.torus_room(1000, 500))):
    for n in (1, 2, 4, 8):
        if n > have:
            break
        cfg = ptb200.config(1920, 1080, 256, 8, 1e-4, -1.0, seed=42, n_gpus=n)
        ms = []
        for i in range(4):
            t0 = time.perf_counter(); lib.render(sc, cfg); ms.append((time.perf_counter() - t0) * 1e3)
        out[f"{name}_n{n}"] = [round(m, 1) for m in ms]
        print(name, "n_gpus", n, "pt_render ms per call", out[f"{name}_n{n}"], flush=True)
json.dump(out, open("gpurun_out/inproc_probe.json", "w"))
