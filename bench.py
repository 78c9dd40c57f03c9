#!/usr/bin/env python
"""bench.py — Mrays/s and frame time of the render loop on N B200s (one process per GPU), with the reference's CPU
renderer timed beside it.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload cornell|torus1m|tor|glossy] [--impl reference]

A step = one full frame of the workload (all samples of all pixels) rendered from a device-resident scene into
device-resident accumulators; with N > 1 ranks the samples of every pixel are split across ranks (sample s -> rank
s mod N) and the per-rank accumulators are summed with one NCCL reduce inside the step.  1 ray = one
Scene::TraceRay-equivalent path segment (scene.cpp:113-157), counted by the kernel.
Prints ONE JSON line (rank 0).  Nothing here reads /root/reference at run time.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "path-tracing_b200"))

import numpy as np  # noqa: E402

WORKLOADS = {
    # name: (description, width, height, spp, depth, error)
    "cornell": ("configs[1]: synthetic Cornell box, 34 tris, 1080p 256 spp depth 8", 1920, 1080, 256, 8, -1.0),
    "torus1m": ("configs[2]: subdivided torus 1,000,014 tris, 1080p 256 spp depth 8, GPU LBVH", 1920, 1080, 256, 8, -1.0),
    "tor": ("configs[0]: models/Tor.obj 512x512 20 spp depth 8, adaptive error 0.001", 512, 512, 20, 8, 0.001),
    "composite4k": ("configs[3]: Tor.obj + the two Cornell boxes, 290 tris, 4K 1024 spp depth 8 (spp split across the ranks)", 3840, 2160, 1024, 8, -1.0),
    "glossy": ("configs[4]: mixed glossy materials, 1080p 64 spp depth 32 (no Russian roulette)", 1920, 1080, 64, 32, -1.0),
    "glossy_rr": ("configs[4]: mixed glossy + refractive materials, 1080p 256 spp depth 32, Russian roulette from bounce 3 "
                  "(extensions: no reference semantics, CPU leg = the oracle port)", 1920, 1080, 256, 32, -1.0),
}
RR_DEPTH = {"glossy_rr": 3}  # PtConfig.rr_depth per workload (0 = off)


def make_scene(name):
    import ptb200
    import scenes
    if name == "cornell":
        return scenes.cornell()
    if name == "torus1m":
        return scenes.torus_room(1000, 500)
    if name == "glossy":
        return scenes.glossy_mix()
    if name == "glossy_rr":
        return scenes.glass_mix()
    if name == "composite4k":
        return scenes.composite(ptb200.HostLoader().load_obj(os.path.join(ROOT, "assets", "models"), "Tor.obj"))
    if name == "tor":
        return ptb200.HostLoader().load_obj(os.path.join(ROOT, "assets", "models"), "Tor.obj")
    raise SystemExit(f"unknown workload {name}")


def peaks():
    try:
        p = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json)", float(p.get("sm_max_mhz", 1965.0))
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)", 1965.0


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--id={index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                       "-lms", "200"], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            pass

    def stop(self):
        if self.p:
            self.p.terminate()
            try:
                self.p.wait(timeout=5)
            except Exception:
                self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        for line in self.f.read().splitlines():
            c = [x.strip() for x in line.split(",")]
            if len(c) < 7:
                continue
            try:
                sm.append(float(c[0])); mx.append(float(c[1]))
            except ValueError:
                continue
            for nm, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), c[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        os.unlink(self.f.name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        # "under load": samples in the upper half of what was seen (idle samples before/after are dropped)
        hi = [v for v in sm if v >= 0.5 * max(sm)]
        return {"sm_mhz": statistics.median(hi), "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm)}


def host_threads():
    """CPUs this process may run on.  NOT omp_get_max_threads(): torchrun exports OMP_NUM_THREADS=1 to its workers, which
    made round 1's multi-rank reference arm single-threaded; the harnesses take an explicit num_threads clause."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


REF_RELEASE_THREADS = 4  # the reference's own release build: THREADS_TO_RUN=4 (CMakeLists.txt:10)


def cpu_reference_leg(name, sc, W, H, spp, depth, error, budget_s=12.0, threads=0, rr_depth=0):
    """The reference's own CPU renderer (oracle/_ref: unmodified scene.cpp/triangles.h/material.h/ray.h + shims) when it
    was compiled, else the oracle's C port, on a bounded sample of the workload: the same scene and parameters at a
    reduced image size / spp, sized by a calibration pass to about `budget_s` seconds."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    from oracle_py import Oracle, Reference
    orc = Oracle()
    planes, squares = (sc["planes"], sc["squares"]) if sc.get("planes") is not None else (None, None)
    n_tri = sc["verts"].shape[0]
    # the unmodified reference has neither roulette nor refraction: those workloads run on the oracle's port
    use_ref = Reference.available() and rr_depth == 0 and sc.get("mats_ext") is None
    if use_ref:
        ref = Reference()
        threads = threads or host_threads()

        def run(w, h, s, nthreads=None):
            ref.set_config(w, h, s, depth, 1e-4, error)
            if planes is None:
                ref.from_arrays(sc["verts"], sc["mat_ids"], sc["mats"])
            else:  # scenes that came from the OBJ loader carry `vn` planes: rebuild through the reference's loader
                ref.load_obj(os.path.join(ROOT, "assets", "models"), "Tor.obj")
            _, _, _, segs, sec = ref.render_accum(seed=42, threads=nthreads or threads)
            return segs, sec
    else:
        threads = threads or host_threads()
        if planes is None:
            planes, squares = np.zeros((n_tri, 4), np.float32), np.zeros(n_tri, np.float32)
            import ctypes as C
            sq = C.c_float()
            zero = np.zeros(3, np.float32)
            v = np.ascontiguousarray(sc["verts"], np.float32)
            for i in range(n_tri):
                orc.lib.or_triangle_setup(v[i].copy(), 0, zero, planes[i], C.byref(sq))
                squares[i] = sq.value
        osc = dict(sc, planes=planes, squares=squares)

        def run(w, h, s, nthreads=None):
            _, _, _, segs, sec = orc.render_accum(osc, orc.config(w, h, s, depth, 1e-4, error, 42, rr_depth=rr_depth), nthreads or threads)
            return segs, sec
    # calibration: a thumbnail with the frame's aspect ratio, 1 sample
    w, h, s = max(16, W // 40), max(9, H // 40), 1
    if n_tri > 100000:
        w, h = 8, 4
    segs, sec = run(w, h, s)
    w2, h2, s2 = w, h, s
    for _ in range(4):  # grow the sample towards the time budget (tiny samples under-estimate the rate)
        if sec >= 0.4 * budget_s or (w2 == W and h2 == H and s2 == spp):
            break
        want = (segs / max(sec, 1e-6)) * budget_s          # segments that fit the budget at the measured rate
        per_px = segs / (w2 * h2 * s2)                      # segments per path
        scale = (want / per_px / (w * h)) ** 0.5            # thumbnail scale if it were 1 spp
        w2, h2 = int(min(W, max(w, w * scale))), int(min(H, max(h, h * scale)))
        s2 = int(max(1, min(spp, want / (per_px * w2 * h2))))
        segs, sec = run(w2, h2, s2)
    # the same sample at the reference's own release thread count (CMakeLists.txt:10: THREADS_TO_RUN=4), scaled down so
    # that it takes about a quarter of the budget
    rel = None
    if threads > REF_RELEASE_THREADS:
        frac = min(1.0, 0.25 * budget_s / max(sec, 1e-6) * REF_RELEASE_THREADS / threads)
        s4 = max(1, int(s2 * frac))
        h4 = h2 if s2 * frac >= 1 else max(1, int(h2 * s2 * frac))
        segs4, sec4 = run(w2, h4, s4, REF_RELEASE_THREADS)
        rel = {"value": segs4 / sec4 / 1e6, "unit": "Mrays/s", "cores": REF_RELEASE_THREADS,
               "what": f"THREADS_TO_RUN=4 (the reference's release setting): {w2}x{h4}, {s4} spp, {segs4} segments in {sec4:.2f} s"}
    return {"value": segs / sec / 1e6, "unit": "Mrays/s", "cores": int(threads), "kind": "reference" if use_ref else "port",
            "release_threads": rel, "host_cpus": host_threads(),
            "sample": f"{name}: same scene/depth at {w2}x{h2}, {s2} spp ({segs} segments in {sec:.2f} s); "
                      f"full frame extrapolates to {(W * H * spp) * (segs / (w2 * h2 * s2)) / (segs / sec) :.0f} s",
            "segments": int(segs), "seconds": sec}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    desc, W, H, spp, depth, error = WORKLOADS[args.workload]
    sc = make_scene(args.workload)
    runs = []
    per_step_budget = max(2.0, min(20.0, 120.0 / max(1, args.steps + args.warmup)))
    for i in range(args.warmup + args.steps):
        r = cpu_reference_leg(args.workload, sc, W, H, spp, depth, error, budget_s=per_step_budget, rr_depth=RR_DEPTH.get(args.workload, 0))
        if i >= args.warmup:
            runs.append(r)
    segs = sum(r["segments"] for r in runs)
    sec = sum(r["seconds"] for r in runs)
    v = segs / sec / 1e6
    last = runs[-1]
    out = {"impl": "reference", "metric": "Mrays/s", "value": v, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps,
           "warmup": args.warmup, "ms_per_step": 1e3 * sec / len(runs), "higher_is_better": True, "scaling": "strong",
           "vs_baseline": None, "dtype": "f32", "data": "synthetic",
           "config": {"workload": desc, "sample": last["sample"]},
           "cpu_baseline": {"value": v, "unit": "Mrays/s", "cores": last["cores"], "kind": last["kind"], "sample": last["sample"],
                            "release_threads": last["release_threads"], "host_cpus": last["host_cpus"]},
           "e2e": {"value": v, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
           "gpu_launches": 0}
    emit(out)


_JSON_FD = None


def quiet_stdout():
    """The contract is ONE JSON line on stdout, but native libraries write there too (NCCL prints its version on the
    first communicator when NCCL_DEBUG is set): from here on fd 1 is stderr, and emit() writes to the real stdout."""
    global _JSON_FD
    if _JSON_FD is None:
        sys.stdout.flush()
        _JSON_FD = os.dup(1)
        os.dup2(2, 1)


INPROC_TIMEOUT_S = 240  # the in-process N-GPU leg of a --gpus N run (an extra beside the line's own numbers)


def emit(obj):
    sys.stdout.flush()
    line = (json.dumps(obj) + "\n").encode()
    os.write(_JSON_FD if _JSON_FD is not None else 1, line)


def main():
    quiet_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default="torus1m", choices=sorted(WORKLOADS),
                    help="default: BASELINE configs[2], the configuration the >= 1 Gray/s target is quoted on")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--spp", type=int, default=0, help="override samples per pixel (profiling runs only)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg (profiling runs only)")
    ap.add_argument("--accel", type=int, default=0)
    ap.add_argument("--mode", type=int, default=0, help="0 wavefront pipeline (default), 1 megakernel variant")
    ap.add_argument("--error", type=float, default=None, help="override the adaptive threshold (tuning runs only)")
    ap.add_argument("--q-shift", type=int, default=0, help="log2 samples per wavefront step (0 = automatic; tuning runs only)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference_arm(args)

    import torch
    import torch.distributed as dist
    import ptb200

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the render library has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    desc, W, H, spp, depth, error = WORKLOADS[args.workload]
    if args.spp:
        spp = args.spp
    if args.error is not None:
        error = args.error
    sc = make_scene(args.workload)
    lib = ptb200.Library()
    rr = RR_DEPTH.get(args.workload, 0)
    cfg = ptb200.config(W, H, spp, depth, 1e-4, error, seed=42, sample_begin=rank, sample_step=world, device=local,
                        accel=args.accel, mode=args.mode, q_shift=args.q_shift, rr_depth=rr)
    ds = lib.scene_create(sc, cfg)
    n_px = W * H
    acc = torch.zeros(n_px * 4, device="cuda")            # [sum rgb (3*n_px) | count as f32 (n_px)] reduced together
    d_sum = acc[:3 * n_px]
    d_cnt_i = torch.zeros(n_px, dtype=torch.int32, device="cuda")
    d_rgb = torch.zeros(n_px * 3, device="cuda")
    d_wt_t = torch.zeros(n_px, device="cuda") if rr else None  # the estimator's divisor under Russian roulette
    d_wt = d_wt_t.data_ptr() if rr else 0
    h_rgb = torch.empty(n_px * 3, pin_memory=True) if world > 1 and rank == 0 else None
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")  # > 126 MB L2
    stream = torch.cuda.current_stream().cuda_stream
    launches_per_step = 0
    lib_launches = [1]

    def step():
        nonlocal launches_per_step
        flush.fill_(1)                                    # L2 flush between iterations
        ds.render_device(cfg, d_sum.data_ptr(), 0, d_cnt_i.data_ptr(), stream, d_wt)
        n = 1
        if world > 1:
            # the divisor travels with the sums: counts (exact in f32 up to 2^24 samples), or under Russian roulette
            # the weight sums (the resolve then reads them instead of the counts, which stay this rank's own)
            acc[3 * n_px:].copy_(d_wt_t if rr else d_cnt_i)
            dist.reduce(acc, 0, op=dist.ReduceOp.SUM)     # one NCCL reduce over NVLink
            (d_wt_t if rr else d_cnt_i).copy_(acc[3 * n_px:])
        if rank == 0:
            ds.resolve_device(d_sum.data_ptr(), d_cnt_i.data_ptr(), n_px, d_rgb.data_ptr(), stream, d_wt)
            n += 1
        launches_per_step = n - 1 + ds_launches()

    def ds_launches():
        return lib_launches[0]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step()
    barrier()
    lib_launches[0] = ds.stats()["launches"]             # kernels one render call launches (counted by the library)
    seg_step = torch.tensor([ds.stats()["segments"]], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(seg_step)
    segments_per_step = float(seg_step.item())

    sampler = ClockSampler(local) if rank == 0 else None
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    kernel_ms = []
    barrier()
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    barrier()
    kernel_ms.append(ds.stats()["kernel_ms"])             # last step's render-kernel time (events on the same stream)
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = float(t.item())
    clocks = sampler.stop() if sampler else None
    ms_per_step = total_ms / args.steps
    mrays = segments_per_step / (ms_per_step * 1e-3) / 1e6

    # ---- end to end through the drop-in entry point: HOST scene -> pt_render -> HOST rgb.  The library keeps the built
    # device scene between calls, keyed by a hash of the caller's arrays, so the timed calls come in two kinds:
    #   e2e         every call gets a scene it has not seen (one vertex moved by an ulp): hash, upload, BVH build, render,
    #               resolve and download are all inside the timed region — the strict reading of the contract;
    #   e2e_cached  the same arrays again, as a frame loop passes them: hash + render + resolve + download.
    # With N ranks every rank renders its sample partition through pt_scene_create + pt_scene_render_device and the
    # accumulators are reduced with NCCL (scene upload and build inside the timed region, on every rank).
    e2e_steps = max(1, min(args.steps, 3))
    h2d = int(sc["verts"].nbytes + sc["mat_ids"].nbytes + sc["mats"].nbytes +
              (sc["planes"].nbytes + sc["squares"].nbytes if sc.get("planes") is not None else 0))
    d2h = n_px * 3 * 4

    def nudge(scene):
        """a scene the library has not seen: one coordinate moved to the neighbouring float (the content hash changes)"""
        v = scene["verts"].reshape(-1)
        k = nudge.flip % v.size  # a different coordinate every call: no two calls of the run see the same arrays
        v[k] = np.nextafter(v[k], np.float32(np.inf))
        nudge.flip += 1
    nudge.flip = 0

    def e2e_call(fresh=True):
        if world == 1:
            if fresh:
                nudge(sc_pinned)
            return lib.render(sc_pinned, cfg, out=rgb_pinned)
        nudge(sc_pinned)  # every rank moves the same coordinate: the ranks keep rendering one scene
        h = lib.scene_create(sc_pinned, cfg)
        h.render_device(cfg, d_sum.data_ptr(), 0, d_cnt_i.data_ptr(), stream, d_wt)
        acc[3 * n_px:].copy_(d_wt_t if rr else d_cnt_i)
        dist.reduce(acc, 0, op=dist.ReduceOp.SUM)
        out = None
        if rank == 0:
            (d_wt_t if rr else d_cnt_i).copy_(acc[3 * n_px:])
            h.resolve_device(d_sum.data_ptr(), d_cnt_i.data_ptr(), n_px, d_rgb.data_ptr(), stream, d_wt)
            h_rgb.copy_(d_rgb, non_blocking=True)  # pinned host buffer
            out = h_rgb
        torch.cuda.synchronize()
        h.close()
        return out

    # the host side of the call: scene arrays and the image live in page-locked host memory, as a host that renders
    # frame after frame would keep them (and as the bench contract words it)
    sc_pinned, rgb_pinned = sc, None
    pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory().numpy()
    sc_pinned = {k: (pin(v) if isinstance(v, np.ndarray) else v) for k, v in sc.items()}
    if world == 1:
        rgb_pinned = torch.zeros((H, W, 3), dtype=torch.float32).pin_memory().numpy()

    def timed_calls(fn, n):
        barrier()
        t0, calls = time.perf_counter(), []
        for _ in range(n):
            tc = time.perf_counter()
            fn()
            calls.append({"ms": round((time.perf_counter() - tc) * 1e3, 1), "phases_ms": lib.last_phases() if world == 1 else None})
        barrier()
        sec = torch.tensor([(time.perf_counter() - t0) / n], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(sec, op=dist.ReduceOp.MAX)
        return float(sec.item()), calls

    if args.warmup > 0:
        for _ in range(2):  # untimed: the library keeps two built scenes per device, each with its path-state workspace
            e2e_call()      # (GBs of HBM); from the third new scene on the oldest one leaves first and its blocks are reused
    e2e_sec, e2e_calls = timed_calls(e2e_call, e2e_steps)
    e2e_mrays = segments_per_step / e2e_sec / 1e6
    e2e_cached = None
    if world == 1:
        e2e_call(fresh=False)
        c_sec, c_calls = timed_calls(lambda: e2e_call(fresh=False), e2e_steps)
        e2e_cached = {"value": segments_per_step / c_sec / 1e6, "unit": "Mrays/s", "frame_ms": c_sec * 1e3, "h2d_bytes_per_step": 0,
                      "d2h_bytes_per_step": d2h, "hashed_host_bytes_per_step": h2d, "calls": c_calls,
                      "what": "pt_render on the SAME host arrays again: the scene is found in the library's device-scene cache"}

    # ---- roofline of the dominant kernel.  Two instrumented frames (same image, one full wavefront step of the production
    # size per rank): one with traversal counters, one with CUDA events around every launch (the counters' atomics would
    # distort the timing).  Algorithmic bytes per ray, SURVEY.md 8d restated for this build's records: 96 B per wide BVH
    # node visited + 64 B per triangle scanned + 64 B per exact test, + the path state the wavefront kernels stream.
    hbm_peak, peak_src, sm_max = peaks()
    roof = None
    if rank == 0:
        cspp = max(world, min(spp, 64 * world))
        ccfg = ptb200.config(W, H, cspp, depth, 1e-4, error, seed=42, sample_begin=rank, sample_step=world, device=local,
                             accel=args.accel, mode=args.mode, q_shift=args.q_shift, rr_depth=rr)
        dsc = lib.scene_create(sc, ccfg, counters=True)
        dsc.render_device(ccfg, d_sum.data_ptr(), 0, d_cnt_i.data_ptr(), stream)
        torch.cuda.synchronize()
        cs = dsc.stats()
        dsc.close()
        dst = lib.scene_create(sc, ccfg, timed=True)
        for _ in range(2):  # the first frame also pays for the workspace allocation
            dst.render_device(ccfg, d_sum.data_ptr(), 0, d_cnt_i.data_ptr(), stream)
            torch.cuda.synchronize()
        ts = dst.stats()
        dst.close()
        segs_c = max(1, cs["segments_traced"])
        n_large = cs["n_large"]
        queued = max(1, cs["rays_queued"])
        nodes_q = cs["nodes_visited"] / queued                       # per ray that walks the tree
        scans_q = max(0.0, cs["tris_tested"] - n_large * segs_c) / queued
        exact_q = cs["exact_tests"] / queued
        paths_c = W * H * cspp // world
        names = ["k_wf_begin", "k_wf_traverse", "k_wf_bounce", "k_wf_commit"]
        class_ms, class_n = ts["class_ms"], ts["class_launches"]
        state = 96.0   # oT, dR, Tn read + written per shaded segment
        # what a launch of each class has to move if every byte is fetched once (HBM-side lower bound of its traffic)
        algo_bytes = [paths_c * 48.0,
                      queued * (96.0 * nodes_q + 64.0 * scans_q + 64.0 * exact_q + 44.0),
                      segs_c * state + queued * 4.0,
                      paths_c * 16.0 + (W * H // world) * 64.0]
        flops = [paths_c * (79.0 * n_large + 40.0),
                 queued * (54.0 * 4 * nodes_q + 40.0 * scans_q + 79.0 * exact_q),  # 8 slab tests per wide node = 4 x the BVH2 node of 8d
                 segs_c * (79.0 * n_large + 40.0), 0.0]
        fp32_peak = 148 * 128 * 2 * sm_max * 1e6 / 1e12
        try:
            tr = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(args.workload, {})
        except Exception:
            tr = {}
        per_kernel = {}
        for i, nm in enumerate(names):
            if class_n[i] == 0 or class_ms[i] <= 0:
                continue
            sec = class_ms[i] * 1e-3
            e = {"ms_per_step": class_ms[i], "launches": class_n[i], "launch_ms": class_ms[i] / class_n[i],
                 "algorithmic_gbs": algo_bytes[i] / sec / 1e9, "frac_of_hbm": algo_bytes[i] / sec / 1e9 / hbm_peak,
                 "fp32_tflops": flops[i] / sec / 1e12, "frac_of_fp32": flops[i] / sec / 1e12 / fp32_peak}
            m = tr.get(nm)
            if m and m.get("dram_bytes_per_launch") and m.get("paths_per_launch"):  # measured by ncu on a launch of this shape (profiles/traffic.json says which capture)
                scale = paths_c / max(1.0, m["paths_per_launch"])
                e.update({"dram_bytes_per_launch": m["dram_bytes_per_launch"] * scale,
                          "dram_gbs": m["dram_bytes_per_launch"] * scale / (class_ms[i] / class_n[i] * 1e-3) / 1e9,
                          "issue_active_pct": m.get("issue_active_pct"), "active_lanes_per_instruction": m.get("lanes"),
                          "l2_hit_pct": m.get("l2_hit_pct"), "l1_hit_pct": m.get("l1_hit_pct"), "ncu_source": m.get("source")})
                e["dram_frac_of_hbm"] = e["dram_gbs"] / hbm_peak
            per_kernel[nm] = e
        if not per_kernel:  # the megakernel variant (--mode 1) has no kernel classes: one kernel is the frame
            st_m = ds.stats()
            sec = st_m["kernel_ms"] * 1e-3
            tot = float(sum(algo_bytes)) * (spp / cspp)
            per_kernel["k_render_mega"] = {"ms_per_step": st_m["kernel_ms"], "launches": 1, "launch_ms": st_m["kernel_ms"],
                                           "algorithmic_gbs": tot / sec / 1e9, "frac_of_hbm": tot / sec / 1e9 / hbm_peak,
                                           "fp32_tflops": sum(flops) * (spp / cspp) / sec / 1e12,
                                           "frac_of_fp32": sum(flops) * (spp / cspp) / sec / 1e12 / fp32_peak}
            names, class_ms, class_n, algo_bytes = ["k_render_mega"], [st_m["kernel_ms"]], [1], [tot]
        k = max(range(len(names)), key=lambda i: class_ms[i])
        top = per_kernel[names[k]]
        roof = {"kernel": names[k], "bound": "hbm", "achieved": top["algorithmic_gbs"], "peak": hbm_peak, "unit": "GB/s",
                "frac": top["frac_of_hbm"], "traffic": top.get("dram_bytes_per_launch"), "peak_source": peak_src,
                "launch_ms": top["launch_ms"], "launches_timed": top["launches"],
                "bytes_per_launch": algo_bytes[k] / max(1, class_n[k]), "share_of_step": class_ms[k] / max(1e-9, sum(class_ms)),
                "binding": "instruction issue and memory latency under divergence, not bandwidth: see issue_active_pct / "
                           "active_lanes_per_instruction (ncu) and frac_of_fp32 beside the HBM fraction",
                "algorithmic_bytes": "per ray that walks the BVH: 96 B x wide nodes visited + 64 B x triangles scanned + 64 B x exact "
                                     "tests + 44 B of queue / path state; per shaded segment: 96 B of path state (the shared-memory "
                                     "triangles cost no HBM bytes)",
                "per_ray_walking_the_bvh": {"wide_nodes": nodes_q, "scan_tests": scans_q, "exact_tests": exact_q,
                                            "share_of_segments": queued / segs_c},
                "fp32_peak_tflops": fp32_peak, "per_kernel": per_kernel, "class_ms": dict(zip(names, class_ms))}
        try:  # the L2 read peak is not in MEASURED_PEAKS.json; measured here, 64 MB footprint
            l2_peak = lib.probe_read_bandwidth(64 << 20, 40)
            hbm_read = lib.probe_read_bandwidth(4 << 30, 2)
            roof["l2"] = {"peak": l2_peak, "unit": "GB/s", "frac": roof["achieved"] / l2_peak, "hbm_read_same_probe": hbm_read,
                          "how": "pt_probe_read_bandwidth: 64 MiB streamed 40 times with 256-bit loads (L2 resident); "
                                 "the same kernel over 4 GiB gives the HBM read figure"}
        except ptb200.PtError as e:
            roof["l2"] = {"error": str(e)}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cpu = cpu_reference_leg(args.workload, sc, W, H, spp, depth, error, rr_depth=RR_DEPTH.get(args.workload, 0))
        cpu = {k: cpu[k] for k in ("value", "unit", "cores", "kind", "sample", "release_threads", "host_cpus")}

    # ---- N GPUs inside ONE pt_render call (PtConfig.n_gpus = N: one host thread per device, NCCL reduce inside the
    # library), timed by rank 0 once the other ranks are done with their devices
    e2e_inproc = None
    inproc_hung = False
    if world > 1:
        barrier()
        dist.destroy_process_group()  # every rank together; rank 0 then has the devices to itself
        if rank == 0:
            time.sleep(2.0)

            def inproc_leg(res):
                try:
                    icfg = ptb200.config(W, H, spp, depth, 1e-4, error, seed=42, device=0, n_gpus=world, accel=args.accel,
                                         mode=args.mode, q_shift=args.q_shift, rr_depth=rr)
                    isc = {k: (pin(v) if isinstance(v, np.ndarray) else v) for k, v in sc.items()}
                    irgb = torch.zeros((H, W, 3), dtype=torch.float32).pin_memory().numpy()
                    lib.render(isc, icfg, out=irgb)                       # communicators, workspaces, scene cache
                    runs = {}
                    for kind in ("fresh", "cached"):
                        t0 = time.perf_counter()
                        for _ in range(e2e_steps):
                            if kind == "fresh":
                                nudge(isc)
                            lib.render(isc, icfg, out=irgb)
                        runs[kind] = (time.perf_counter() - t0) / e2e_steps
                    res["out"] = {"api": f"pt_render with PtConfig.n_gpus = {world} (one process, NCCL reduce inside the library)",
                                  "value": segments_per_step / runs["fresh"] / 1e6, "unit": "Mrays/s", "frame_ms": runs["fresh"] * 1e3,
                                  "cached": {"value": segments_per_step / runs["cached"] / 1e6, "frame_ms": runs["cached"] * 1e3},
                                  "h2d_bytes_per_step": h2d * world, "d2h_bytes_per_step": d2h}
                except Exception as e:  # noqa: BLE001 — the line of the torchrun leg must be printed whatever happens here
                    res["out"] = {"error": f"{type(e).__name__}: {e}"}

            # this leg is an extra: if it hangs (a collective that never completes) the line is printed without it
            res = {}
            th = threading.Thread(target=inproc_leg, args=(res,), daemon=True)
            th.start()
            th.join(INPROC_TIMEOUT_S)
            inproc_hung = th.is_alive()
            e2e_inproc = {"error": f"no answer within {INPROC_TIMEOUT_S} s"} if inproc_hung else res.get("out")

    if rank == 0:
        st = ds.stats()
        out = {"metric": "Mrays/s", "value": mrays, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps,
               "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong",
               "vs_baseline": None, "dtype": "f32", "data": "synthetic",
               "config": {"workload": desc, "width": W, "height": H, "spp": spp, "max_depth": depth, "triangles": int(sc["verts"].shape[0]),
                          "accel": {1: "brute-force (shared memory)", 2: "GPU LBVH"}[st["accel"]], "bvh_nodes": st["n_nodes"],
                          "large_tris_in_smem": st["n_large"], "quad_scan_units": st["n_quads"], "intersector": "reference-exact (triangles.h:48-73)",
                          "l2": "flushed between iterations (256 MiB fill inside the timed region)",
                          "parallelism": f"spp split over {world} GPU(s), one NCCL reduce" if world > 1 else "1 GPU",
                          "segments_per_frame": segments_per_step},
               "frame_ms": ms_per_step, "frame_ms_e2e": e2e_sec * 1e3,
               "e2e": {"value": e2e_mrays, "unit": "Mrays/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                       "api": "pt_render(const PtScene*, const PtConfig*, float* rgb)" if world == 1 else "pt_scene_create + pt_scene_render_device + NCCL reduce + pt_resolve_device",
                       "calls": e2e_calls,
                       "what": "every timed call gets a scene the library has not seen: hash + upload + BVH build + render + resolve + download"},
               "e2e_cached": e2e_cached, "e2e_inproc": e2e_inproc,
               "gpu_launches": launches_per_step * args.steps, "clocks": clocks, "roofline": roof, "cpu_baseline": cpu}
        emit(out)
        if world > 1 and inproc_hung:
            sys.stdout.flush()
            os._exit(0)  # the hung leg holds device handles: no orderly teardown
    ds.close()


if __name__ == "__main__":
    main()
