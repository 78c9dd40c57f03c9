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
    print(json.dumps(out))


def main():
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

    # ---- end to end through the drop-in entry point: HOST scene -> pt_render -> HOST rgb (upload, LBVH build, render,
    # resolve, download all inside the timed region).  One call per step; with N ranks every rank renders its sample
    # partition through pt_render_accum and the accumulators are reduced from pinned host memory.
    e2e_steps = max(1, min(args.steps, 3))
    h2d = int(sc["verts"].nbytes + sc["mat_ids"].nbytes + sc["mats"].nbytes +
              (sc["planes"].nbytes + sc["squares"].nbytes if sc.get("planes") is not None else 0))
    d2h = n_px * 3 * 4
    def e2e_call():
        """HOST scene arrays -> HOST rgb through the C-ABI.  One GPU: pt_render.  N ranks: what a multi-process host does
        with the device-resident entry points — pt_scene_create (upload + LBVH build), pt_scene_render_device on this
        rank's sample partition, one NCCL reduce of (sum, count), pt_resolve_device, download on rank 0."""
        if world == 1:
            return lib.render(sc_pinned, cfg, out=rgb_pinned)
        h = lib.scene_create(sc, cfg)
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
    if world == 1:
        pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory().numpy()
        sc_pinned = {k: (pin(v) if isinstance(v, np.ndarray) else v) for k, v in sc.items()}
        rgb_pinned = torch.zeros((H, W, 3), dtype=torch.float32).pin_memory().numpy()
    if args.warmup > 0:
        e2e_call()  # one untimed call: the library keeps its path-state workspace (GBs of HBM) between calls
    barrier()
    t0 = time.perf_counter()
    e2e_calls = []
    for _ in range(e2e_steps):
        tc = time.perf_counter()
        rgb = e2e_call()
        e2e_calls.append({"ms": round((time.perf_counter() - tc) * 1e3, 1), "phases_ms": lib.last_phases() if world == 1 else None})
    barrier()
    e2e_s = torch.tensor([(time.perf_counter() - t0) / e2e_steps], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    e2e_mrays = segments_per_step / float(e2e_s.item()) / 1e6

    # ---- roofline of the dominant kernel (k_render): algorithmic bytes per ray = 64 B per BVH node visited + 64 B per
    # triangle tested, counted by an instrumented render of the same frame at reduced spp
    hbm_peak, peak_src, sm_max = peaks()
    roof = None
    if rank == 0:
        cspp = max(world, min(spp, 64 * world))  # one full wavefront step of the production size per rank
        ccfg = ptb200.config(W, H, cspp, depth, 1e-4, error, seed=42, sample_begin=rank, sample_step=world, device=local,
                             accel=args.accel, mode=args.mode, q_shift=args.q_shift, rr_depth=rr)
        # two instrumented frames (same image, cspp samples): one with traversal counters, one with CUDA events around
        # every launch (the counters' atomics would distort the timing)
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
        cs["class_ms"], cs["class_launches"] = ts["class_ms"], ts["class_launches"]
        segs_c = max(1, cs["segments_traced"])
        n_large = cs["n_large"]
        nodes_ray = cs["nodes_visited"] / segs_c
        tris_ray = cs["tris_tested"] / segs_c
        leaf_tris_ray = max(0.0, tris_ray - n_large)
        bytes_ray = 64.0 * (nodes_ray + tris_ray)
        flops_ray = 54.0 * nodes_ray + 79.0 * tris_ray + 40.0
        st = ds.stats()
        k_ms, rays_kernel = st["kernel_ms"], st["segments_traced"]
        fp32_peak = 148 * 128 * 2 * sm_max * 1e6 / 1e12
        pipeline_gbs = rays_kernel * bytes_ray / (k_ms * 1e-3) / 1e9
        roof = {"bound": "hbm", "peak": hbm_peak, "unit": "GB/s", "peak_source": peak_src, "traffic": None,
                "bytes_per_ray": bytes_ray, "nodes_per_ray": nodes_ray, "tris_per_ray": tris_ray,
                "rays_reaching_lbvh": cs["rays_queued"] / segs_c,
                "pipeline": {"achieved": pipeline_gbs, "frac": pipeline_gbs / hbm_peak, "ms": k_ms,
                             "what": "all kernels of one frame: rays x 64 B x (nodes visited + triangles tested)"},
                "note": "algorithmic bytes (SURVEY 8d: 64 B per node visited / triangle tested, + path state for the "
                        "wavefront kernels) are mostly served by shared memory and L2, so frac against the HBM peak can "
                        "exceed 1; fp32 gives the compute-side ceiling",
                "fp32": {"achieved_tflops": rays_kernel * flops_ray / (k_ms * 1e-3) / 1e12, "peak_tflops": fp32_peak,
                         "frac": rays_kernel * flops_ray / (k_ms * 1e-3) / 1e12 / fp32_peak, "flops_per_ray": flops_ray}}
        if args.mode == 0 and sum(cs["class_launches"]) > 0:
            # dominant kernel class of the wavefront pipeline, timed live with CUDA events around every launch of the
            # instrumented frame (same image, cspp samples)
            names = ["k_wf_begin", "k_wf_traverse", "k_wf_bounce", "k_wf_commit"]
            paths_c = W * H * cspp // world
            state = 96.0  # oT,dR,Tn read + written per live segment
            per_class_bytes = [paths_c * (64.0 * n_large + 48.0),
                               segs_c * 64.0 * (nodes_ray + leaf_tris_ray),
                               (segs_c - paths_c) * 64.0 * n_large + segs_c * state,
                               paths_c * 16.0 + (W * H // world) * 64.0]
            k = max(range(4), key=lambda i: cs["class_ms"][i])
            ach = per_class_bytes[k] / (cs["class_ms"][k] * 1e-3) / 1e9
            roof.update({"kernel": names[k], "achieved": ach, "frac": ach / hbm_peak,
                         "launch_ms": cs["class_ms"][k] / max(1, cs["class_launches"][k]),
                         "launches_timed": cs["class_launches"][k],
                         "bytes_per_launch": per_class_bytes[k] / max(1, cs["class_launches"][k]),
                         "share_of_step": cs["class_ms"][k] / max(1e-9, sum(cs["class_ms"])),
                         "class_ms": dict(zip(names, cs["class_ms"]))})
            try:  # measured DRAM traffic of that kernel (one ncu --set full capture), scaled to the timed launch
                tr = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
                per_slot = tr.get(args.workload, {}).get(names[k], {}).get("dram_bytes_per_slot")
                if per_slot:
                    roof["traffic"] = per_slot * paths_c / max(1, cs["class_launches"][0])
                    roof["traffic_source"] = tr[args.workload][names[k]]["source"]
            except Exception:
                pass
        else:
            roof.update({"kernel": "k_render_mega", "achieved": pipeline_gbs, "frac": pipeline_gbs / hbm_peak})
        try:  # the L2 read peak is not in MEASURED_PEAKS.json; measured here, 64 MB footprint
            l2_peak = lib.probe_read_bandwidth(64 << 20, 40)
            hbm_read = lib.probe_read_bandwidth(4 << 30, 2)
            roof["l2"] = {"peak": l2_peak, "unit": "GB/s", "frac": roof["achieved"] / l2_peak,
                          "hbm_read_same_probe": hbm_read,
                          "how": "pt_probe_read_bandwidth: 64 MiB streamed 40 times with 256-bit loads (L2 resident); "
                                 "the same kernel over 4 GiB gives the HBM read figure"}
        except ptb200.PtError as e:
            roof["l2"] = {"error": str(e)}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cpu = cpu_reference_leg(args.workload, sc, W, H, spp, depth, error, rr_depth=RR_DEPTH.get(args.workload, 0))
        cpu = {k: cpu[k] for k in ("value", "unit", "cores", "kind", "sample", "release_threads", "host_cpus")}

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
               "frame_ms": ms_per_step, "frame_ms_e2e": float(e2e_s.item()) * 1e3,
               "e2e": {"value": e2e_mrays, "unit": "Mrays/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                       "api": "pt_render(const PtScene*, const PtConfig*, float* rgb)" if world == 1 else "pt_scene_create + pt_scene_render_device + NCCL reduce + pt_resolve_device",
                       "calls": e2e_calls},
               "gpu_launches": launches_per_step * args.steps, "clocks": clocks, "roofline": roof, "cpu_baseline": cpu}
        print(json.dumps(out))
    ds.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
