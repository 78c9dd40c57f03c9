"""ctypes binding of the B200 render library (include/pt_render.h) and of the host-side C++ loader.

This module is plumbing only: it loads lib/libptrender.so (CUDA kernels + C-ABI) and lib/libpthost.so (the C++
Scene/Config mirror of the reference's scene.cpp / config.h) and exposes their entry points on numpy arrays.  It
contains no rendering arithmetic and no CPU fallback: if the CUDA library is missing, import fails loudly.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
RENDER_SO = os.environ.get("PTB200_RENDER_SO") or os.path.join(HERE, "lib", "libptrender.so")
HOST_SO = os.path.join(HERE, "lib", "libpthost.so")

PT_ACCEL_AUTO, PT_ACCEL_BRUTE, PT_ACCEL_LBVH = 0, 1, 2
PT_ISECT_REFERENCE, PT_ISECT_WATERTIGHT = 0, 1
ERRORS = {0: "PT_OK", 1: "PT_EINVAL", 2: "PT_ENODEVICE", 3: "PT_ECUDA", 4: "PT_ENCCL", 5: "PT_ENOMEM"}

# every symbol include/pt_render.h declares (tests check that the library exports all of them)
ABI_SYMBOLS = ["pt_render", "pt_render_accum", "pt_scene_create", "pt_scene_destroy", "pt_scene_render_device",
               "pt_resolve_device", "pt_scene_stats", "pt_intersect", "pt_scene_intersect", "pt_triangle_setup",
               "pt_philox", "pt_sincos", "pt_primary_rays", "pt_last_error", "pt_device_count", "pt_version",
               "pt_render_accum_w", "pt_scene_render_device_w", "pt_resolve_device_w", "pt_release_workspace", "pt_postprocess", "pt_postprocess_device", "pt_render_image", "pt_powf", "pt_last_phases", "pt_probe_read_bandwidth", "pt_render_progressive", "pt_debug_checks",
               "pt_scene_render_device_static"]


class PtScene(C.Structure):
    _fields_ = [("n_triangles", C.c_int32), ("vertices", C.c_void_p), ("planes", C.c_void_p),
                ("squares", C.c_void_p), ("material_ids", C.c_void_p), ("n_materials", C.c_int32),
                ("materials", C.c_void_p), ("sky_width", C.c_int32), ("sky_height", C.c_int32),
                ("sky_bgr", C.c_void_p), ("materials_ext", C.c_void_p)]


class PtConfig(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("rays_per_pixel", C.c_int32),
                ("max_ray_reflections", C.c_int32), ("eps", C.c_float), ("error", C.c_float), ("seed", C.c_uint32),
                ("sample_begin", C.c_int32), ("sample_step", C.c_int32), ("intersector", C.c_int32),
                ("accel", C.c_int32), ("rr_depth", C.c_int32), ("n_gpus", C.c_int32), ("device", C.c_int32),
                ("reserved", C.c_int32 * 4), ("row_begin", C.c_int32), ("row_end", C.c_int32)]


class PtStats(C.Structure):
    _fields_ = [("segments", C.c_uint64), ("segments_traced", C.c_uint64), ("nodes_visited", C.c_uint64),
                ("tris_tested", C.c_uint64), ("kernel_ms", C.c_float), ("build_ms", C.c_float),
                ("upload_ms", C.c_float), ("n_nodes", C.c_int32), ("n_large", C.c_int32), ("accel", C.c_int32),
                ("launches", C.c_int32), ("class_ms", C.c_float * 4), ("class_launches", C.c_int32 * 4),
                ("rays_queued", C.c_uint64), ("node_hist", C.c_uint64 * 16), ("n_quads", C.c_int32), ("n_levels", C.c_int32),
                ("exact_tests", C.c_uint64)]


class PtPostConfig(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("gamma_correction", C.c_float), ("gauss", C.c_int32),
                ("median", C.c_int32), ("reserved", C.c_int32 * 3)]


class PtDispersion(C.Structure):
    _fields_ = [("max_dispersion", C.c_float), ("min_dispersion", C.c_float), ("average_dispersion", C.c_float)]


class PtError(RuntimeError):
    def __init__(self, code, text):
        super().__init__(f"{ERRORS.get(code, code)}: {text}")
        self.code = code


def config(width=512, height=512, spp=20, max_depth=8, eps=1e-4, error=0.001, seed=42, sample_begin=0,
           sample_step=1, accel=PT_ACCEL_AUTO, intersector=PT_ISECT_REFERENCE, device=0, n_gpus=1, rr_depth=0,
           mode=0, q_shift=0, resume=0, alive=0, rows=None):
    """PtConfig with the reference's defaults (config.h:10,16-29)."""
    c = PtConfig()
    c.width, c.height, c.rays_per_pixel, c.max_ray_reflections = width, height, spp, max_depth
    c.eps, c.error, c.seed = eps, error, seed
    c.sample_begin, c.sample_step = sample_begin, sample_step
    c.intersector, c.accel, c.rr_depth, c.n_gpus, c.device = intersector, accel, rr_depth, n_gpus, device
    c.reserved[0] = mode  # 0 = wavefront pipeline, 1 = megakernel variant
    c.reserved[1] = q_shift  # 0 = automatic, else log2(samples per pixel per wavefront step)
    c.reserved[3] = alive  # alive lists of the wavefront pipeline: 0 automatic, 1 on, 2 off
    c.reserved[2] = resume  # 1 = continue from the accumulators passed in (samples before sample_begin are in them)
    if rows is not None:    # image rows [begin, end) traced by this call
        c.row_begin, c.row_end = rows
    return c


PREVIEW_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_void_p, C.c_int32)


def _point_at_bundled_nccl():
    """PtConfig.n_gpus > 1 loads NCCL at run time (dlopen, $PT_NCCL_LIB first).  A Python process that also imports
    PyTorch should end up with ONE libnccl.so.2: point the library at the wheel's copy (nvidia/nccl/lib) when there is
    one and the caller has not chosen; otherwise the system library is used."""
    if os.environ.get("PT_NCCL_LIB"):
        return
    try:
        import importlib.util
        spec = importlib.util.find_spec("nvidia.nccl")
        for root in (spec.submodule_search_locations if spec else []):
            cand = os.path.join(root, "lib", "libnccl.so.2")
            if os.path.exists(cand):
                os.environ["PT_NCCL_LIB"] = cand
                return
    except Exception:
        pass


def _load(path):
    if not os.path.exists(path):
        raise ImportError(f"{path} is missing: build it with `make -C path-tracing_b200` (nvcc, sm_100a). "
                          "There is no CPU fallback.")
    return C.CDLL(path)


class Library:
    """The C-ABI of include/pt_render.h on numpy arrays.  A scene is a dict of arrays:
    verts (n,9) f32, planes (n,4) f32 | None, squares (n,) f32 | None, mat_ids (n,) i32, mats (m,10) f32."""

    def __init__(self):
        _point_at_bundled_nccl()
        L = self.lib = _load(RENDER_SO)
        L.pt_last_error.restype = C.c_char_p
        L.pt_version.restype = C.c_char_p
        L.pt_scene_create.argtypes = [C.POINTER(PtScene), C.POINTER(PtConfig), C.c_int32, C.POINTER(C.c_void_p)]
        L.pt_scene_destroy.argtypes = [C.c_void_p]
        L.pt_scene_destroy.restype = None
        L.pt_scene_render_device.argtypes = [C.c_void_p, C.POINTER(PtConfig), C.c_void_p, C.c_void_p, C.c_void_p,
                                             C.c_void_p]
        L.pt_resolve_device.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]
        L.pt_scene_render_device_w.argtypes = [C.c_void_p, C.POINTER(PtConfig), C.c_void_p, C.c_void_p, C.c_void_p,
                                               C.c_void_p, C.c_void_p]
        L.pt_resolve_device_w.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]
        L.pt_scene_render_device_static.argtypes = [C.c_void_p, C.POINTER(PtConfig), C.c_void_p, C.c_void_p, C.c_void_p,
                                                    C.c_void_p, C.c_void_p]
        L.pt_render_accum_w.argtypes = [C.POINTER(PtScene), C.POINTER(PtConfig), C.c_void_p, C.c_void_p, C.c_void_p,
                                        C.c_void_p]
        L.pt_scene_stats.argtypes = [C.c_void_p, C.POINTER(PtStats)]
        L.pt_render.argtypes = [C.POINTER(PtScene), C.POINTER(PtConfig), C.c_void_p]
        L.pt_render_accum.argtypes = [C.POINTER(PtScene), C.POINTER(PtConfig), C.c_void_p, C.c_void_p, C.c_void_p]
        L.pt_intersect.argtypes = [C.POINTER(PtScene), C.POINTER(PtConfig), C.c_void_p, C.c_int64, C.c_void_p,
                                   C.c_void_p]
        L.pt_scene_intersect.argtypes = [C.c_void_p, C.POINTER(PtConfig), C.c_void_p, C.c_int64, C.c_void_p,
                                         C.c_void_p]
        L.pt_triangle_setup.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]
        L.pt_philox.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]
        L.pt_sincos.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]
        L.pt_primary_rays.argtypes = [C.POINTER(PtConfig), C.c_void_p, C.c_int64, C.c_void_p]
        L.pt_postprocess.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(PtPostConfig), C.c_void_p,
                                     C.POINTER(PtDispersion)]
        L.pt_postprocess_device.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(PtPostConfig),
                                            C.c_void_p, C.POINTER(PtDispersion), C.c_void_p]
        L.pt_render_image.argtypes = [C.POINTER(PtScene), C.POINTER(PtConfig), C.POINTER(PtPostConfig), C.c_void_p,
                                      C.POINTER(PtDispersion)]
        L.pt_powf.argtypes = [C.c_void_p, C.c_int64, C.c_float, C.c_void_p]
        L.pt_probe_read_bandwidth.argtypes = [C.c_size_t, C.c_int, C.POINTER(C.c_double)]
        L.pt_release_workspace.restype = None

    def last_phases(self):
        ms = (C.c_float * 5)()
        self.lib.pt_last_phases(ms)
        return dict(zip(("scene_create", "frame_buffers", "render_issue", "resolve_d2h", "release"), [round(float(v), 2) for v in ms]))

    def check(self, rc):
        if rc != 0:
            raise PtError(rc, self.lib.pt_last_error().decode())

    def debug_checks(self):
        """violation counters of the checked build: dict name -> count (reading resets them)"""
        out = (C.c_uint64 * 8)()
        self.check(self.lib.pt_debug_checks(out))
        return dict(zip(("queue", "alive", "stack", "node", "record", "slot", "pixel", "checks_run"), [int(v) for v in out]))

    def device_count(self):
        return int(self.lib.pt_device_count())

    @staticmethod
    def scene_struct(sc):
        keep = {"verts": np.ascontiguousarray(sc["verts"], np.float32).reshape(-1, 9),
                "mat_ids": np.ascontiguousarray(sc["mat_ids"], np.int32),
                "mats": np.ascontiguousarray(sc["mats"], np.float32).reshape(-1, 10)}
        s = PtScene()
        s.n_triangles = keep["verts"].shape[0]
        s.vertices = keep["verts"].ctypes.data
        if sc.get("planes") is not None:
            keep["planes"] = np.ascontiguousarray(sc["planes"], np.float32).reshape(-1, 4)
            keep["squares"] = np.ascontiguousarray(sc["squares"], np.float32)
            s.planes, s.squares = keep["planes"].ctypes.data, keep["squares"].ctypes.data
        s.material_ids = keep["mat_ids"].ctypes.data
        s.n_materials = keep["mats"].shape[0]
        s.materials = keep["mats"].ctypes.data
        if sc.get("sky") is not None:  # (H, W, 3) uint8, blue green red, rows top-down (bitmap_image's layout)
            keep["sky"] = np.ascontiguousarray(sc["sky"], np.uint8)
            s.sky_height, s.sky_width = keep["sky"].shape[:2]
            s.sky_bgr = keep["sky"].ctypes.data
        if sc.get("mats_ext") is not None:  # extension: (m, 2) f32 = Ni, Tr per material
            keep["mats_ext"] = np.ascontiguousarray(sc["mats_ext"], np.float32).reshape(-1, 2)
            assert keep["mats_ext"].shape[0] == keep["mats"].shape[0]
            s.materials_ext = keep["mats_ext"].ctypes.data
        s._keep = keep
        return s

    # ---- host-buffer entry points (the drop-in boundary) ----
    def render(self, sc, cfg, out=None):
        """out: a caller-owned (H, W, 3) float32 array to fill (e.g. page-locked memory reused across frames)"""
        rgb = np.zeros((cfg.height, cfg.width, 3), np.float32) if out is None else out
        assert rgb.dtype == np.float32 and rgb.flags.c_contiguous and rgb.size == cfg.height * cfg.width * 3
        s = self.scene_struct(sc)
        self.check(self.lib.pt_render(C.byref(s), C.byref(cfg), rgb.ctypes.data))
        return rgb

    def render_accum(self, sc, cfg, state=None):
        """state: (sum, sum2, count) of an earlier call to continue from (cfg built with resume=1)"""
        H, W = cfg.height, cfg.width
        sm = np.zeros((H, W, 3), np.float32)
        s2 = np.zeros((H, W, 3), np.float32)
        cnt = np.zeros((H, W), np.int32)
        if state is not None:
            sm, s2, cnt = (np.ascontiguousarray(a).copy() for a in state[:3])
        s = self.scene_struct(sc)
        self.check(self.lib.pt_render_accum(C.byref(s), C.byref(cfg), sm.ctypes.data, s2.ctypes.data, cnt.ctypes.data))
        return sm, s2, cnt

    def render_accum_w(self, sc, cfg, state=None):
        """-> sum, sum2, count, weight (the divisor under Russian roulette; == count otherwise)"""
        H, W = cfg.height, cfg.width
        sm = np.zeros((H, W, 3), np.float32)
        s2 = np.zeros((H, W, 3), np.float32)
        cnt = np.zeros((H, W), np.int32)
        wt = np.zeros((H, W), np.float32)
        if state is not None:
            sm, s2, cnt, wt = (np.ascontiguousarray(a).copy() for a in state)
        s = self.scene_struct(sc)
        self.check(self.lib.pt_render_accum_w(C.byref(s), C.byref(cfg), sm.ctypes.data, s2.ctypes.data, cnt.ctypes.data,
                                              wt.ctypes.data))
        return sm, s2, cnt, wt

    def intersect(self, sc, rays6, cfg=None):
        cfg = cfg or config()
        rays6 = np.ascontiguousarray(rays6, np.float32).reshape(-1, 6)
        n = rays6.shape[0]
        ids = np.empty(n, np.int32)
        t = np.empty(n, np.float32)
        s = self.scene_struct(sc)
        self.check(self.lib.pt_intersect(C.byref(s), C.byref(cfg), rays6.ctypes.data, n, ids.ctypes.data, t.ctypes.data))
        return ids, t

    def triangle_setup(self, verts, vn=None, has_vn=None):
        v = np.ascontiguousarray(verts, np.float32).reshape(-1, 9)
        n = v.shape[0]
        planes = np.zeros((n, 4), np.float32)
        squares = np.zeros(n, np.float32)
        hp = vp = None
        if vn is not None:
            vn = np.ascontiguousarray(vn, np.float32).reshape(-1, 3)
            has_vn = np.ascontiguousarray(has_vn if has_vn is not None else np.ones(n), np.uint8)
            hp, vp = has_vn.ctypes.data, vn.ctypes.data
        self.check(self.lib.pt_triangle_setup(v.ctypes.data, hp, vp, n, planes.ctypes.data, squares.ctypes.data))
        return planes, squares

    def philox(self, ctr4, key2):
        ctr4 = np.ascontiguousarray(ctr4, np.uint32).reshape(-1, 4)
        key2 = np.ascontiguousarray(key2, np.uint32).reshape(-1, 2)
        out = np.zeros_like(ctr4)
        self.check(self.lib.pt_philox(ctr4.ctypes.data, key2.ctypes.data, ctr4.shape[0], out.ctypes.data))
        return out

    def sincos(self, x):
        x = np.ascontiguousarray(x, np.float32)
        s = np.empty_like(x)
        c = np.empty_like(x)
        self.check(self.lib.pt_sincos(x.ctypes.data, x.size, s.ctypes.data, c.ctypes.data))
        return s, c

    def powf(self, x, y):
        x = np.ascontiguousarray(x, np.float32)
        out = np.empty_like(x)
        self.check(self.lib.pt_powf(x.ctypes.data, x.size, y, out.ctypes.data))
        return out

    def probe_read_bandwidth(self, footprint_bytes, passes=20):
        """GB/s of streaming a buffer of that size `passes` times (L2 bandwidth when it fits in L2)"""
        g = C.c_double(0.0)
        self.check(self.lib.pt_probe_read_bandwidth(footprint_bytes, passes, C.byref(g)))
        return g.value

    # ---- after the loop (main.cpp:162-201) ----
    @staticmethod
    def post_config(width, height, gamma=1 / 2.2, gauss=0, median=0):
        return PtPostConfig(width, height, gamma, gauss, median)

    def postprocess(self, sm, s2, cnt, post, weight=None):
        """HOST accumulators -> (rgb8 (H, W, 3) uint8, (max, min, average) dispersion)"""
        sm, s2 = np.ascontiguousarray(sm, np.float32), np.ascontiguousarray(s2, np.float32)
        cnt = np.ascontiguousarray(cnt, np.int32)
        w = None if weight is None else np.ascontiguousarray(weight, np.float32)
        img = np.zeros((post.height, post.width, 3), np.uint8)
        st = PtDispersion()
        self.check(self.lib.pt_postprocess(sm.ctypes.data, s2.ctypes.data, cnt.ctypes.data, None if w is None else w.ctypes.data,
                                           C.byref(post), img.ctypes.data, C.byref(st)))
        return img, (st.max_dispersion, st.min_dispersion, st.average_dispersion)

    def render_image(self, sc, cfg, post):
        img = np.zeros((post.height, post.width, 3), np.uint8)
        st = PtDispersion()
        s = self.scene_struct(sc)
        self.check(self.lib.pt_render_image(C.byref(s), C.byref(cfg), C.byref(post), img.ctypes.data, C.byref(st)))
        return img, (st.max_dispersion, st.min_dispersion, st.average_dispersion)

    def render_progressive(self, sc, cfg, post, update=32, time_limit_ms=0, on_update=None):
        """pt_render_progressive; on_update(image HxWx3 u8, samples) -> truthy stops the loop.
        -> image, (max, min, average dispersion), samples done"""
        H, W = post.height, post.width
        img = np.zeros((H, W, 3), np.uint8)
        st = PtDispersion()
        done = C.c_int32(0)
        s = self.scene_struct(sc)

        def tramp(user, ptr, samples):
            view = np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_uint8)), shape=(H, W, 3))
            return 1 if on_update(view.copy(), samples) else 0

        cb = PREVIEW_FN(tramp) if on_update else C.cast(None, PREVIEW_FN)
        self.lib.pt_render_progressive.argtypes = [C.POINTER(PtScene), C.POINTER(PtConfig), C.POINTER(PtPostConfig),
                                                   C.c_int, C.c_int64, PREVIEW_FN, C.c_void_p, C.c_void_p,
                                                   C.POINTER(PtDispersion), C.POINTER(C.c_int32)]
        self.check(self.lib.pt_render_progressive(C.byref(s), C.byref(cfg), C.byref(post), update, time_limit_ms, cb,
                                                  None, img.ctypes.data, C.byref(st), C.byref(done)))
        return img, (st.max_dispersion, st.min_dispersion, st.average_dispersion), done.value

    def primary_rays(self, cfg, xys3):
        xys3 = np.ascontiguousarray(xys3, np.int32).reshape(-1, 3)
        rays = np.zeros((xys3.shape[0], 6), np.float32)
        self.check(self.lib.pt_primary_rays(C.byref(cfg), xys3.ctypes.data, xys3.shape[0], rays.ctypes.data))
        return rays

    # ---- device-resident API ----
    def scene_create(self, sc, cfg, counters=False, timed=False):
        return DeviceScene(self, sc, cfg, counters, timed)


class DeviceScene:
    """Handle on a device-resident scene (SoA triangles + LBVH)."""

    def __init__(self, lib, sc, cfg, counters=False, timed=False):
        self.L = lib
        self.h = C.c_void_p()
        s = Library.scene_struct(sc)
        lib.check(lib.lib.pt_scene_create(C.byref(s), C.byref(cfg), (1 if counters else 0) | (2 if timed else 0),
                                          C.byref(self.h)))

    def render_device(self, cfg, d_sum3=0, d_sum2=0, d_count=0, stream=0, d_weight=0):
        """Device pointers (ints, e.g. torch.Tensor.data_ptr()); asynchronous on `stream`."""
        self.L.check(self.L.lib.pt_scene_render_device_w(self.h, C.byref(cfg), d_sum3 or None, d_sum2 or None,
                                                         d_count or None, d_weight or None, stream or None))

    def render_device_static(self, cfg, d_sum3=0, d_sum2=0, d_count=0, stream=0, d_weight=0):
        """pt_scene_render_device_static: the same render out of the library's static device workspace (small frames,
        flat spp; allocates nothing).  Calls must not overlap on the GPU."""
        self.L.check(self.L.lib.pt_scene_render_device_static(self.h, C.byref(cfg), d_sum3 or None, d_sum2 or None,
                                                              d_count or None, d_weight or None, stream or None))

    def resolve_device(self, d_sum3, d_count, n_pixels, d_rgb, stream=0, d_weight=0):
        self.L.check(self.L.lib.pt_resolve_device_w(d_sum3, d_count, d_weight or None, n_pixels, d_rgb, stream or None))

    def intersect(self, rays6, cfg=None):
        cfg = cfg or config()
        rays6 = np.ascontiguousarray(rays6, np.float32).reshape(-1, 6)
        n = rays6.shape[0]
        ids = np.empty(n, np.int32)
        t = np.empty(n, np.float32)
        self.L.check(self.L.lib.pt_scene_intersect(self.h, C.byref(cfg), rays6.ctypes.data, n, ids.ctypes.data,
                                                   t.ctypes.data))
        return ids, t

    def stats(self):
        st = PtStats()
        self.L.check(self.L.lib.pt_scene_stats(self.h, C.byref(st)))
        out = {k: getattr(st, k) for k, _ in PtStats._fields_}
        out["class_ms"] = list(st.class_ms)
        out["class_launches"] = list(st.class_launches)
        out["node_hist"] = list(st.node_hist)
        return out

    def close(self):
        if self.h:
            self.L.lib.pt_scene_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class HostLoader:
    """The C++ host side (host/scene.cpp: Scene::LoadModel with the reference's OBJ/MTL semantics)."""

    def __init__(self):
        _load(RENDER_SO)
        L = self.lib = _load(HOST_SO)
        L.pth_load_obj.argtypes = [C.c_char_p, C.c_char_p]
        L.pth_copy_scene.argtypes = [C.c_void_p] * 5
        L.pth_copy_materials_ext.argtypes = [C.c_void_p]
        L.pth_triangle_setup.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        L.pth_triangle_setup_n.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]
        L.pth_intersect.argtypes = [C.c_void_p, C.c_int64, C.POINTER(PtConfig), C.c_void_p, C.c_void_p]
        L.pth_ray_reflect.argtypes = [C.c_void_p] * 7
        L.pth_parse_config.argtypes = [C.c_int, C.POINTER(C.c_char_p), C.POINTER(PtConfig)]
        L.pth_set_cache_dir.argtypes = [C.c_char_p]

    def load_obj(self, model_path, model_name, cache_dir=""):
        """cache_dir: the host side's -CACHE <dir> (binary image of the parsed model); sc["from_cache"] tells whether
        this load took it"""
        if not model_path.endswith("/"):
            model_path += "/"
        self.lib.pth_set_cache_dir(cache_dir.encode())
        n = self.lib.pth_load_obj(model_path.encode(), model_name.encode())
        self.lib.pth_set_cache_dir(b"")
        if n < 0:
            raise FileNotFoundError(model_path + model_name)
        m = self.lib.pth_material_count()
        sc = {"verts": np.zeros((n, 9), np.float32), "planes": np.zeros((n, 4), np.float32),
              "squares": np.zeros(n, np.float32), "mat_ids": np.zeros(n, np.int32),
              "mats": np.zeros((m, 10), np.float32)}
        self.lib.pth_copy_scene(sc["verts"].ctypes.data, sc["planes"].ctypes.data, sc["squares"].ctypes.data,
                                sc["mat_ids"].ctypes.data, sc["mats"].ctypes.data)
        ext = np.zeros((m, 2), np.float32)  # Ni, Tr as read from the MTL (extension; d -> Tr = 1 - d)
        self.lib.pth_copy_materials_ext(ext.ctypes.data)
        sc["mats_ext"] = ext if self.lib.pth_has_refraction() else None
        sc["mats_ext_raw"] = ext
        sc["from_cache"] = bool(self.lib.pth_loaded_from_cache())
        sc["parse_pieces"] = int(self.lib.pth_parsed_in_pieces())  # threads the OBJ was parsed with (0 = sequential)
        return sc

    def check_fast_strtof(self, text):
        """The loader's fast float conversion against strtof on every token of `text`: (tokens, fast, mismatches)."""
        counts = (C.c_int64 * 3)()
        self.lib.pth_check_fast_strtof(C.c_char_p(text.encode()), counts)
        return tuple(counts)

    def triangle_setup(self, verts, vn=None):
        v = np.ascontiguousarray(verts, np.float32).reshape(-1, 9)
        n = v.shape[0]
        planes = np.zeros((n, 4), np.float32)
        squares = np.zeros(n, np.float32)
        sq = C.c_float()
        zero = np.zeros(3, np.float32)
        for i in range(n):
            has = vn is not None
            nn = np.ascontiguousarray(vn[i], np.float32) if has else zero
            self.lib.pth_triangle_setup(v[i].ctypes.data, int(has), nn.ctypes.data, planes[i].ctypes.data, C.byref(sq))
            squares[i] = sq.value
        return planes, squares

    def triangle_setup_bulk(self, verts, vn=None, has_vn=None):
        """Triangle's constructor (+ SetNormal where has_vn) for n triangles in one call"""
        v = np.ascontiguousarray(verts, np.float32).reshape(-1, 9)
        n = v.shape[0]
        planes = np.zeros((n, 4), np.float32)
        squares = np.zeros(n, np.float32)
        hp = vp = None
        if vn is not None:
            vn = np.ascontiguousarray(vn, np.float32).reshape(-1, 3)
            has_vn = np.ascontiguousarray(has_vn if has_vn is not None else np.ones(n), np.uint8)
            hp, vp = has_vn.ctypes.data, vn.ctypes.data
        self.lib.pth_triangle_setup_n(v.ctypes.data, hp, vp, n, planes.ctypes.data, squares.ctypes.data)
        return planes, squares

    def intersect_loaded(self, rays6, cfg=None):
        """Scene::Intersect(std::vector<Ray>) on the scene of the last load_obj (needs a GPU)"""
        cfg = cfg or config()
        rays6 = np.ascontiguousarray(rays6, np.float32).reshape(-1, 6)
        n = rays6.shape[0]
        ids = np.empty(n, np.int32)
        t = np.empty(n, np.float32)
        rc = self.lib.pth_intersect(rays6.ctypes.data, n, C.byref(cfg), ids.ctypes.data, t.ctypes.data)
        if rc:
            raise PtError(rc, "Scene::Intersect")
        return ids, t

    def ray_reflect(self, begin4, dir4, color3, depth, nbegin4, ndir4, tint3):
        """Ray::Reflect on a Ray built from (begin, direction, colour, depth): -> the Ray's state afterwards"""
        b, d, c = (np.array(a, np.float32) for a in (begin4, dir4, color3))
        nb, nd, tn = (np.ascontiguousarray(a, np.float32) for a in (nbegin4, ndir4, tint3))
        dep = C.c_int(depth)
        self.lib.pth_ray_reflect(b.ctypes.data, d.ctypes.data, c.ctypes.data, C.addressof(dep), nb.ctypes.data, nd.ctypes.data,
                                 tn.ctypes.data)
        return b, d, c, dep.value

    def parse_config(self, argv):
        arr = (C.c_char_p * len(argv))(*[a.encode() for a in argv])
        out = PtConfig()
        self.lib.pth_parse_config(len(argv), arr, C.byref(out))
        return out
