"""TEST INFRASTRUCTURE ONLY: ctypes access to oracle/libptoracle.so (plain-C restatement) and, when it
has been built, oracle/_ref/libptref.so (the unmodified reference + shims).  Imported by tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs only."""
import ctypes as C
import os
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(HERE, "libptoracle.so")
REF_SO = os.path.join(HERE, "_ref", "libptref.so")

_f32p = np.ctypeslib.ndpointer(dtype=np.float32, flags="C_CONTIGUOUS")
_i32p = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")
_u8p = np.ctypeslib.ndpointer(dtype=np.uint8, flags="C_CONTIGUOUS")


class OrScene(C.Structure):
    _fields_ = [("n_tri", C.c_int), ("verts9", C.c_void_p), ("planes4", C.c_void_p), ("squares", C.c_void_p),
                ("mat_ids", C.c_void_p), ("n_mat", C.c_int), ("mats10", C.c_void_p),
                ("sky_w", C.c_int), ("sky_h", C.c_int), ("sky_bgr", C.c_void_p), ("mats_ext2", C.c_void_p)]


class OrConfig(C.Structure):
    _fields_ = [("width", C.c_int), ("height", C.c_int), ("spp", C.c_int), ("max_depth", C.c_int),
                ("eps", C.c_float), ("error", C.c_float), ("seed", C.c_uint32),
                ("s_begin", C.c_int), ("s_step", C.c_int), ("rr_depth", C.c_int)]


def _load(path):
    if not os.path.exists(path):
        raise FileNotFoundError(f"{path} missing: run `make -C oracle`")
    return C.CDLL(path)


class Oracle:
    """The plain-C restatement.  A scene is a dict of float32/int32 arrays:
    verts (n,9), planes (n,4), squares (n,), mat_ids (n,), mats (m,10)."""

    def __init__(self):
        L = self.lib = _load(ORACLE_SO)
        L.or_render_accum.restype = C.c_ulonglong
        L.or_render_accum.argtypes = [C.POINTER(OrScene), C.POINTER(OrConfig), C.c_void_p, C.c_void_p, C.c_void_p,
                                      C.c_void_p, C.c_int, C.POINTER(C.c_double)]
        L.or_intersect.argtypes = [C.POINTER(OrScene), C.c_float, _f32p, C.c_long, _i32p, _f32p, C.c_int]
        L.or_triangle_setup.argtypes = [_f32p, C.c_int, _f32p, _f32p, C.POINTER(C.c_float)]
        L.or_primary_ray.argtypes = [C.POINTER(OrConfig), C.c_int, C.c_int, C.c_int, _f32p, _f32p]
        L.or_trace_path.restype = C.c_int
        L.or_trace_path.argtypes = [C.POINTER(OrScene), C.POINTER(OrConfig), C.c_int, C.c_int, C.c_int, C.c_int, _f32p]
        L.or_resolve.argtypes = [_f32p, _i32p, C.c_void_p, C.c_long, _f32p]
        L.or_libm_sincos.argtypes = [_f32p, C.c_long, _f32p, _f32p]
        L.or_max_threads.restype = C.c_int
        L.or_resolve_display.argtypes = [_f32p, _f32p, _i32p, C.c_void_p, C.c_int, C.c_int, C.c_float, _f32p, _u8p, _f32p]
        L.or_to_rgb8.argtypes = [_f32p, _i32p, C.c_long, _u8p]
        L.or_powf_check.restype = C.c_long
        L.or_powf_check.argtypes = [_f32p, C.c_long, C.c_float, C.POINTER(C.c_float)]
        L.or_libm_powf.argtypes = [_f32p, C.c_long, C.c_float, _f32p]
        L.or_refract.restype = C.c_int
        L.or_refract.argtypes = [_f32p, _f32p, C.c_float, _f32p]

    @staticmethod
    def scene_struct(sc):
        keep = {k: np.ascontiguousarray(sc[k], dtype=(np.int32 if k == "mat_ids" else np.float32))
                for k in ("verts", "planes", "squares", "mat_ids", "mats")}
        s = OrScene(int(keep["verts"].shape[0]), keep["verts"].ctypes.data, keep["planes"].ctypes.data,
                    keep["squares"].ctypes.data, keep["mat_ids"].ctypes.data, int(keep["mats"].shape[0]),
                    keep["mats"].ctypes.data)
        if sc.get("sky") is not None:  # (H, W, 3) uint8, blue green red, rows top-down
            keep["sky"] = np.ascontiguousarray(sc["sky"], np.uint8)
            s.sky_h, s.sky_w = keep["sky"].shape[:2]
            s.sky_bgr = keep["sky"].ctypes.data
        if sc.get("mats_ext") is not None:  # extension: (m, 2) = Ni, Tr
            keep["mats_ext"] = np.ascontiguousarray(sc["mats_ext"], np.float32)
            s.mats_ext2 = keep["mats_ext"].ctypes.data
        s._keep = keep
        return s

    @staticmethod
    def config(width, height, spp, max_depth=8, eps=1e-4, error=-1.0, seed=42, s_begin=0, s_step=1, rr_depth=0):
        return OrConfig(width, height, spp, max_depth, eps, error, seed, s_begin, s_step, rr_depth)

    def max_threads(self):
        return int(self.lib.or_max_threads())

    def triangle_setup(self, verts9, vn=None):
        v = np.ascontiguousarray(verts9, np.float32).reshape(-1, 9)
        n = v.shape[0]
        planes = np.zeros((n, 4), np.float32)
        squares = np.zeros(n, np.float32)
        sq = C.c_float()
        zero = np.zeros(3, np.float32)
        for i in range(n):
            has = vn is not None and vn[i] is not None
            self.lib.or_triangle_setup(v[i].copy(), int(has), np.ascontiguousarray(vn[i], np.float32) if has else zero,
                                       planes[i], C.byref(sq))
            squares[i] = sq.value
        return planes, squares

    def intersect(self, sc, rays6, eps=1e-4, threads=0):
        rays6 = np.ascontiguousarray(rays6, np.float32).reshape(-1, 6)
        n = rays6.shape[0]
        ids = np.empty(n, np.int32)
        t = np.empty(n, np.float32)
        s = self.scene_struct(sc)
        self.lib.or_intersect(C.byref(s), eps, rays6, n, ids, t, threads or self.max_threads())
        return ids, t

    def primary_ray(self, cfg, x, y, s):
        o = np.zeros(3, np.float32)
        d = np.zeros(3, np.float32)
        self.lib.or_primary_ray(C.byref(cfg), x, y, s, o, d)
        return o, d

    def render_accum(self, sc, cfg, threads=0):
        """-> (sum[H,W,3], sum2[H,W,3], count[H,W] int32, segments, seconds)"""
        H, W = cfg.height, cfg.width
        sm = np.zeros((H, W, 3), np.float32)
        s2 = np.zeros((H, W, 3), np.float32)
        cnt = np.zeros((H, W), np.int32)
        sec = C.c_double()
        s = self.scene_struct(sc)
        segs = self.lib.or_render_accum(C.byref(s), C.byref(cfg), sm.ctypes.data, s2.ctypes.data, cnt.ctypes.data,
                                        None, threads or self.max_threads(), C.byref(sec))
        return sm, s2, cnt, int(segs), sec.value

    def render_accum_w(self, sc, cfg, threads=0):
        """Extension outputs too: -> (sum, sum2, count, weight[H,W] f32, segments)"""
        H, W = cfg.height, cfg.width
        sm = np.zeros((H, W, 3), np.float32)
        s2 = np.zeros((H, W, 3), np.float32)
        cnt = np.zeros((H, W), np.int32)
        wt = np.zeros((H, W), np.float32)
        sec = C.c_double()
        s = self.scene_struct(sc)
        segs = self.lib.or_render_accum(C.byref(s), C.byref(cfg), sm.ctypes.data, s2.ctypes.data, cnt.ctypes.data,
                                        wt.ctypes.data, threads or self.max_threads(), C.byref(sec))
        return sm, s2, cnt, wt, int(segs)

    def trace_path(self, sc, cfg, x, y, s, max_seg=64):
        rec = np.zeros((max_seg, 9), np.float32)
        st = self.scene_struct(sc)
        n = self.lib.or_trace_path(C.byref(st), C.byref(cfg), x, y, s, max_seg, rec)
        return rec[:min(n, max_seg)], n

    def resolve(self, sm, cnt, weight=None):
        rgb = np.zeros_like(sm)
        w = None if weight is None else np.ascontiguousarray(weight, np.float32)
        self.lib.or_resolve(np.ascontiguousarray(sm), np.ascontiguousarray(cnt), None if w is None else w.ctypes.data,
                            cnt.size, rgb)
        return rgb

    def resolve_display(self, sm, s2, cnt, gamma=1 / 2.2, weight=None):
        """main.cpp:162-185,193-201 -> (float image (H,W,3), rgb8 (H,W,3), (max, min, average) dispersion)"""
        H, W = cnt.shape
        color = np.zeros((H, W, 3), np.float32)
        img = np.zeros((H, W, 3), np.uint8)
        st = np.zeros(3, np.float32)
        w = None if weight is None else np.ascontiguousarray(weight, np.float32)
        self.lib.or_resolve_display(np.ascontiguousarray(sm, np.float32), np.ascontiguousarray(s2, np.float32),
                                    np.ascontiguousarray(cnt, np.int32), None if w is None else w.ctypes.data, W, H, gamma,
                                    color, img, st)
        return color, img, tuple(float(v) for v in st)

    def to_rgb8(self, color, cnt):
        img = np.zeros(color.shape, np.uint8)
        self.lib.or_to_rgb8(np.ascontiguousarray(color, np.float32), np.ascontiguousarray(cnt, np.int32), cnt.size, img)
        return img

    def powf_check(self, x, y):
        """number of inputs where the oracle's restatement of glibc powf differs from the host libm"""
        x = np.ascontiguousarray(x, np.float32)
        fb = C.c_float()
        return int(self.lib.or_powf_check(x, x.size, y, C.byref(fb))), fb.value

    def libm_powf(self, x, y):
        x = np.ascontiguousarray(x, np.float32)
        out = np.empty_like(x)
        self.lib.or_libm_powf(x, x.size, y, out)
        return out

    def refract(self, d, n, eta):
        out = np.zeros(3, np.float32)
        tir = self.lib.or_refract(np.ascontiguousarray(d, np.float32), np.ascontiguousarray(n, np.float32), eta, out)
        return out, bool(tir)

    def libm_sincos(self, x):
        x = np.ascontiguousarray(x, np.float32)
        s = np.empty_like(x)
        c = np.empty_like(x)
        self.lib.or_libm_sincos(x, x.size, s, c)
        return s, c

    def philox(self, ctr, key):
        out = (C.c_uint32 * 4)()
        self.lib.or_philox((C.c_uint32 * 4)(*ctr), (C.c_uint32 * 2)(*key), out)
        return [int(v) for v in out]


class Reference:
    """oracle/_ref/libptref.so: the reference's own Scene/Triangle/Material/Ray code (global scene)."""

    def __init__(self):
        L = self.lib = _load(REF_SO)
        L.ref_set_config.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, C.c_char_p]
        L.ref_scene_load_obj.restype = C.c_int
        L.ref_scene_load_obj.argtypes = [C.c_char_p, C.c_char_p]
        L.ref_scene_from_arrays.restype = C.c_int
        L.ref_scene_from_arrays.argtypes = [C.c_int, _f32p, C.c_void_p, C.c_void_p, _i32p, C.c_int, _f32p]
        L.ref_scene_size.restype = C.c_int
        L.ref_scene_dump.argtypes = [_f32p, _f32p, _f32p]
        L.ref_intersect.argtypes = [_f32p, C.c_long, _i32p, _f32p, C.c_int]
        L.ref_primary_ray.argtypes = [C.c_uint, C.c_int, C.c_int, C.c_int, _f32p, _f32p]
        L.ref_render_accum.restype = C.c_ulonglong
        L.ref_render_accum.argtypes = [C.c_uint, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                       C.c_void_p, C.POINTER(C.c_double)]
        L.ref_trace_path.restype = C.c_int
        L.ref_trace_path.argtypes = [C.c_uint, C.c_int, C.c_int, C.c_int, C.c_int, _f32p]
        L.ref_max_threads.restype = C.c_int
        L.ref_gauss_blur.argtypes = [C.c_int, C.c_int, _f32p, C.c_float, _f32p]
        L.ref_median_filter.argtypes = [C.c_int, C.c_int, _f32p, C.c_int, _f32p]
        L.ref_refract.argtypes = [_f32p, _f32p, C.c_float, _f32p]
        self.cfg = None

    @staticmethod
    def available():
        return os.path.exists(REF_SO)

    def max_threads(self):
        return int(self.lib.ref_max_threads())

    def set_config(self, width, height, spp, max_depth=8, eps=1e-4, error=-1.0, skybox=""):
        self.cfg = (width, height, spp, max_depth)
        self.lib.ref_set_config(width, height, spp, max_depth, eps, error, skybox.encode())

    def load_obj(self, model_path, model_name):
        if not model_path.endswith("/"):
            model_path += "/"
        return self.lib.ref_scene_load_obj(model_path.encode(), model_name.encode())

    def from_arrays(self, verts, mat_ids, mats, vn=None, has_vn=None):
        verts = np.ascontiguousarray(verts, np.float32).reshape(-1, 9)
        mat_ids = np.ascontiguousarray(mat_ids, np.int32)
        mats = np.ascontiguousarray(mats, np.float32).reshape(-1, 10)
        hv = vp = None
        if vn is not None:
            vn = np.ascontiguousarray(vn, np.float32).reshape(-1, 3)
            has_vn = np.ascontiguousarray(has_vn if has_vn is not None else np.ones(len(vn)), np.uint8)
            hv, vp = has_vn.ctypes.data, vn.ctypes.data
        return self.lib.ref_scene_from_arrays(verts.shape[0], verts, hv, vp, mat_ids, mats.shape[0], mats)

    def dump(self):
        n = self.lib.ref_scene_size()
        planes = np.zeros((n, 4), np.float32)
        verts = np.zeros((n, 9), np.float32)
        squares = np.zeros(n, np.float32)
        self.lib.ref_scene_dump(planes, verts, squares)
        return planes, verts, squares

    def intersect(self, rays6, threads=0):
        rays6 = np.ascontiguousarray(rays6, np.float32).reshape(-1, 6)
        n = rays6.shape[0]
        ids = np.empty(n, np.int32)
        t = np.empty(n, np.float32)
        self.lib.ref_intersect(rays6, n, ids, t, threads or self.max_threads())
        return ids, t

    def primary_ray(self, seed, x, y, s):
        o = np.zeros(3, np.float32)
        d = np.zeros(3, np.float32)
        self.lib.ref_primary_ray(seed, x, y, s, o, d)
        return o, d

    # main.cpp's free functions; images are (H, W, 3) float32
    def gauss_blur(self, img, r):
        img = np.ascontiguousarray(img, np.float32)
        out = np.zeros_like(img)
        self.lib.ref_gauss_blur(img.shape[1], img.shape[0], img, float(r), out)
        return out

    def median_filter(self, img, window):
        img = np.ascontiguousarray(img, np.float32)
        out = np.zeros_like(img)
        self.lib.ref_median_filter(img.shape[1], img.shape[0], img, int(window), out)
        return out

    def refract(self, d, n, refractive):
        out = np.zeros(3, np.float32)
        self.lib.ref_refract(np.ascontiguousarray(d, np.float32), np.ascontiguousarray(n, np.float32), refractive, out)
        return out

    def render_accum(self, seed=42, s_begin=0, s_end=None, s_step=1, threads=0, accumulate=False):
        W, H, spp, _ = self.cfg
        s_end = spp if s_end is None else s_end
        sm = np.zeros((H, W, 3), np.float32)
        s2 = np.zeros((H, W, 3), np.float32)
        cnt = np.zeros((H, W), np.int32)
        sec = C.c_double()
        segs = self.lib.ref_render_accum(seed, s_begin, s_end, s_step, threads or self.max_threads(), int(accumulate),
                                         sm.ctypes.data, s2.ctypes.data, cnt.ctypes.data, C.byref(sec))
        return sm, s2, cnt, int(segs), sec.value

    def trace_path(self, seed, x, y, s, max_seg=64):
        rec = np.zeros((max_seg, 9), np.float32)
        n = self.lib.ref_trace_path(seed, x, y, s, max_seg, rec)
        return rec[:min(n, max_seg)], n
