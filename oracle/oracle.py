"""ctypes binding of the CPU oracle (oracle/liboracle.so).

TEST INFRASTRUCTURE ONLY — imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs; never by the product package.  See oracle/rt_oracle.hpp for what it restates.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

PREC_F32, PREC_MIXED, PREC_F64 = 0, 1, 2
KIND_NAMES = ["Unhittable", "Sphere", "MovingSphere", "RectXY", "RectXZ", "RectYZ", "Cuboid", "BVHNode", "Translate",
              "Rotate", "ConstantMedium"]
STAT_NAMES = ["paths", "bounces", "box_tests", "sphere_tests", "moving_tests", "rect_tests", "cuboid_tests",
              "translate_visits", "rotate_visits", "medium_visits", "sphere_hits", "noise_calls", "texels", "queries",
              "visits", "unhittable_visits", "scatter_lambertian", "scatter_metal", "scatter_dielectric", "scatter_isotropic",
              "emitted", "light_samples", "light_pdf_leaves", "checker_evals"]


def build(force=False):
    so = os.path.join(_HERE, "liboracle.so")
    srcs = [os.path.join(_HERE, f) for f in ("oracle_capi.cpp", "rt_oracle.hpp", "rt_scenes.hpp", "../ray-tracing_b200/csrc/rt_strict_math.h")]
    if force or not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
        subprocess.check_call(["make", "-C", _HERE, "liboracle.so"], stdout=subprocess.DEVNULL)
    return so


def lib():
    global _LIB
    if _LIB is None:
        L = C.CDLL(build())
        L.orc_last_error.restype = C.c_char_p
        L.orc_scene_create.restype = C.c_void_p
        L.orc_scene_create.argtypes = [C.c_char_p, C.c_int64, C.c_double, C.c_double, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int]
        L.orc_scene_destroy.argtypes = [C.c_void_p]
        L.orc_scene_counts.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_scene_roots.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_camera.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_void_p]
        L.orc_export.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        L.orc_trace.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_double, C.c_double, C.c_uint64, C.c_int] + [C.c_void_p] * 8
        L.orc_render.argtypes = [C.c_void_p, C.c_void_p] + [C.c_int] * 7 + [C.c_uint64] + [C.c_int] * 5 + [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_uint32]
        L.orc_texture_value.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        L.orc_strict_math_eval.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        L.orc_splitmix.argtypes = [C.c_int64, C.c_int, C.c_void_p, C.c_void_p]
        L.orc_stream.argtypes = [C.c_uint64, C.c_uint32, C.c_uint32, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_albedo_to_color.argtypes = [C.c_int, C.c_void_p, C.c_void_p]
        _LIB = L
    return _LIB


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


# per-kind export layouts: what -> (doubles per object, uint32s per object)
_EXPORT = {1: (4, 1), 2: (9, 1), 3: (5, 2), 6: (6, 1), 7: (6, 3), 8: (3, 1), 9: (2, 2), 10: (1, 2), 11: (1, 2), 12: (4, 3), 13: (768, 768)}


class OracleScene:
    def __init__(self, name, seed=1024, t0=0.0, t1=1.0, earth=None, lights_mode=0, synth_n=1000):
        L = lib()
        self.name, self.synth_n = name, synth_n
        if earth is not None:
            earth = np.ascontiguousarray(earth, dtype=np.uint8)
            eh, ew = earth.shape[:2]
            ep = _p(earth)
        else:
            eh = ew = 0
            ep = None
        self.h = L.orc_scene_create(name.encode(), seed, t0, t1, ep, ew, eh, lights_mode, synth_n)
        if not self.h:
            raise RuntimeError(L.orc_last_error().decode())

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_scene_destroy(self.h)
            self.h = None

    def counts(self):
        c = np.zeros(15, np.uint32)
        lib().orc_scene_counts(self.h, _p(c))
        return c

    def roots(self):
        r = np.zeros(2, np.uint32); bg = np.zeros(3); g1 = np.zeros(2, np.uint64)
        lib().orc_scene_roots(self.h, _p(r), _p(bg), _p(g1))
        return r, bg, g1

    def camera(self, W, H):
        c = np.zeros(24)
        lib().orc_camera(self.name.encode(), W, H, self.synth_n, _p(c))
        return c

    def export(self, what):
        cnt = self.counts()
        n = int(cnt[what]) if what <= 10 else int(cnt[what])
        nd, nu = _EXPORT[what]
        d = np.zeros((max(n, 1), nd)); u = np.zeros((max(n, 1), nu), np.uint32)
        got = lib().orc_export(self.h, what, _p(d), _p(u))
        assert got == n, (what, got, n)
        return d[:n], u[:n]

    def trace(self, rays7, prec=PREC_F64, tmin=1e-4, tmax=np.inf, seed=0, log_stride=0, strict=False):
        rays7 = np.ascontiguousarray(rays7, dtype=np.float64)
        n = rays7.shape[0]
        out = dict(hit=np.zeros(n, np.int32), rec=np.zeros((n, 9)), prim=np.zeros(n, np.uint32), mat=np.zeros(n, np.uint32),
                   front=np.zeros(n, np.uint8), nvisit=np.zeros(n, np.uint64), vhash=np.zeros(n, np.uint64),
                   vlog=np.zeros((n, max(log_stride, 1)), np.uint32))
        rc = lib().orc_trace(self.h, prec, 1 if strict else 0, n, _p(rays7), tmin, tmax, seed, log_stride, _p(out["hit"]), _p(out["rec"]), _p(out["prim"]),
                             _p(out["mat"]), _p(out["front"]), _p(out["nvisit"]), _p(out["vhash"]), _p(out["vlog"]))
        if rc:
            raise RuntimeError(lib().orc_last_error().decode())
        return out

    def render(self, W, H, spp, depth=50, prec=PREC_F64, mode=1, seed=1, s0=0, region=None, threads=None, stats=False, cam=None,
               strict=False, order="reference", pathlog=0):
        """Returns (sum[H,W,3] float64 radiance sums, stats dict or None) — or (sum, prim_log, t_log) with pathlog = max bounces.
        strict: the shared strict-math functions (= the device under RT_FLAG_STRICT_MATH); order "beta": throughput accumulated
        forward like the device (mode 1 only); both default to the reference's literal form."""
        cam = self.camera(W, H) if cam is None else np.ascontiguousarray(cam, dtype=np.float64)
        r0, r1, c0, c1 = region if region else (0, H, 0, W)
        threads = threads or os.cpu_count() or 1
        out = np.zeros((H, W, 3)); st = np.zeros(24, np.uint64)
        flags = (1 if strict else 0) | (2 if order == "beta" else 0)
        if (flags or pathlog) and mode != 1:
            raise ValueError("strict / beta / pathlog need mode=1")
        prim = t = None
        if pathlog:
            paths = (r1 - r0) * (c1 - c0) * spp
            prim = np.full((paths, pathlog), 0xFFFFFFFE, np.uint32); t = np.zeros((paths, pathlog))
        rc = lib().orc_render(self.h, _p(cam), prec, W, H, s0, spp, depth, mode, seed, r0, r1, c0, c1, threads, _p(out), _p(st) if stats else None,
                              flags, _p(prim) if pathlog else None, _p(t) if pathlog else None, pathlog)
        if rc:
            raise RuntimeError(lib().orc_last_error().decode())
        if pathlog:
            return out, prim, t
        return out, (dict(zip(STAT_NAMES, (int(x) for x in st))) if stats else None)

    def texture_value(self, tex_id, q5, prec=PREC_F64, strict=False):
        q5 = np.ascontiguousarray(q5, dtype=np.float64)
        out = np.zeros((q5.shape[0], 3))
        rc = lib().orc_texture_value(self.h, tex_id, prec, 1 if strict else 0, q5.shape[0], _p(q5), _p(out))
        if rc:
            raise RuntimeError(lib().orc_last_error().decode())
        return out


def strict_math_eval(fn, x, y=None, as_float=False):
    x = np.ascontiguousarray(x, dtype=np.float64)
    y = np.zeros_like(x) if y is None else np.ascontiguousarray(y, dtype=np.float64)
    out = np.zeros_like(x)
    lib().orc_strict_math_eval(["sin", "cos", "atan2", "asin", "log"].index(fn), x.size, _p(x), _p(y), _p(out), int(as_float))
    return out


def splitmix(seed, n):
    w = np.zeros(n, np.uint64); d = np.zeros(n)
    lib().orc_splitmix(seed, n, _p(w), _p(d))
    return w, d


def stream(seed, pixel, sample, n):
    st = np.zeros(2, np.uint64); d = np.zeros(n); f = np.zeros(n, np.float32)
    lib().orc_stream(seed, pixel, sample, n, _p(st), _p(d), _p(f))
    return st, d, f


def albedo_to_color(mean):
    mean = np.ascontiguousarray(mean, dtype=np.float64)
    out = np.zeros(mean.shape, np.uint8)
    lib().orc_albedo_to_color(mean.size, _p(mean), _p(out))
    return out


def primary_rays(cam24, W, H, time=0.0):
    """Pixel-centre pinhole rays (lens offset 0) in row-major, top row first: the fixed ray set for K1 parity."""
    c = np.asarray(cam24, dtype=np.float64)
    o, llc, hz, vt = c[0:3], c[3:6], c[6:9], c[9:12]
    ys = (H - 1 - np.arange(H))[:, None]; xs = np.arange(W)[None, :]
    u = ((xs + 0.5) / W)[..., None]; v = ((ys + 0.5) / H)[..., None]
    d = (llc + u * hz + v * vt) - o
    rays = np.zeros((H, W, 7)); rays[..., 0:3] = o; rays[..., 3:6] = d; rays[..., 6] = time
    return rays.reshape(-1, 7)
