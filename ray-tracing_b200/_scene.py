"""HostScene (the mirrored Scenes.hs constructors + flattener) and DeviceScene (a scene handle behind the C ABI)."""
import ctypes as C

import numpy as np

from . import (PREC_DEFAULT, REDUCE_KINDS, RtRenderParams, RtSceneDesc, RtStats, COUNTER_NAMES, _check, _p, core, host, load_earth)


def _arr(ptr, n, dtype):
    if not ptr or n == 0:
        return np.zeros(0, dtype)
    ct = {np.float64: C.c_double, np.uint32: C.c_uint32, np.int32: C.c_int32}[dtype]
    return np.ctypeslib.as_array(C.cast(ptr, C.POINTER(ct)), shape=(n,)).copy()


def desc_arrays(d):
    """Copies the arrays of an RtSceneDesc into numpy (dict keyed by the field names of include/rt_abi.h)."""
    f, u = np.float64, np.uint32
    out = {
        "node_box": _arr(d.node_box, d.n_nodes * 6, f).reshape(-1, 6), "node_left": _arr(d.node_left, d.n_nodes, u),
        "node_right": _arr(d.node_right, d.n_nodes, u), "node_size": _arr(d.node_size, d.n_nodes, u),
        "sphere_cr": _arr(d.sphere_cr, d.n_spheres * 4, f).reshape(-1, 4), "sphere_mat": _arr(d.sphere_mat, d.n_spheres, u),
        "moving": _arr(d.moving, d.n_moving * 9, f).reshape(-1, 9), "moving_mat": _arr(d.moving_mat, d.n_moving, u),
        "rect": _arr(d.rect, d.n_rects * 5, f).reshape(-1, 5), "rect_mat": _arr(d.rect_mat, d.n_rects, u),
        "cuboid": _arr(d.cuboid, d.n_cuboids * 6, f).reshape(-1, 6), "cuboid_mat": _arr(d.cuboid_mat, d.n_cuboids, u),
        "translate_off": _arr(d.translate_off, d.n_translates * 3, f).reshape(-1, 3), "translate_child": _arr(d.translate_child, d.n_translates, u),
        "rotate_sc": _arr(d.rotate_sc, d.n_rotates * 2, f).reshape(-1, 2), "rotate_axis": _arr(d.rotate_axis, d.n_rotates, u),
        "rotate_child": _arr(d.rotate_child, d.n_rotates, u),
        "medium_neg_inv_density": _arr(d.medium_neg_inv_density, d.n_media, f), "medium_mat": _arr(d.medium_mat, d.n_media, u),
        "medium_boundary": _arr(d.medium_boundary, d.n_media, u),
        "mat_kind": _arr(d.mat_kind, d.n_materials, u), "mat_tex": _arr(d.mat_tex, d.n_materials, u), "mat_param": _arr(d.mat_param, d.n_materials, f),
        "tex_kind": _arr(d.tex_kind, d.n_textures, u), "tex_rgb": _arr(d.tex_rgb, d.n_textures * 3, f).reshape(-1, 3),
        "tex_a": _arr(d.tex_a, d.n_textures, u), "tex_b": _arr(d.tex_b, d.n_textures, u), "tex_scale": _arr(d.tex_scale, d.n_textures, f),
        "perlin_ranvec": _arr(d.perlin_ranvec, d.n_perlin * 768, f).reshape(-1, 768),
        "perlin_perm": _arr(d.perlin_perm, d.n_perlin * 768, np.int32).reshape(-1, 768),
        "image_w": _arr(d.image_w, d.n_images, u), "image_h": _arr(d.image_h, d.n_images, u),
        "world_root": int(d.world_root), "lights_root": int(d.lights_root),
        "background": np.array(list(d.background)), "camera": np.array(list(d.camera)),
    }
    return out


class HostScene:
    """A reference scene built with the mirrored constructors (librt_host.so): name, seed (randGen), shutter, lit flag."""

    def __init__(self, name, seed=1024, t0=0.0, t1=1.0, earth="auto", lit=False, synth_n=1000):
        L = host()
        self.name = name
        if isinstance(earth, str):
            earth = load_earth() if name in ("earth", "random", "next_week_final") else None
        if earth is not None:
            earth = np.ascontiguousarray(earth, dtype=np.uint8)
            eh, ew = earth.shape[:2]
            ep = _p(earth)
        else:
            eh = ew = 0
            ep = None
        self.h = L.rth_scene_create(name.encode(), seed, t0, t1, ep, ew, eh, 1 if lit else 0, synth_n)
        if not self.h:
            raise RuntimeError(L.rth_last_error().decode())

    def __del__(self):
        if getattr(self, "h", None):
            host().rth_scene_destroy(self.h)
            self.h = None

    def desc(self, W, H):
        """Pointer to the flattened RtSceneDesc (valid until the next desc() call on this scene)."""
        d = host().rth_scene_desc(self.h, W, H)
        if not d:
            raise RuntimeError(host().rth_last_error().decode())
        return d

    def g1(self):
        g = np.zeros(2, np.uint64)
        host().rth_scene_g1(self.h, _p(g))
        return g

    def run_render(self, W, H, num_samples, max_depth=50):
        """runRender through the mirrored API -> uint8 [H, W, 3] (albedoToColor applied, rows top first)."""
        out = np.zeros((H, W, 3), np.uint8)
        if host().rth_run_render(self.h, W, H, num_samples, max_depth, _p(out)) != 0:
            raise RuntimeError(host().rth_last_error().decode())
        return out

    def upload(self, W, H, device=-1):
        return DeviceScene(self.desc(W, H), device)


class DeviceScene:
    """A scene handle behind rt_scene_create; every method is one C-ABI call."""

    def __init__(self, desc_ptr, device=-1):
        sid = C.c_uint64(0)
        _check(core().rt_scene_create(desc_ptr, device, C.byref(sid)))
        self.id = sid

    def destroy(self):
        if getattr(self, "id", None) is not None:
            core().rt_scene_destroy(self.id)
            self.id = None

    __del__ = destroy

    @staticmethod
    def params(W, H, spp, depth=50, seed=1, s0=0, region=None, precision=PREC_DEFAULT, flags=0, total=None, n_gpus=0):
        r0, r1, c0, c1 = region if region else (0, 0, 0, 0)
        return RtRenderParams(W, H, s0, spp, total or spp, depth, r0, r1, c0, c1, seed, precision, flags, -1, n_gpus)

    def render(self, W, H, spp, counted=False, out=None, **kw):
        """-> (radiance sums float64 [H,W,3], stats dict)"""
        p = self.params(W, H, spp, **kw)
        out = np.zeros((H, W, 3)) if out is None else out
        st = RtStats()
        fn = core().rt_render_counted if counted else core().rt_render
        _check(fn(self.id, C.byref(p), _p(out), C.byref(st)))
        stats = dict(kernel_ms=st.kernel_ms, d2h_ms=st.d2h_ms, samples=int(st.samples), launches=int(st.launches), gpus=int(st.gpus),
                     reduce_ms=st.reduce_ms, replicate_ms=st.replicate_ms, tail_ms=st.tail_ms, kernel=["renderKernel", "renderKernelCoop", "wavefront"][st.kernel_kind], rounds=int(st.rounds), reduce_kind=REDUCE_KINDS[st.reduce_kind],
                     gpu_kernel_ms=[float(x) for x in st.gpu_kernel_ms[:max(int(st.gpus), 1)]])
        if counted:
            stats["counters"] = dict(zip(COUNTER_NAMES, (int(x) for x in st.counters)))
            stats["sched"] = [int(x) for x in st.sched]
        return out, stats

    def render_device(self, d_ptr, W, H, spp, stream=0, timed=False, **kw):
        p = self.params(W, H, spp, **kw)
        st = RtStats()
        _check(core().rt_render_device(self.id, C.byref(p), C.c_void_p(d_ptr), C.c_void_p(stream), C.byref(st) if timed else None))
        return dict(kernel_ms=st.kernel_ms, launches=int(st.launches), rounds=int(st.rounds), kernel=["renderKernel", "renderKernelCoop", "wavefront"][st.kernel_kind]) if timed else None

    def render_device_static(self, d_ptr, W, H, spp, stream=0, **kw):
        """rt_render_device_static: no allocation, scratch in the library's static device array (include/rt_abi.h); asynchronous on `stream`"""
        p = self.params(W, H, spp, **kw)
        _check(core().rt_render_device_static(self.id, C.byref(p), C.c_void_p(d_ptr), C.c_void_p(stream)))

    def render_pathlog(self, W, H, spp, max_bounces=50, **kw):
        """-> (radiance sums, prim_log uint32 [paths, max_bounces] (0xfffffffe = bounce not reached, 0xffffffff = miss), t_log float64)"""
        p = self.params(W, H, spp, **kw)
        r0, r1, c0, c1 = kw.get("region") or (0, H, 0, W)
        paths = (r1 - r0) * (c1 - c0) * spp
        out = np.zeros((H, W, 3))
        prim = np.full((paths, max_bounces), 0xFFFFFFFE, np.uint32); t = np.zeros((paths, max_bounces))
        _check(core().rt_render_pathlog(self.id, C.byref(p), _p(out), max_bounces, _p(prim), _p(t)))
        return out, prim, t

    def trace(self, rays7, precision=PREC_DEFAULT, flags=0, tmin=1e-4, tmax=np.inf, seed=0, log_stride=0):
        rays7 = np.ascontiguousarray(rays7, dtype=np.float64)
        n = rays7.shape[0]
        o = dict(hit=np.zeros(n, np.int32), rec=np.zeros((n, 9)), prim=np.zeros(n, np.uint32), mat=np.zeros(n, np.uint32),
                 front=np.zeros(n, np.uint8), nvisit=np.zeros(n, np.uint64), vhash=np.zeros(n, np.uint64),
                 vlog=np.zeros((n, max(log_stride, 1)), np.uint32))
        _check(core().rt_trace(self.id, precision, flags, n, _p(rays7), tmin, tmax, seed, log_stride, _p(o["hit"]), _p(o["rec"]),
                               _p(o["prim"]), _p(o["mat"]), _p(o["front"]), _p(o["nvisit"]), _p(o["vhash"]), _p(o["vlog"])))
        return o

    def texture_value(self, tex, q5, precision=PREC_DEFAULT, flags=0):
        q5 = np.ascontiguousarray(q5, dtype=np.float64)
        out = np.zeros((q5.shape[0], 3))
        _check(core().rt_texture_value(self.id, tex, precision, flags, q5.shape[0], _p(q5), _p(out)))
        return out


def strict_math_eval(fn, x, y=None, as_float=False):
    """rt_strict_math_eval: the shared strict-math functions evaluated on the device (fn: sin cos atan2 asin log)."""
    x = np.ascontiguousarray(x, dtype=np.float64)
    y = np.zeros_like(x) if y is None else np.ascontiguousarray(y, dtype=np.float64)
    out = np.zeros_like(x)
    _check(core().rt_strict_math_eval(["sin", "cos", "atan2", "asin", "log"].index(fn), x.size, _p(x), _p(y), _p(out), int(as_float)))
    return out


def tonemap(rgb_sum, total_samples):
    rgb_sum = np.ascontiguousarray(rgb_sum, dtype=np.float64)
    out = np.zeros(rgb_sum.shape, np.uint8)
    _check(core().rt_tonemap(_p(rgb_sum), rgb_sum.size, total_samples, _p(out)))
    return out


def write_image(path, rgb8, binary=False):
    """The image file app/Main.hs:58-62 prints for a rendered frame (uint8 [H, W, 3], top row first): "P3" text rows in
    printRow's format (Lib.hs:299-305), or "P6" binary with the same header."""
    rgb8 = np.ascontiguousarray(rgb8, dtype=np.uint8)
    H, W, _ = rgb8.shape
    if host().rth_write_image(str(path).encode(), _p(rgb8), W, H, int(bool(binary))) != 0:
        raise RuntimeError(host().rth_last_error().decode())


def bvh_build(leaf_box, node_axis, device=-1):
    """rt_bvh_build: the reference's makeBVH on the device.  leaf_box [n, 6] float64, node_axis [N(n)] uint8 in pre-order.
    Returns dict(box [N, 6], left [N], right [N], size [N], ms); children with bit 31 set are pre-order node ids."""
    leaf_box = np.ascontiguousarray(leaf_box, dtype=np.float64)
    node_axis = np.ascontiguousarray(node_axis, dtype=np.uint8)
    n = leaf_box.shape[0]
    N = core().rt_bvh_node_count(n)
    if node_axis.shape[0] != N:
        raise ValueError("node_axis must have rt_bvh_node_count(n) = %d entries" % N)
    box = np.zeros((N, 6)); left = np.zeros(N, np.uint32); right = np.zeros(N, np.uint32); size = np.zeros(N, np.uint32)
    ms = C.c_double(0.0)
    _check(core().rt_bvh_build(n, _p(leaf_box), _p(node_axis), N, _p(box), _p(left), _p(right), _p(size), device, C.byref(ms)))
    return dict(box=box, left=left, right=right, size=size, ms=ms.value)
