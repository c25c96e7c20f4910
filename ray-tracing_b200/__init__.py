"""ray-tracing_b200 — Python host binding of the B200 path-tracing core.

Plumbing only: ctypes over
  librt_b200.so  (CUDA kernels + the C ABI of include/rt_abi.h) and
  librt_host.so  (C++ mirror of the reference's Lib/Scenes host API + flattener).
All rendering happens in the CUDA kernels; there is no CPU fallback here, and nothing in this package imports
oracle/.  Load as a module named `ray_tracing_b200` through rtb200_loader.py (the directory name has a hyphen).
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_HERE)

PREC_F32, PREC_MIXED, PREC_F64 = 0, 1, 2
PREC_DEFAULT = PREC_MIXED  # RT_PREC_DEFAULT: the production arithmetic (include/rt_abi.h)
FLAG_FAST_MATH = 1
FLAG_STRICT_MATH = 2
FLAG_COOP = 4
FLAG_NO_COOP = 8
FLAG_WAVEFRONT = 16
FLAG_NO_WAVEFRONT = 32
STATIC_MAX_PIXELS = 262144  # RT_STATIC_MAX_PIXELS (rt_render_device_static)
INIT_NCCL = 1
REDUCE_KINDS = ["none", "peer-kernel", "nccl", "staged"]
RT_NONE = 0xFFFFFFFF
KIND_NAMES = ["Unhittable", "Sphere", "MovingSphere", "RectXY", "RectXZ", "RectYZ", "Cuboid", "BVHNode", "Translate",
              "Rotate", "ConstantMedium"]
COUNTER_NAMES = ["paths", "bounces", "box_tests", "sphere_tests", "moving_tests", "rect_tests", "cuboid_tests",
                 "translate_visits", "rotate_visits", "medium_visits", "sphere_hits", "noise_calls", "texels", "queries",
                 "visits", "unhittable_visits", "scatter_lambertian", "scatter_metal", "scatter_dielectric", "scatter_isotropic",
                 "emitted", "light_samples", "light_pdf_leaves", "checker_evals"]
SCENES = ["cornell", "cornell_smoke", "simple_light", "earth", "two_perlin", "two_spheres", "random_book1", "random",
          "next_week_final", "synthetic", "quirks"]


class RtSceneDesc(C.Structure):
    _p, _u, _d = C.c_void_p, C.c_uint32, C.c_double
    _fields_ = [
        ("abi_version", _u),
        ("n_nodes", _u), ("node_box", _p), ("node_left", _p), ("node_right", _p), ("node_size", _p),
        ("n_spheres", _u), ("sphere_cr", _p), ("sphere_mat", _p),
        ("n_moving", _u), ("moving", _p), ("moving_mat", _p),
        ("n_rects", _u), ("rect", _p), ("rect_mat", _p),
        ("n_cuboids", _u), ("cuboid", _p), ("cuboid_mat", _p),
        ("n_translates", _u), ("translate_off", _p), ("translate_child", _p),
        ("n_rotates", _u), ("rotate_sc", _p), ("rotate_axis", _p), ("rotate_child", _p),
        ("n_media", _u), ("medium_neg_inv_density", _p), ("medium_mat", _p), ("medium_boundary", _p),
        ("n_materials", _u), ("mat_kind", _p), ("mat_tex", _p), ("mat_param", _p),
        ("n_textures", _u), ("tex_kind", _p), ("tex_rgb", _p), ("tex_a", _p), ("tex_b", _p), ("tex_scale", _p),
        ("n_perlin", _u), ("perlin_ranvec", _p), ("perlin_perm", _p),
        ("n_images", _u), ("image_rgb", _p), ("image_w", _p), ("image_h", _p),
        ("world_root", _u), ("lights_root", _u), ("background", _d * 3), ("camera", _d * 24),
    ]


class RtRenderParams(C.Structure):
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("sample_offset", C.c_uint32), ("sample_count", C.c_uint32),
                ("total_samples", C.c_uint32), ("max_depth", C.c_uint32), ("row0", C.c_uint32), ("row1", C.c_uint32),
                ("col0", C.c_uint32), ("col1", C.c_uint32), ("seed", C.c_uint64), ("precision", C.c_uint32),
                ("flags", C.c_uint32), ("device", C.c_int32), ("n_gpus", C.c_uint32)]


class RtStats(C.Structure):
    _fields_ = [("kernel_ms", C.c_double), ("h2d_ms", C.c_double), ("d2h_ms", C.c_double), ("samples", C.c_uint64),
                ("launches", C.c_uint64), ("counters", C.c_uint64 * 24), ("sched", C.c_uint64 * 16),
                ("reduce_ms", C.c_double), ("replicate_ms", C.c_double), ("gpus", C.c_uint32), ("reduce_kind", C.c_uint32),
                ("gpu_kernel_ms", C.c_double * 16), ("kernel_kind", C.c_uint32), ("rounds", C.c_uint32), ("tail_ms", C.c_double)]


class RtLimits(C.Structure):
    _fields_ = [("max_stack", C.c_uint32), ("max_frames", C.c_uint32), ("max_medium_depth", C.c_uint32),
                ("max_light_depth", C.c_uint32), ("max_tex_depth", C.c_uint32)]


ABI_SYMBOLS = ["rt_abi_version", "rt_last_error", "rt_device_count", "rt_get_limits", "rt_init", "rt_scene_create", "rt_scene_destroy",
               "rt_render", "rt_render_device", "rt_render_device_static", "rt_render_counted", "rt_render_pathlog", "rt_trace", "rt_texture_value", "rt_tonemap",
               "rt_selftest_division", "rt_selftest_box_filter", "rt_strict_math_eval", "rt_bvh_node_count", "rt_bvh_build", "rt_shutdown"]

_core = None
_host = None


def build(force=False):
    """Compile both shared libraries in-tree (nvcc for sm_100a; g++ for the host mirror)."""
    args = ["-B"] if force else []
    subprocess.check_call(["make", "-C", os.path.join(_HERE, "csrc"), "-j4"] + args, stdout=subprocess.DEVNULL)
    subprocess.check_call(["make", "-C", os.path.join(_HERE, "host")] + args, stdout=subprocess.DEVNULL)


def core():
    """librt_b200.so.  Raises if it has not been built: there is no other implementation to fall back to."""
    global _core
    if _core is None:
        path = os.path.join(_HERE, os.environ.get("RTB_CORE_LIB", "librt_b200.so"))  # override: A/B builds in tools/ only
        if not os.path.exists(path):
            raise RuntimeError("librt_b200.so is missing: run __graft_entry__.build() (no CPU fallback exists)")
        L = C.CDLL(path, mode=C.RTLD_GLOBAL)
        L.rt_last_error.restype = C.c_char_p
        L.rt_scene_create.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_uint64)]
        L.rt_scene_destroy.argtypes = [C.c_uint64]
        L.rt_get_limits.argtypes = [C.POINTER(RtLimits)]
        for f in (L.rt_render, L.rt_render_counted):
            f.argtypes = [C.c_uint64, C.POINTER(RtRenderParams), C.c_void_p, C.POINTER(RtStats)]
        L.rt_render_device.argtypes = [C.c_uint64, C.POINTER(RtRenderParams), C.c_void_p, C.c_void_p, C.POINTER(RtStats)]
        L.rt_render_device_static.argtypes = [C.c_uint64, C.POINTER(RtRenderParams), C.c_void_p, C.c_void_p]
        L.rt_trace.argtypes = [C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p, C.c_double, C.c_double, C.c_uint64,
                               C.c_uint32] + [C.c_void_p] * 8
        L.rt_texture_value.argtypes = [C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p]
        L.rt_init.argtypes = [C.c_int, C.c_uint32]
        L.rt_selftest_box_filter.argtypes = [C.c_uint64, C.c_uint64, C.POINTER(C.c_uint64)]
        L.rt_render_pathlog.argtypes = [C.c_uint64, C.POINTER(RtRenderParams), C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p]
        L.rt_strict_math_eval.argtypes = [C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        L.rt_tonemap.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p]
        L.rt_bvh_node_count.restype = C.c_uint32
        L.rt_bvh_node_count.argtypes = [C.c_uint32]
        L.rt_bvh_build.argtypes = [C.c_uint32, C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        L.rt_selftest_division.argtypes = [C.c_uint64, C.c_uint64, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
        _core = L
    return _core


def host():
    global _host
    if _host is None:
        core()
        path = os.path.join(_HERE, "librt_host.so")
        if not os.path.exists(path):
            raise RuntimeError("librt_host.so is missing: run __graft_entry__.build()")
        L = C.CDLL(path)
        L.rth_last_error.restype = C.c_char_p
        L.rth_scene_create.restype = C.c_void_p
        L.rth_scene_create.argtypes = [C.c_char_p, C.c_int64, C.c_double, C.c_double, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int]
        L.rth_scene_destroy.argtypes = [C.c_void_p]
        L.rth_scene_desc.restype = C.POINTER(RtSceneDesc)
        L.rth_scene_desc.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.rth_scene_g1.argtypes = [C.c_void_p, C.c_void_p]
        L.rth_run_render.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]
        L.rth_set_device_bvh.argtypes = [C.c_int]
        L.rth_last_device_bvh_ms.restype = C.c_double
        L.rth_write_image.argtypes = [C.c_char_p, C.c_void_p, C.c_int, C.c_int, C.c_int]
        _host = L
    return _host


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _check(rc):
    if rc != 0:
        raise RuntimeError("rt error %d: %s" % (rc, core().rt_last_error().decode()))


def load_earth():
    """The reference's earthmap.jpg decoded to RGB8 (Scenes.hs:157-165 reads it with JuicyPixels; no JPEG decoder is assumed
    here, so the decoded texels ship with the package: ray-tracing_b200/data/earthmap_rgb8.npz, made by tests/golden/make_fixtures.py in the build container)."""
    return np.load(os.path.join(_HERE, "data", "earthmap_rgb8.npz"))["rgb"]


def init_gpus(n=0, nccl=False):
    """rt_init: make n devices (0 = all visible) the default of every later render(); returns the number in use."""
    rc = core().rt_init(n, INIT_NCCL if nccl else 0)
    if rc < 0:
        _check(rc)
    return rc


from ._scene import HostScene, DeviceScene, desc_arrays  # noqa: E402,F401
from ._progressive import ProgressiveRender  # noqa: E402,F401
