"""CPU tests of the drop-in boundary: the C-ABI library loads and exports every symbol of include/rt_abi.h, the struct layouts
seen from Python match the header, the host mirror's flattener agrees bit-for-bit with the oracle's independent scene
construction, and the library validates scenes and fails loudly without a GPU (no CPU fallback)."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SCENES = ["cornell", "cornell_smoke", "simple_light", "earth", "two_perlin", "two_spheres", "random_book1", "random", "next_week_final", "synthetic", "quirks"]


def test_abi_exports_every_declared_symbol(R):
    hdr = open(os.path.join(ROOT, "include", "rt_abi.h")).read()
    declared = set(re.findall(r"^(?:int|void|uint32_t|const char\*)\s+(rt_[a-z_]+)\s*\(", hdr, re.M))
    assert declared == set(R.ABI_SYMBOLS), declared ^ set(R.ABI_SYMBOLS)
    L = R.core()
    for s in declared:
        assert hasattr(L, s), s
    assert L.rt_abi_version() == 2
    lim = R.RtLimits()
    assert L.rt_get_limits(C.byref(lim)) == 0 and lim.max_frames >= 2 and lim.max_stack >= 48


def test_struct_layouts_match_header(R, tmp_path):
    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "rt_abi.h"\nint main(){printf("%zu %zu %zu %zu %zu %zu\\n",'
                   'sizeof(RtSceneDesc),sizeof(RtRenderParams),sizeof(RtStats),offsetof(RtSceneDesc,camera),offsetof(RtRenderParams,seed),sizeof(RtLimits));return 0;}')
    exe = tmp_path / "sz"
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    got = [int(x) for x in subprocess.check_output([str(exe)]).split()]
    assert got == [C.sizeof(R.RtSceneDesc), C.sizeof(R.RtRenderParams), C.sizeof(R.RtStats), R.RtSceneDesc.camera.offset,
                   R.RtRenderParams.seed.offset, C.sizeof(R.RtLimits)]


@pytest.mark.parametrize("name", SCENES)
@pytest.mark.parametrize("lit", [False, True])
def test_flattener_matches_oracle_scene(R, O, oracle_scene, name, lit):
    """Two independent restatements of Scenes.hs + makeBVH + makePerlin + newCamera (pointer tree in oracle/, index arena in
    ray-tracing_b200/host/) must produce the same canonical flattened scene, bit for bit, from randGen 1024."""
    hs = R.HostScene(name, lit=lit, synth_n=30)
    d = R.desc_arrays(hs.desc(60, 40).contents)
    os_ = oracle_scene(name, lit=lit, synth_n=30)
    roots, bg, g1 = os_.roots()
    assert roots[0] == d["world_root"] and roots[1] == d["lights_root"]
    assert np.array_equal(bg, d["background"]) and np.array_equal(g1, hs.g1())
    assert np.array_equal(os_.camera(60, 40), d["camera"])

    def two(a, b):
        return np.stack([a, b], 1) if len(a) else np.zeros((0, 2), np.uint32)
    D, U = os_.export(1); assert np.array_equal(D, d["sphere_cr"]) and np.array_equal(U[:, 0], d["sphere_mat"])
    D, U = os_.export(2); assert np.array_equal(D, d["moving"]) and np.array_equal(U[:, 0], d["moving_mat"])
    D, U = os_.export(3); assert np.array_equal(D, d["rect"]) and np.array_equal(U[:, 1], d["rect_mat"])
    D, U = os_.export(6); assert np.array_equal(D, d["cuboid"]) and np.array_equal(U[:, 0], d["cuboid_mat"])
    D, U = os_.export(7)
    assert np.array_equal(D, d["node_box"]) and np.array_equal(U[:, 0], d["node_left"]) and np.array_equal(U[:, 1], d["node_right"]) and np.array_equal(U[:, 2], d["node_size"])
    D, U = os_.export(8); assert np.array_equal(D, d["translate_off"]) and np.array_equal(U[:, 0], d["translate_child"])
    D, U = os_.export(9); assert np.array_equal(D, d["rotate_sc"]) and np.array_equal(U, two(d["rotate_axis"], d["rotate_child"]))
    D, U = os_.export(10); assert np.array_equal(D[:, 0], d["medium_neg_inv_density"]) and np.array_equal(U, two(d["medium_mat"], d["medium_boundary"]))
    D, U = os_.export(11); assert np.array_equal(D[:, 0], d["mat_param"]) and np.array_equal(U, two(d["mat_kind"], d["mat_tex"]))
    D, U = os_.export(12)
    assert np.array_equal(D[:, :3], d["tex_rgb"]) and np.array_equal(D[:, 3], d["tex_scale"])
    assert np.array_equal(U, np.stack([d["tex_kind"], d["tex_a"], d["tex_b"]], 1))
    D, U = os_.export(13); assert np.array_equal(D, d["perlin_ranvec"]) and np.array_equal(U, d["perlin_perm"].astype(np.uint32))
    # the rect plane travels in the ref kind: every rect ref found in the graph must carry the plane the oracle recorded
    planes = os_.export(3)[1][:, 0]
    refs = np.concatenate([d["node_left"], d["node_right"], d["translate_child"], d["rotate_child"], d["medium_boundary"], [d["world_root"], d["lights_root"]]]).astype(np.uint32)
    for r in refs:
        k, i = int(r) >> 28, int(r) & 0x0fffffff
        if 3 <= k <= 5:
            assert planes[i] == k - 3


def test_scene_validation_errors(R):
    """Bad descriptions are rejected with RT_ERR_ARG before any device work."""
    hs = R.HostScene("cornell")
    d = hs.desc(16, 16).contents
    sid = C.c_uint64()
    L = R.core()
    bad = R.RtSceneDesc.from_buffer_copy(d); bad.abi_version = 99
    assert L.rt_scene_create(C.byref(bad), -1, C.byref(sid)) == -1 and b"abi_version" in L.rt_last_error()
    bad = R.RtSceneDesc.from_buffer_copy(d); bad.world_root = (7 << 28) | 1000
    assert L.rt_scene_create(C.byref(bad), -1, C.byref(sid)) == -1 and b"root" in L.rt_last_error()
    bad = R.RtSceneDesc.from_buffer_copy(d); bad.node_box = None
    assert L.rt_scene_create(C.byref(bad), -1, C.byref(sid)) == -1
    left = np.ctypeslib.as_array(C.cast(d.node_left, C.POINTER(C.c_uint32)), shape=(d.n_nodes,)).copy()
    left[0] = (7 << 28) | 0  # node 0 is its own child
    bad = R.RtSceneDesc.from_buffer_copy(d); bad.node_left = left.ctypes.data_as(C.c_void_p)
    assert L.rt_scene_create(C.byref(bad), -1, C.byref(sid)) == -1 and b"cycle" in L.rt_last_error()
    assert L.rt_scene_destroy(C.c_uint64(12345)) == -4
    assert L.rt_render(C.c_uint64(12345), None, None, None) == -4


def test_no_cpu_fallback(R):
    """On a box without a CUDA device every compute entry point must fail loudly with RT_ERR_CUDA."""
    if R.core().rt_device_count() > 0:
        pytest.skip("a CUDA device is present")
    hs = R.HostScene("two_spheres")
    with pytest.raises(RuntimeError, match="no CUDA device"):
        hs.upload(8, 8)
    with pytest.raises(RuntimeError, match="no CUDA device"):
        hs.run_render(8, 8, 1)
    with pytest.raises(RuntimeError, match="no CUDA device"):
        from ray_tracing_b200._scene import tonemap
        tonemap(np.zeros(3), 1)


def test_product_does_not_touch_oracle():
    """The product package must not include, import or link anything under oracle/."""
    pkg = os.path.join(ROOT, "ray-tracing_b200")
    for dp, _, fs in os.walk(pkg):
        if "build" in dp:
            continue
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp", "Makefile")):
                txt = open(os.path.join(dp, f)).read()
                for pat in (r'#include\s+"[^"]*oracle', r"^\s*(import|from)\s+oracle", r"liboracle", r"-loracle", r"oracle\.py", r"rt_oracle\.hpp", r"rt_scenes\.hpp"):
                    assert not re.search(pat, txt, re.M), (pat, os.path.join(dp, f))
    for so in ("librt_b200.so", "librt_host.so"):
        out = subprocess.check_output(["ldd", os.path.join(pkg, so)]).decode()
        assert "oracle" not in out


def test_write_image_formats(R, tmp_path):
    """writeImage: P3 text in printRow's format (Lib.hs:299-305, app/Main.hs:58-62) and P6 binary; host-only, no GPU."""
    from ray_tracing_b200._scene import write_image
    rng = np.random.default_rng(5)
    img = rng.integers(0, 256, size=(3, 4, 3), dtype=np.uint8)
    p3, p6 = tmp_path / "a.ppm", tmp_path / "b.ppm"
    write_image(p3, img, binary=False)
    write_image(p6, img, binary=True)
    lines = p3.read_text().split("\n")
    assert lines[:3] == ["P3", "4 3", "255"] and lines[-1] == ""
    for r in range(3):
        assert lines[3 + r] == " ".join(str(int(v)) for v in img[r].reshape(-1))  # unwords (show <$> row), show RGB = "r g b"
    raw = p6.read_bytes()
    assert raw.startswith(b"P6\n4 3\n255\n") and raw[len(b"P6\n4 3\n255\n"):] == img.tobytes()


def test_haskell_binding_matches_header():
    """integration/haskell/LibB200.hsc cannot be compiled here (no GHC): at least every struct field it pokes and every
    constant it names must exist in include/rt_abi.h, every RtSceneDesc member must be filled, and every entry point it
    imports must be exported by the built library."""
    h = open(os.path.join(ROOT, "include", "rt_abi.h")).read()
    s = open(os.path.join(ROOT, "integration", "haskell", "LibB200.hsc")).read()
    for st, f in re.findall(r"#\{(?:poke|ptr)\s+(\w+),\s*(\w+)\}", s):
        body = h[h.index("typedef struct " + st):h.index("} " + st + ";")]
        assert re.search(r"\b" + f + r"\b", body), (st, f)
    for c in re.findall(r"#\{const\s+(\w+)\}", s):
        assert re.search(r"\b" + c + r"\b", h), c
    body = h[h.index("typedef struct RtSceneDesc"):h.index("} RtSceneDesc;")]
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    declared = set(re.findall(r"(\w+)(?:\[\d+\])?;", body))
    filled = set(re.findall(r"#\{(?:poke|ptr)\s+RtSceneDesc,\s*(\w+)\}", s))
    assert declared == filled, declared ^ filled
    import ctypes
    lib = ctypes.CDLL(os.path.join(ROOT, "ray-tracing_b200", "librt_b200.so"))
    for sym in re.findall(r'foreign import ccall \w+\s+"(\w+)"', s):
        assert hasattr(lib, sym), sym


def _mini_desc(R, keep, **kw):
    """A hand-written RtSceneDesc: one Lambertian material on a constant texture, one sphere; kw adds / overrides arrays
    (numpy arrays are kept alive in `keep`)."""
    f, u = np.float64, np.uint32
    arrays = dict(sphere_cr=np.array([[0, 0, -1, 0.5]], f), sphere_mat=np.zeros(1, u), mat_kind=np.zeros(1, u), mat_tex=np.zeros(1, u),
                  mat_param=np.zeros(1, f), tex_kind=np.zeros(1, u), tex_rgb=np.full((1, 3), 0.5, f), tex_a=np.zeros(1, u), tex_b=np.zeros(1, u),
                  tex_scale=np.ones(1, f))
    arrays.update({k: v for k, v in kw.items() if isinstance(v, np.ndarray)})
    d = R.RtSceneDesc()
    d.abi_version = 2
    counts = dict(n_spheres="sphere_mat", n_materials="mat_kind", n_textures="tex_kind", n_nodes="node_left", n_translates="translate_child",
                  n_rotates="rotate_child", n_media="medium_mat", n_rects="rect_mat", n_cuboids="cuboid_mat", n_moving="moving_mat")
    for cnt, arr in counts.items():
        setattr(d, cnt, len(arrays[arr]) if arr in arrays else 0)
    for k, a in arrays.items():
        a = np.ascontiguousarray(a); keep.append(a)
        setattr(d, k, a.ctypes.data_as(C.c_void_p))
    d.world_root = kw.get("world_root", (1 << 28) | 0)
    d.lights_root = kw.get("lights_root", 0)  # Unhittable
    return d


def test_traversal_budget_limits_are_reported(R):
    """Nesting beyond the compiled traversal budget (rt_get_limits) is RT_ERR_UNSUPPORTED (-2) with a message, decided on the host
    before any device work; at the limit the description passes validation (and then stops at the missing device here)."""
    L = R.core()
    lim = R.RtLimits(); assert L.rt_get_limits(C.byref(lim)) == 0
    assert lim.max_frames >= 4 and lim.max_stack >= 32 and lim.max_light_depth >= 4 and lim.max_medium_depth == 1
    sid = C.c_uint64()
    u, f = np.uint32, np.float64
    SPH, BVH, TRN, MED = 1 << 28, 7 << 28, 8 << 28, 10 << 28

    def create(**kw):
        keep = []
        rc = L.rt_scene_create(C.byref(_mini_desc(R, keep, **kw)), -1, C.byref(sid))
        msg = L.rt_last_error().decode()
        if rc == 0:
            L.rt_scene_destroy(sid)
        return rc, msg

    def chain(n):  # translate i -> translate i+1 -> ... -> the sphere
        return dict(translate_off=np.zeros((n, 3), f), translate_child=np.array([TRN | (i + 1) for i in range(n - 1)] + [SPH], u), world_root=TRN | 0)
    rc, msg = create(**chain(int(lim.max_frames) + 1))
    assert rc == -2 and "nested Translate/Rotate frames" in msg
    rc, msg = create(**chain(int(lim.max_frames)))
    assert rc in (0, -3), msg  # valid: created on a GPU box, "no CUDA device" here

    # a medium whose boundary is a medium
    rc, msg = create(medium_neg_inv_density=np.array([-100.0, -100.0], f), medium_mat=np.zeros(2, u), medium_boundary=np.array([MED | 1, SPH], u), world_root=MED | 0)
    assert rc == -2 and "ConstantMedium nested" in msg

    def left_chain(n):  # node i = (node i+1, sphere): a degenerate BVH of depth n
        return dict(node_box=np.tile(np.array([-1, -1, -2, 1, 1, 0], f), (n, 1)), node_left=np.array([BVH | (i + 1) for i in range(n - 1)] + [SPH], u),
                    node_right=np.full(n, SPH, u), node_size=np.arange(n + 1, 1, -1).astype(u))
    rc, msg = create(**left_chain(int(lim.max_stack) + 4), world_root=BVH | 0)
    assert rc == -2 and "traversal stack" in msg
    rc, msg = create(**left_chain(int(lim.max_light_depth) + 2), world_root=SPH, lights_root=BVH | 0)
    assert rc == -2 and "lights BVH too deep" in msg
    rc, msg = create(**left_chain(3), world_root=BVH | 0, lights_root=BVH | 1)
    assert rc in (0, -3), msg


def test_main_example_builds(R, tmp_path):
    """examples/Main.cpp = app/Main.hs against the C++ mirror: it must compile and link against the two libraries; without a CUDA device it
    fails loudly (no CPU path), with one it prints the P3 image of the Cornell box."""
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    pkg_dir = os.path.join(root, "ray-tracing_b200")
    exe = str(tmp_path / "Main")
    subprocess.check_call(["g++", "-O1", "-std=c++17", "-I" + os.path.join(pkg_dir, "host"), "-I" + os.path.join(root, "include"), os.path.join(root, "examples", "Main.cpp"),
                           "-L" + pkg_dir, "-lrt_host", "-lrt_b200", "-Wl,-rpath," + pkg_dir, "-o", exe])
    if R.core().rt_device_count() > 0:
        return  # rendering 500x500x1000 belongs to the GPU suite (test_gpu_parity.py renders the same frame through the same mirror)
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode == 1 and "no CUDA device" in r.stderr and r.stdout == ""
