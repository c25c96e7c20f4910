"""GPU parity of the wavefront path (ray-tracing_b200/csrc/rt_wavefront.cuh): traversal and shading as separate
kernels over path-state slots in HBM.  Per path the steps, RNG draws and arithmetic are renderKernel's (the same step functions, the
same shadeStep / lpdfAdvance), so the bar is EQUALITY with renderKernel's frame on the same streams, up to the order of the double
additions into a pixel (1e-12) — on every scene, media and nested frames included — plus the sharding / edge-case behaviour of
tests/test_gpu_parity.py.  renderKernel itself is held to the oracle there and in tests/test_gpu_strict.py."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ALL = ["cornell", "cornell_smoke", "simple_light", "earth", "two_perlin", "two_spheres", "random_book1", "random", "next_week_final", "synthetic", "quirks"]
W, H = 60, 40


def _same(a, b):
    nan = np.isnan(b)
    return np.array_equal(np.isnan(a), nan) and np.allclose(a[~nan], b[~nan], rtol=1e-12, atol=1e-12)


@pytest.mark.parametrize("name,lit", [(n, False) for n in ALL] + [("random", True), ("next_week_final", True), ("quirks", True)])
def test_wavefront_frame_equals_render_kernel_frame(gpu, name, lit):
    hs = gpu.HostScene(name, lit=lit, synth_n=30)
    ds = hs.upload(W, H)
    for prec in (gpu.PREC_MIXED, gpu.PREC_F32):
        ref, sr = ds.render(W, H, 8, seed=21, precision=prec, flags=gpu.FLAG_NO_WAVEFRONT | gpu.FLAG_NO_COOP)
        img, sw = ds.render(W, H, 8, seed=21, precision=prec, flags=gpu.FLAG_WAVEFRONT | gpu.FLAG_NO_COOP)
        assert sr["kernel"] == "renderKernel" and sw["kernel"] == "wavefront" and sw["rounds"] > 0 and sw["samples"] == W * H * 8
        assert _same(img, ref), (name, prec, float(np.nanmax(np.abs(img - ref))))
    ds.destroy()


def test_wavefront_full_size_cornell(gpu):
    """cfg4's own frame size (more paths than slots: regeneration through many rounds), sample slices and row bands additive."""
    hs = gpu.HostScene("cornell")
    ds = hs.upload(500, 500)
    ref, _ = ds.render(500, 500, 12, seed=5, flags=gpu.FLAG_NO_WAVEFRONT)
    img, st = ds.render(500, 500, 12, seed=5, flags=gpu.FLAG_WAVEFRONT)
    assert st["kernel"] == "wavefront" and _same(img, ref)
    a, _ = ds.render(500, 500, 5, seed=5, s0=0, flags=gpu.FLAG_WAVEFRONT)
    b, _ = ds.render(500, 500, 7, seed=5, s0=5, flags=gpu.FLAG_WAVEFRONT)
    assert _same(a + b, ref)
    top = np.zeros((500, 500, 3)); ds.render(500, 500, 12, seed=5, region=(0, 170, 0, 500), out=top, flags=gpu.FLAG_WAVEFRONT)
    bot = np.zeros((500, 500, 3)); ds.render(500, 500, 12, seed=5, region=(170, 500, 0, 500), out=bot, flags=gpu.FLAG_WAVEFRONT)
    assert np.all(top[170:] == 0) and _same(top + bot, ref)
    part = np.zeros((500, 500, 3)); ds.render(500, 500, 12, seed=5, region=(33, 300, 100, 411), out=part, flags=gpu.FLAG_WAVEFRONT)
    assert _same(part[33:300, 100:411], ref[33:300, 100:411]) and part[:33].sum() == 0 and part[:, :100].sum() == 0
    ds.destroy()


def test_wavefront_edge_cases_and_selection(gpu):
    hs = gpu.HostScene("cornell")
    ds = hs.upload(W, H)
    fl = gpu.FLAG_WAVEFRONT
    img, st = ds.render(W, H, 0, flags=fl)
    assert st["launches"] == 0 and np.all(img == 0)
    img, _ = ds.render(W, H, 3, depth=0, flags=fl)  # rayColor with depth <= 0 is black (Lib.hs:1303-1304)
    assert np.all(img == 0)
    a, _ = ds.render(W, H, 2, depth=1, flags=fl)
    b, _ = ds.render(W, H, 2, depth=1, flags=gpu.FLAG_NO_WAVEFRONT)
    assert _same(a, b)
    a, _ = ds.render(W, H, 2, depth=2, flags=fl)    # the light queries of the last bounce run, the bounce's hit does not
    b, _ = ds.render(W, H, 2, depth=2, flags=gpu.FLAG_NO_WAVEFRONT)
    assert _same(a, b)
    a, _ = ds.render(1, 1, 7, flags=fl)             # ragged: a single pixel
    b, _ = ds.render(1, 1, 7, flags=gpu.FLAG_NO_WAVEFRONT)
    assert a.shape == (1, 1, 3) and _same(a, b)
    with pytest.raises(RuntimeError, match="RT_FLAG_WAVEFRONT"):
        ds.render(W, H, 1, precision=gpu.PREC_F64, flags=fl)
    with pytest.raises(RuntimeError, match="RT_FLAG_WAVEFRONT"):
        ds.render(W, H, 1, flags=fl | gpu.FLAG_STRICT_MATH)
    _, st = ds.render(W, H, 4, counted=True)        # instrumented launches stay on renderKernel
    assert st["kernel"] == "renderKernel"
    ds.destroy()
