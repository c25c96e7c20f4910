"""Progressive accumulation / checkpoint / resume on the device path: a frame rendered in uneven steps,
checkpointed, and finished from the file on a fresh scene handle equals the single rt_render call."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

W, H = 60, 40


@pytest.mark.parametrize("name", ["cornell", "cornell_smoke"])
def test_progressive_equals_one_shot(gpu, tmp_path, name):
    hs = gpu.HostScene(name)
    ds = hs.upload(W, H)
    full, _ = ds.render(W, H, 13, depth=50, seed=5)
    pr = gpu.ProgressiveRender(ds, W, H, 13, depth=50, seed=5, tag=name)
    pr.step(1); pr.step(4)
    half = pr.image()
    assert half.shape == (H, W, 3) and half.dtype == np.uint8 and half.max() > 0
    ck = tmp_path / "frame.npz"
    pr.save(ck)
    ds.destroy()
    ds2 = gpu.HostScene(name).upload(W, H)  # a new handle, as after a restart
    pr2 = gpu.ProgressiveRender.resume(ds2, ck, width=W, height=H, total_spp=13, seed=5, tag=name)
    assert pr2.done == 5
    while pr2.remaining:
        pr2.step(3)
    ds2.destroy()
    assert pr2.done == 13 and pr2.kernel_ms > 0
    assert np.allclose(pr2.sums, full, rtol=1e-12, atol=1e-12, equal_nan=True)
    from ray_tracing_b200._scene import tonemap
    assert np.array_equal(pr2.image(), tonemap(full, 13))
