"""Multi-GPU INSIDE the C ABI: one process, one rt_render call, n devices.  The reference parallelises inside
its one pure entry point (runRender sparks, Lib.hs:1519-1520; called once from app/Main.hs:58), so the drop-in shards inside the
call: the scene is replicated, samples are split into contiguous slices (interleaved rows when there are fewer samples than devices) and
the partial double framebuffers are summed onto the scene's device by the library's peer-memory kernel over NVLink, or by
ncclReduce after rt_init(.., RT_INIT_NCCL).  Because sample s of pixel p owns its stream, the n-device frame equals the 1-device
frame up to the order of the double additions (1e-12).  Needs >= 2 visible devices: the tests skip on a single-GPU box (the
sharding arithmetic itself is covered there by test_gpu_parity.py::test_sharding_invariance and on CPU by tests/test_dist_cpu.py)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

W, H = 96, 64


@pytest.fixture(scope="module")
def ndev(gpu):
    n = gpu.core().rt_device_count()
    if n < 2:
        pytest.skip("needs >= 2 CUDA devices (this box has %d)" % n)
    return n


def _close(a, b):
    nan = np.isnan(b)
    return np.array_equal(np.isnan(a), nan) and np.allclose(a[~nan], b[~nan], rtol=1e-12, atol=1e-12)


@pytest.mark.parametrize("name", ["cornell", "random", "next_week_final"])
def test_n_gpu_frame_equals_one_gpu_frame(gpu, ndev, name):
    hs = gpu.HostScene(name)
    ds = hs.upload(W, H)
    one, s1 = ds.render(W, H, 24, seed=5, n_gpus=1)
    assert s1["gpus"] == 1 and s1["reduce_kind"] == "none"
    for n in sorted({2, ndev}):
        img, st = ds.render(W, H, 24, seed=5, n_gpus=n)
        assert st["gpus"] == n and st["reduce_kind"] in ("peer-kernel", "staged") and st["samples"] == W * H * 24
        assert len(st["gpu_kernel_ms"]) == n and min(st["gpu_kernel_ms"]) > 0 and st["reduce_ms"] > 0
        assert _close(img, one), float(np.nanmax(np.abs(img - one)))
        again, st2 = ds.render(W, H, 24, seed=5, n_gpus=n)  # replicas are cached: no second replication
        assert st2["replicate_ms"] == 0.0 and _close(again, one)
    ds.destroy()


def test_row_bands_when_fewer_samples_than_devices(gpu, ndev):
    hs = gpu.HostScene("cornell")
    ds = hs.upload(W, H)
    one, _ = ds.render(W, H, 1, seed=9, n_gpus=1)
    img, st = ds.render(W, H, 1, seed=9, n_gpus=ndev)
    assert st["gpus"] == ndev and _close(img, one)
    # a sub-region, sharded, leaves the rest of the caller's buffer untouched
    out = np.full((H, W, 3), -7.0)
    ds.render(W, H, 16, seed=9, n_gpus=ndev, region=(8, 40, 10, 70), out=out)
    ref = np.full((H, W, 3), -7.0)
    ds.render(W, H, 16, seed=9, n_gpus=1, region=(8, 40, 10, 70), out=ref)
    assert _close(out, ref) and np.all(out[:8] == -7.0) and np.all(out[:, :10] == -7.0)
    ds.destroy()


def test_rt_init_default_and_nccl_reduce(gpu, ndev):
    """rt_init(n) makes n devices the default of rt_render (n_gpus = 0); with RT_INIT_NCCL the reduction is an ncclReduce over the
    communicators of ncclCommInitAll.  The frame is the same either way."""
    hs = gpu.HostScene("cornell")
    ds = hs.upload(W, H)
    try:
        one, _ = ds.render(W, H, 16, seed=3, n_gpus=1)
        assert gpu.init_gpus(0) == ndev
        img, st = ds.render(W, H, 16, seed=3)  # n_gpus = 0: the rt_init default
        assert st["gpus"] == ndev and _close(img, one)
        assert gpu.init_gpus(ndev, nccl=True) == ndev
        img, st = ds.render(W, H, 16, seed=3)
        assert st["gpus"] == ndev and st["reduce_kind"] == "nccl" and _close(img, one)
    finally:
        gpu.init_gpus(1)
        ds.destroy()
    ds2 = gpu.HostScene("cornell").upload(W, H)
    _, st = ds2.render(W, H, 2, seed=1)
    ds2.destroy()
    assert st["gpus"] == 1


def test_scene_on_another_device_is_the_reduction_root(gpu, ndev):
    hs = gpu.HostScene("cornell")
    d1 = hs.upload(W, H, device=1)
    d0 = hs.upload(W, H, device=0)
    a, _ = d0.render(W, H, 8, seed=4, n_gpus=1)
    b, st = d1.render(W, H, 8, seed=4, n_gpus=2)
    assert st["gpus"] == 2 and _close(b, a)
    d0.destroy(); d1.destroy()
