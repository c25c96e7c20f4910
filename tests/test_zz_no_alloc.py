"""rt_render_device_static (include/rt_abi.h): the allocation-free variant of the render entry point — renderKernel's scratch (work
counters + the double framebuffer it accumulates into) lives in a static __device__ array of the library.  No counterpart in the
reference.  Checked through the C ABI against the existing entry point (rt_render) on
the same streams: per path the arithmetic is identical, only the order of the double additions into a pixel differs (1e-12, the
tolerance of tests/test_gpu_multi.py / test_gpu_wavefront.py).  Collected last on purpose (pytest -x)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _close(a, b):
    nan = np.isnan(b)
    return np.array_equal(np.isnan(a), nan) and np.allclose(a[~nan], b[~nan], rtol=1e-12, atol=1e-12)


@pytest.fixture(scope="module")
def torch_cuda(gpu):
    import torch
    assert torch.cuda.is_available()
    return torch


@pytest.mark.parametrize("name,W,H", [("cornell", 100, 100), ("random", 120, 80)])
def test_static_scratch_equals_rt_render_two_calls_in_a_row(gpu, torch_cuda, name, W, H):
    torch = torch_cuda
    ds = gpu.HostScene(name).upload(W, H)
    d_out = torch.full((H, W, 3), -7.0, dtype=torch.float64, device="cuda:0")
    try:
        for seed, spp in ((5, 16), (11, 9)):  # the second call finds the first call's sums in the scratch and in d_out: both are replaced
            ds.render_device_static(d_out.data_ptr(), W, H, spp, seed=seed)
            torch.cuda.synchronize()
            got = d_out.cpu().numpy()
            ref, st = ds.render(W, H, spp, seed=seed)  # the library's own choice of kernel (wavefront for cornell)
            same, st2 = ds.render(W, H, spp, seed=seed, flags=gpu.FLAG_NO_WAVEFRONT | gpu.FLAG_NO_COOP)  # renderKernel, as the static variant
            assert st2["kernel"] == "renderKernel"
            assert _close(got, same), float(np.nanmax(np.abs(got - same)))
            assert _close(got, ref), float(np.nanmax(np.abs(got - ref)))
            assert np.nansum(got) > 0
    finally:
        ds.destroy()


def test_static_scratch_region_and_stale_contents(gpu, torch_cuda):
    torch = torch_cuda
    W, H = 96, 64
    ds = gpu.HostScene("cornell").upload(W, H)
    try:
        d_out = torch.full((H, W, 3), -7.0, dtype=torch.float64, device="cuda:0")
        ds.render_device_static(d_out.data_ptr(), W, H, 8, seed=2)  # fills the whole scratch framebuffer
        ds.render_device_static(d_out.data_ptr(), W, H, 12, seed=9, region=(8, 40, 10, 70), precision=gpu.PREC_F32)
        torch.cuda.synchronize()
        got = d_out.cpu().numpy()
        full, _ = ds.render(W, H, 8, seed=2)
        ref = full.copy()
        ds.render(W, H, 12, seed=9, region=(8, 40, 10, 70), precision=gpu.PREC_F32, out=ref)
        assert _close(got, ref)
        assert not _close(got, full)
    finally:
        ds.destroy()


def test_static_scratch_limits_and_errors(gpu, torch_cuda):
    torch = torch_cuda
    ds = gpu.HostScene("cornell").upload(64, 64)
    try:
        assert gpu.STATIC_MAX_PIXELS == 512 * 512
        d_out = torch.zeros((520, 512, 3), dtype=torch.float64, device="cuda:0")
        ds.render_device_static(d_out.data_ptr(), 512, 512, 1, seed=1)  # exactly the maximum is accepted
        with pytest.raises(RuntimeError, match="RT_STATIC_MAX_PIXELS"):
            ds.render_device_static(d_out.data_ptr(), 512, 513, 1, seed=1)
        with pytest.raises(RuntimeError, match="renderKernel only"):
            ds.render_device_static(d_out.data_ptr(), 64, 64, 1, seed=1, flags=gpu.FLAG_WAVEFRONT)
        with pytest.raises(RuntimeError, match="null"):
            ds.render_device_static(0, 64, 64, 1, seed=1)
        # fp64: the scene is uploaded in that arithmetic by the first fp64 call of an ALLOCATING entry point, never by this one
        with pytest.raises(RuntimeError, match="not on the device"):
            ds.render_device_static(d_out.data_ptr(), 64, 64, 2, seed=1, precision=gpu.PREC_F64)
        ref, _ = ds.render(64, 64, 2, seed=1, precision=gpu.PREC_F64)
        ds.render_device_static(d_out.data_ptr(), 64, 64, 2, seed=1, precision=gpu.PREC_F64)
        torch.cuda.synchronize()
        assert _close(d_out.cpu().numpy().reshape(-1)[:64 * 64 * 3].reshape(64, 64, 3), ref)
    finally:
        ds.destroy()


def test_static_scratch_allocates_nothing(gpu, torch_cuda):
    """device memory in use does not move across calls (after the warm-up call that resolves the symbol and sizes the kernel's stack)"""
    torch = torch_cuda
    W, H = 128, 96
    ds = gpu.HostScene("random").upload(W, H)
    try:
        d_out = torch.zeros((H, W, 3), dtype=torch.float64, device="cuda:0")
        ds.render_device_static(d_out.data_ptr(), W, H, 4, seed=1)
        torch.cuda.synchronize()
        free0, _ = torch.cuda.mem_get_info(0)
        for s in range(3):
            ds.render_device_static(d_out.data_ptr(), W, H, 4, seed=2 + s)
        torch.cuda.synchronize()
        free1, _ = torch.cuda.mem_get_info(0)
        assert free1 == free0, (free0, free1)
    finally:
        ds.destroy()
