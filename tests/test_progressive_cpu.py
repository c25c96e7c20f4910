"""Host logic of progressive accumulation / checkpoint / resume (ray-tracing_b200/_progressive.py) with a stand-in for the device
scene: the arithmetic of the stand-in is irrelevant, only that a sample range's contribution depends on (pixel, sample) alone,
as the device's streams guarantee (tests/test_gpu_progressive.py checks the real thing on the GPU)."""
import numpy as np
import pytest


class _RangeScene:
    """render() = sum over s in [s0, s0+n) of f(pixel, s, seed): what rt_render's sample-range contract promises."""

    def __init__(self):
        self.calls = []

    def render(self, W, H, spp, depth=50, seed=1, s0=0, precision=0, flags=0, total=None):
        self.calls.append((s0, spp, total))
        y, x = np.mgrid[0:H, 0:W]
        out = np.zeros((H, W, 3))
        for s in range(s0, s0 + spp):
            for c in range(3):
                out[:, :, c] += np.sin(0.37 * x + 0.11 * y + 1.7 * s + c + seed) ** 2
        return out, dict(samples=W * H * spp, kernel_ms=0.5 * spp, launches=1)


def test_steps_partition_the_frame(R, tmp_path):
    sc = _RangeScene()
    full, _ = sc.render(16, 9, 20, seed=5)
    pr = R.ProgressiveRender(sc, 16, 9, 20, seed=5, tag="fake")
    assert pr.step(3) == 3 and pr.step(8) == 8 and pr.done == 11 and pr.remaining == 9
    assert pr.step(100) == 9 and pr.step() == 0 and pr.step(4) == 0  # clipped to the frame, then nothing left
    assert np.allclose(pr.sums, full, rtol=1e-13, atol=0)
    assert [c[:2] for c in sc.calls[1:]] == [(0, 3), (3, 8), (11, 9)] and all(c[2] == 20 for c in sc.calls[1:])
    assert np.allclose(pr.mean(), full / 20) and pr.kernel_ms == 10.0


def test_checkpoint_resume_in_another_object(R, tmp_path):
    sc = _RangeScene()
    full, _ = sc.render(8, 5, 12, seed=9)
    pr = R.ProgressiveRender(sc, 8, 5, 12, depth=7, seed=9, precision=1, tag="fake")
    pr.step(5)
    ck = tmp_path / "frame.npz"
    pr.save(ck)
    pr.step(2)          # work after the checkpoint is lost with the "crash"
    pr.save(ck)         # a second save replaces the file atomically
    assert not (tmp_path / "frame.npz.tmp").exists()
    sc2 = _RangeScene()
    pr2 = R.ProgressiveRender.resume(sc2, ck, width=8, height=5, seed=9, tag="fake")
    assert pr2.done == 7 and pr2.meta["depth"] == 7 and pr2.meta["precision"] == 1 and pr2.kernel_ms == 3.5
    pr2.step()
    assert pr2.done == 12 and sc2.calls == [(7, 5, 12)]
    assert np.allclose(pr2.sums, full, rtol=1e-13, atol=0)


def test_resume_refuses_a_different_frame(R, tmp_path):
    pr = R.ProgressiveRender(_RangeScene(), 8, 5, 12, seed=9, tag="cornell")
    pr.step(1)
    ck = tmp_path / "f.npz"
    pr.save(ck)
    for kw in (dict(seed=10), dict(width=9), dict(total_spp=13), dict(tag="random")):
        with pytest.raises(ValueError, match="checkpoint was made with"):
            R.ProgressiveRender.resume(_RangeScene(), ck, **kw)
    with pytest.raises(TypeError):
        R.ProgressiveRender.resume(_RangeScene(), ck, colour="red")
    with pytest.raises(ValueError):
        R.ProgressiveRender(_RangeScene(), 0, 5, 1)
    # an empty frame and a zero-sample frame are legal and do nothing
    z = R.ProgressiveRender(_RangeScene(), 4, 4, 0)
    assert z.step() == 0 and z.done == 0 and np.all(z.mean() == 0)


def test_short_render_is_detected(R):
    class Short(_RangeScene):
        def render(self, *a, **k):
            out, st = super().render(*a, **k)
            st["samples"] -= 1
            return out, st
    with pytest.raises(RuntimeError, match="traced"):
        R.ProgressiveRender(Short(), 4, 4, 2).step()
