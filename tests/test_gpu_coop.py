"""GPU parity of the warp-cooperative scan-stream traversal (ray-tracing_b200/csrc/rt_coop.cuh): one camera path
per WARP, 32 lanes testing 32 consecutive records of the BVH laid out in the reference's visit order (Lib.hs:971-988: a node, its
left subtree, its right subtree), first state-changing lane committed.  Visit order is part of the parity contract, so the new
kernels are held to exactly the bars of the one-path-per-lane kernels: K1 bit-exact against the oracle (visit count, hash, log, hit,
t, p, normal), K2 equal to renderKernel's image / path log / counters — same visits, same draws, same arithmetic.  cfg5's scene
family (`synthetic`: a makeBVH over a jittered sphere grid + two rects, light list = one XZ rect) is the eligible one."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

W, H = 60, 40


def _precs(gpu, O):
    return {"f32": (gpu.PREC_F32, O.PREC_F32), "mixed": (gpu.PREC_MIXED, O.PREC_MIXED), "f64": (gpu.PREC_F64, O.PREC_F64)}


@pytest.fixture(scope="module")
def synth(gpu):
    made = {}

    def get(n):
        if n not in made:
            hs = gpu.HostScene("synthetic", synth_n=n)
            made[n] = (hs, hs.upload(W, H))
        return made[n][1]
    yield get
    for _, ds in made.values():
        ds.destroy()


@pytest.mark.parametrize("n", [30, 300])
@pytest.mark.parametrize("prec", ["f64", "f32", "mixed"])
def test_coop_trace_bit_exact(gpu, O, synth, n, prec):
    """K1 through the scan stream: hit, primitive, material, frontFace, t / p / normal and the ordered visits (count, hash, first 64)
    bit-identical to the oracle's recursion, at 900 and at 90 000 spheres (~80 000 `hit` calls per ray)."""
    from test_gpu_parity import _rays
    gp, op = _precs(gpu, O)[prec]
    osc = O.OracleScene("synthetic", synth_n=n)
    rays = _rays(O, osc, n_rand=600 if n == 300 else 3000)
    if prec != "f64":
        rays = rays.astype(np.float32).astype(np.float64)
    g = synth(n).trace(rays, precision=gp, flags=gpu.FLAG_COOP, seed=9, log_stride=64)
    o = osc.trace(rays, prec=op, seed=9, log_stride=64)
    for k in ("hit", "prim", "mat", "front", "nvisit", "vhash", "vlog"):
        assert np.array_equal(g[k], o[k]), (k, int((g[k] != o[k]).sum()))
    assert np.array_equal(g["rec"][:, :7], o["rec"][:, :7])
    assert g["hit"].sum() > 100 and (o["nvisit"].mean() > 20000 or n == 30)
    # and identical to the one-path-per-lane kernel, u / v included (same device functions)
    p = synth(n).trace(rays, precision=gp, flags=gpu.FLAG_NO_COOP, seed=9, log_stride=64)
    for k in ("hit", "prim", "mat", "front", "nvisit", "vhash", "vlog"):
        assert np.array_equal(g[k], p[k]), k
    assert np.array_equal(g["rec"].view(np.uint64), p["rec"].view(np.uint64))


@pytest.mark.parametrize("prec", ["mixed", "f32", "f64"])
def test_coop_render_equals_per_lane_render(gpu, O, synth, prec):
    """Same streams, same visits, same arithmetic: the two kernels' frames differ only by the order of the double additions into a
    pixel; their per-path bounce logs and their algorithmic-work counters are identical."""
    gp, _ = _precs(gpu, O)[prec]
    ds = synth(30)
    a, sa = ds.render(W, H, 8, seed=21, precision=gp, flags=gpu.FLAG_COOP)
    b, sb = ds.render(W, H, 8, seed=21, precision=gp, flags=gpu.FLAG_NO_COOP)
    assert sa["kernel"] == "renderKernelCoop" and sb["kernel"] == "renderKernel"
    assert np.isfinite(a).all() and np.allclose(a, b, rtol=1e-12, atol=1e-12), float(np.abs(a - b).max())
    _, pa, ta = ds.render_pathlog(W, H, 4, 50, seed=5, precision=gp, flags=gpu.FLAG_COOP)
    _, pb, tb = ds.render_pathlog(W, H, 4, 50, seed=5, precision=gp, flags=gpu.FLAG_NO_COOP)
    assert np.array_equal(pa, pb) and np.array_equal(ta.view(np.uint64), tb.view(np.uint64))
    _, ca = ds.render(W, H, 4, counted=True, seed=77, precision=gp, flags=gpu.FLAG_COOP)
    _, cb = ds.render(W, H, 4, counted=True, seed=77, precision=gp, flags=gpu.FLAG_NO_COOP)
    assert ca["counters"] == cb["counters"]


@pytest.mark.parametrize("prec", ["mixed", "f64"])
def test_coop_render_strict_matches_oracle(gpu, O, synth, prec):
    """...and, with the shared strict-math functions, the oracle's frame: identical NaN masks, >= 99.9 % of pixels within 1e-6 / 1e-12,
    identical path logs, exactly equal counters."""
    gp, op = _precs(gpu, O)[prec]
    ds = synth(30)
    osc = O.OracleScene("synthetic", synth_n=30)
    fl = gpu.FLAG_COOP | gpu.FLAG_STRICT_MATH
    img, _ = ds.render(W, H, 8, seed=21, precision=gp, flags=fl)
    ref, _ = osc.render(W, H, 8, prec=op, mode=1, seed=21, strict=True, order="beta")
    assert np.isfinite(img).all() and np.isfinite(ref).all()
    rel = (np.abs(img - ref) / (np.abs(ref) + 1e-3)).max(axis=2)
    assert (rel < (1e-12 if prec == "f64" else 1e-6)).mean() >= 0.999, float(rel.max())
    _, gp_, gt = ds.render_pathlog(W, H, 4, 50, seed=5, precision=gp, flags=fl)
    _, op_, ot = osc.render(W, H, 4, prec=op, mode=1, seed=5, strict=True, order="beta", pathlog=50)
    assert np.array_equal(gp_, op_) and np.array_equal(gt.view(np.uint64), ot.view(np.uint64))
    _, st = ds.render(W, H, 4, counted=True, seed=77, precision=gp, flags=fl)
    _, os_ = osc.render(W, H, 4, prec=op, mode=1, seed=77, stats=True, strict=True, order="beta")
    for k, v in os_.items():
        assert st["counters"][k] == v, (k, st["counters"][k], v)


def test_coop_selection_and_eligibility(gpu, synth):
    """Chosen automatically for a large eligible BVH (131 071 nodes), not for a small one; forcing it on a scene with cuboids /
    frames / media / a BVH light list is RT_ERR_UNSUPPORTED (-2); sharding (sample slices, regions) is invariant as for renderKernel."""
    _, st = synth(30).render(W, H, 1, seed=1)
    assert st["kernel"] == "renderKernel"
    big = synth(300)
    full, st = big.render(W, H, 2, seed=1)
    assert st["kernel"] == "renderKernelCoop" and st["tail_ms"] >= 0
    a, _ = big.render(W, H, 1, seed=1, s0=0)
    b, _ = big.render(W, H, 1, seed=1, s0=1)
    assert np.allclose(a + b, full, rtol=1e-12, atol=1e-12)
    top = np.zeros((H, W, 3)); big.render(W, H, 2, seed=1, region=(0, 17, 0, W), out=top)
    bot = np.zeros((H, W, 3)); big.render(W, H, 2, seed=1, region=(17, H, 0, W), out=bot)
    assert np.allclose(top + bot, full, rtol=1e-12, atol=1e-12)
    img, st = big.render(W, H, 1, depth=0)
    assert np.all(img == 0)
    for name in ("cornell", "random", "next_week_final"):
        ds = gpu.HostScene(name).upload(W, H)
        with pytest.raises(RuntimeError, match="not eligible"):
            ds.render(W, H, 1, flags=gpu.FLAG_COOP)
        with pytest.raises(RuntimeError, match="not eligible"):
            ds.trace(np.zeros((4, 7)) + 1.0, flags=gpu.FLAG_COOP)
        ds.destroy()
