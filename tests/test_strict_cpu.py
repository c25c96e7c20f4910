"""CPU tests of the parity instruments added in round 2: the shared strict-math functions (accuracy against libm), the oracle's
beta-order throughput (the device's multiplication order) against the reference's multiply-on-return chain, the per-path bounce
log, and the host-side validation of texture nesting."""
import ctypes as C

import numpy as np
import pytest


def _ulps(a, b):
    u = np.abs(np.nextafter(b, np.inf) - b)
    with np.errstate(invalid="ignore"):
        r = np.abs(a - b) / u
    return np.where(a == b, 0.0, r)


def test_strict_math_is_accurate(O):
    """rt_strict_math.h is not libm, but it must be a faithful sin / cos / atan2 / asin / log: <= 2 ulp in double over the argument
    ranges the renderer uses (angles 2 pi xi, checker sin(10 x) at |x| up to 1e4, asin / atan2 of unit normals, log of xi)."""
    rng = np.random.default_rng(1)
    x = np.concatenate([rng.uniform(-7, 7, 200000), rng.uniform(-1e5, 1e5, 200000), rng.uniform(-1e-6, 1e-6, 1000)])
    assert _ulps(O.strict_math_eval("sin", x), np.sin(x)).max() <= 2
    assert _ulps(O.strict_math_eval("cos", x), np.cos(x)).max() <= 2
    y, xx = rng.uniform(-2, 2, 300000), rng.uniform(-2, 2, 300000)
    y[::5] *= 1e-7; xx[::7] *= 1e-9
    assert _ulps(O.strict_math_eval("atan2", y, xx), np.arctan2(y, xx)).max() <= 2
    a = np.concatenate([rng.uniform(-1, 1, 200000), 1 - rng.uniform(0, 1e-6, 1000), [1.0, -1.0, 0.0]])
    assert _ulps(O.strict_math_eval("asin", a), np.arcsin(a)).max() <= 3
    l = np.concatenate([rng.uniform(0, 1, 200000), np.ldexp(rng.uniform(0.5, 1, 5000), -rng.integers(0, 60, 5000)), 1 - rng.uniform(0, 1e-5, 1000)])
    assert _ulps(O.strict_math_eval("log", l), np.log(l)).max() <= 2
    # special values the renderer can reach
    assert O.strict_math_eval("log", np.array([0.0]))[0] == -np.inf
    assert np.isnan(O.strict_math_eval("asin", np.array([1.0000001]))[0])
    assert O.strict_math_eval("atan2", np.array([0.0, -0.0]), np.array([-1.0, -1.0])).tolist() == [np.pi, -np.pi]
    assert np.isnan(O.strict_math_eval("sin", np.array([np.inf]))[0])
    # fp32 entry points: evaluated through double and rounded once -> correctly rounded except at double-rounding ties
    xf = rng.uniform(-7, 7, 200000).astype(np.float32).astype(np.float64)
    got = O.strict_math_eval("sin", xf, as_float=True)
    want = np.sin(xf).astype(np.float32).astype(np.float64)
    assert (got == want).mean() > 0.9999


def test_beta_order_is_the_same_estimator(oracle_scene, O):
    """Forward-accumulated throughput (the device's order) and the reference's multiply-on-return chain take the same decisions on
    the same streams and differ only in the rounding of the final product (and never in which pixels are NaN)."""
    for name, lit in (("cornell", False), ("random", True), ("next_week_final", False), ("quirks", False)):
        s = oracle_scene(name, lit)
        for prec, tol in ((O.PREC_F64, 1e-12), (O.PREC_MIXED, 2e-5)):
            a, _ = s.render(40, 30, 4, prec=prec, mode=1, seed=8)
            b, _ = s.render(40, 30, 4, prec=prec, mode=1, seed=8, order="beta")
            assert np.array_equal(np.isnan(a), np.isnan(b))
            ok = ~np.isnan(a)
            inf = np.isinf(a) | np.isinf(b)
            assert np.array_equal(a[inf], b[inf])
            ok &= ~inf
            assert np.all(np.abs(a[ok] - b[ok]) <= tol * (np.abs(a[ok]) + 1e-3)), np.abs(a[ok] - b[ok]).max()


def test_strict_oracle_tracks_libm_oracle(oracle_scene, O):
    """Strict math changes last-ulp values only: same NaN structure, nearly every pixel within 1e-9 in double."""
    s = oracle_scene("cornell")
    a, _ = s.render(40, 40, 8, prec=O.PREC_F64, mode=1, seed=4)
    b, _ = s.render(40, 40, 8, prec=O.PREC_F64, mode=1, seed=4, strict=True)
    rel = np.abs(a - b) / (np.abs(a) + 1e-3)
    assert (rel.max(axis=2) < 1e-9).mean() > 0.99
    with pytest.raises(ValueError):
        s.render(8, 8, 1, mode=0, strict=True)


def test_oracle_path_log(oracle_scene, O):
    """One row per path (region pixels row-major, samples innermost), one column per bounce: primitive ref or 0xffffffff for a miss,
    0xfffffffe beyond the path's end; the radiance is unchanged by logging."""
    s = oracle_scene("cornell")
    img, prim, t = s.render(12, 10, 3, prec=O.PREC_F64, mode=1, seed=2, pathlog=50)
    plain, _ = s.render(12, 10, 3, prec=O.PREC_F64, mode=1, seed=2)
    assert np.array_equal(img, plain)
    assert prim.shape == (12 * 10 * 3, 50)
    assert np.all(prim[:, 0] != 0xFFFFFFFE)                      # every path has a first query
    nb = (prim != 0xFFFFFFFE).sum(axis=1)
    assert nb.min() >= 1 and nb.max() <= 50
    kinds = prim[prim < 0xFFFFFFFE] >> 28
    assert set(np.unique(kinds)) <= {1, 3, 4, 5, 6}              # spheres, rects, cuboid faces
    assert np.all(t[prim < 0xFFFFFFFE] > 1e-4)
    # a miss ends its path: it is always the last logged bounce
    rows, cols = np.nonzero(prim == 0xFFFFFFFF)
    assert np.array_equal(cols + 1, nb[rows])
    # a sub-region logs the same paths as the full frame
    _, p2, t2 = s.render(12, 10, 3, prec=O.PREC_F64, mode=1, seed=2, pathlog=50, region=(2, 5, 3, 9))
    full = prim.reshape(10, 12, 3, 50)[2:5, 3:9].reshape(-1, 50)
    assert np.array_equal(p2, full)


def test_checker_nesting_is_validated(R):
    """A checker deeper than the compiled textureValue walk, or a checker cycle, is reported by rt_scene_create before any device
    work (the device used to return black after 16 levels)."""
    from test_host_abi import _mini_desc
    L = R.core()
    lim = R.RtLimits(); assert L.rt_get_limits(C.byref(lim)) == 0 and lim.max_tex_depth >= 8
    sid = C.c_uint64()
    u, f = np.uint32, np.float64

    def create(n, cyc=False):
        keep = []
        kind = np.ones(n, u); kind[-1] = 0                        # checker ... checker -> constant
        a = np.arange(1, n + 1).astype(u); a[-1] = 0
        if cyc:
            a[n - 2] = 0                                          # the last checker points back at the first
        d = _mini_desc(R, keep, tex_kind=kind, tex_rgb=np.full((n, 3), 0.5, f), tex_a=a, tex_b=np.full(n, n - 1, u), tex_scale=np.ones(n, f))
        rc = L.rt_scene_create(C.byref(d), -1, C.byref(sid))
        msg = L.rt_last_error().decode()
        if rc == 0:
            L.rt_scene_destroy(sid)
        return rc, msg
    rc, msg = create(int(lim.max_tex_depth) + 2)
    assert rc == -2 and "nested CheckerTexture" in msg
    rc, msg = create(int(lim.max_tex_depth))
    assert rc in (0, -3), msg
    rc, msg = create(5, cyc=True)
    assert rc == -1 and "cycle" in msg


def test_rt_init_without_device(R):
    if R.core().rt_device_count() > 0:
        pytest.skip("a CUDA device is present")
    assert R.core().rt_init(0, 0) == -3 and b"no CUDA device" in R.core().rt_last_error()
