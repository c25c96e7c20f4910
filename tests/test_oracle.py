"""CPU tests of the oracle itself (oracle/): RNG restatement, scene construction, HEAD semantics, and the one reference
artefact that pins anything: cornellBox1000.png (statistically)."""
import os

import numpy as np
import pytest

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_splitmix_restatement(O):
    # splitmix-0.1 mkSMGen / nextWord64 restated independently in Python ints
    M = (1 << 64) - 1

    def mix64(z):
        z = ((z ^ (z >> 33)) * 0xff51afd7ed558ccd) & M
        z = ((z ^ (z >> 33)) * 0xc4ceb9fe1a85ec53) & M
        return z ^ (z >> 33)

    def mix13(z):
        z = ((z ^ (z >> 30)) * 0xbf58476d1ce4e5b9) & M
        z = ((z ^ (z >> 27)) * 0x94d049bb133111eb) & M
        return z ^ (z >> 31)

    def mk(s):
        g = mix13((s + 0x9e3779b97f4a7c15) & M) | 1
        if bin(g ^ (g >> 1)).count("1") < 24:
            g ^= 0xaaaaaaaaaaaaaaaa
        return mix64(s), g
    for seed in (0, 1, 1024, 42, 2**40 + 7):
        s, g = mk(seed)
        words, dbl = O.splitmix(seed, 16)
        for i in range(16):
            s = (s + g) & M
            w = mix64(s)
            assert int(words[i]) == w
            # random-1.2.0 word64ToDoubleUpper: 52 mantissa bits into [1,2), minus 1
            assert dbl[i] == ((w >> 12) * 2.0**-52)
    assert np.all((dbl >= 0) & (dbl < 1))


def test_splitmix_published_known_answers(O):
    """Known answers PUBLISHED with the dependency itself — the doctest examples in the haddocks of splitmix-0.1
    (System.Random.SplitMix, pinned by the reference's stack.yaml:42-44): `mkSMGen 42` shows as
    `SMGen 9297814886316923340 13679457532755275413`, and `unfoldr (Just . nextWord64) (mkSMGen 1337)` starts
    b5c19e300e8b07b3, d600e0e216c0ac76, c54efc3b3cc5af29.  These pin mix64 / mixGamma / nextWord64 of the restatement (the
    reference itself has no tests: test/Spec.hs:1-2); `randGen` is `mkStdGen` = `mkSMGen . fromIntegral` (Random.hs:21)."""
    w, _ = O.splitmix(1337, 3)
    assert [hex(int(x)) for x in w] == ["0xb5c19e300e8b07b3", "0xd600e0e216c0ac76", "0xc54efc3b3cc5af29"]
    # seed / gamma of mkSMGen 42 through the exported generator state: stream(...) is our own keyed construction, so read the
    # state by inverting the first output instead: nextWord64 g = mix64 (seed + gamma)
    M = (1 << 64) - 1

    def mix64(z):
        z = ((z ^ (z >> 33)) * 0xff51afd7ed558ccd) & M
        z = ((z ^ (z >> 33)) * 0xc4ceb9fe1a85ec53) & M
        return z ^ (z >> 33)
    w42, d42 = O.splitmix(42, 2)
    seed, gamma = 9297814886316923340, 13679457532755275413
    assert int(w42[0]) == mix64((seed + gamma) & M) and int(w42[1]) == mix64((seed + 2 * gamma) & M)
    # random-1.2.0 `uniformRM (0, 1) :: Double` = word64ToDoubleUpper: the top 52 bits as a mantissa in [1, 2), minus 1
    assert d42[0] == (int(w42[0]) >> 12) * 2.0 ** -52


def test_stream_float_is_truncated_double(O):
    st, d, f = O.stream(7, 123, 45, 64)
    assert st[1] % 2 == 1  # gamma odd
    assert np.all((d >= 0) & (d < 1)) and np.all((f >= 0) & (f < 1))
    assert np.array_equal(f, np.floor(d * 2.0**23).astype(np.float64) / 2.0**23)
    st2, d2, _ = O.stream(7, 123, 46, 64)
    assert not np.array_equal(d, d2)


def test_scene_shapes(oracle_scene):
    # BVH sizes follow makeBVH's recursion N(n) = 1 + N(n/2) + N(n - n/2), N(1) = N(2) = 1 (Lib.hs:941-961)
    def N(n):
        return 1 if n <= 2 else 1 + N(n // 2) + N(n - n // 2)
    c = oracle_scene("cornell").counts()
    assert c[7] == N(8) + 1 and c[3] == 6 and c[6] == 1 and c[1] == 1 and c[8] == 1 and c[9] == 1
    c = oracle_scene("random_book1").counts()
    assert c[7] == N(int(c[1])) and 480 <= c[1] <= 488
    c = oracle_scene("random").counts()
    assert c[7] == N(int(c[1] + c[2] + c[6])) and c[6] == 1 and c[14] == 1
    c = oracle_scene("next_week_final").counts()
    assert c[6] == 400 and c[7] == N(400) + N(1000) + N(11) and c[10] == 2 and c[13] == 1
    c = oracle_scene("synthetic", synth_n=30).counts()
    assert c[1] == 900 and c[3] == 2 and c[7] == N(902)


def test_head_nan_hazard(oracle_scene, O):
    """with an Unhittable light list the HEAD mixture pdf is 0/0 where n.x <= 0 -> NaN pixels; lit variants are NaN-free."""
    img, _ = oracle_scene("two_spheres").render(24, 16, 8, prec=O.PREC_F64, mode=1, seed=3)
    assert np.isnan(img).any()
    img, _ = oracle_scene("two_spheres", lit=True).render(24, 16, 8, prec=O.PREC_F64, mode=1, seed=3)
    assert np.isfinite(img).all()
    img, _ = oracle_scene("cornell").render(24, 24, 8, prec=O.PREC_F64, mode=1, seed=3)
    assert np.isfinite(img).all()


def test_reference_rng_mode_matches_stream_mode_statistically(oracle_scene, O):
    """runRender's per-column SplitMix plumbing (Lib.hs:1491-1523) and the per-(pixel,sample) streams sample the same estimator."""
    s = oracle_scene("cornell")
    a, _ = s.render(40, 40, 64, prec=O.PREC_F64, mode=0, seed=11)
    b, _ = s.render(40, 40, 64, prec=O.PREC_F64, mode=1, seed=12)
    c, _ = s.render(40, 40, 64, prec=O.PREC_F64, mode=1, seed=13)
    ma, mb, mc = a.mean() / 64, b.mean() / 64, c.mean() / 64
    assert abs(ma - mb) < 3 * abs(mb - mc) + 0.02 * mb
    # and determinism
    a2, _ = s.render(40, 40, 64, prec=O.PREC_F64, mode=0, seed=11, threads=3)
    assert np.array_equal(a, a2)


def test_float_oracle_tracks_double_oracle(oracle_scene, O):
    s = oracle_scene("cornell")
    d, _ = s.render(32, 32, 16, prec=O.PREC_F64, mode=1, seed=5)
    f, _ = s.render(32, 32, 16, prec=O.PREC_F32, mode=1, seed=5)
    rel = np.abs(d - f) / (np.abs(d) + 1e-3)
    # same streams: a path coincides until one comparison (Schlick test, rejection sampler, slab test) flips
    assert (rel < 1e-3).mean() > 0.5
    assert abs(d.mean() - f.mean()) < 0.03 * d.mean()


def test_cornell_matches_reference_png(oracle_scene, O):
    """The only reference render that matches HEAD Main.hs (500x500, 1000 spp, Cornell): compare 8-bit channel means and the
    box-filtered image.  125x125 at 256 spp here; tolerance covers Monte-Carlo noise + the sqrt-gamma bias of fewer samples."""
    gold = np.load(os.path.join(GOLD, "cornellBox1000_box4.npy"))
    s = oracle_scene("cornell")
    img, _ = s.render(125, 125, 256, prec=O.PREC_F64, mode=0, seed=2024)
    c8 = O.albedo_to_color(img / 256).astype(np.float32)
    for ch in range(3):
        assert abs(c8[..., ch].mean() - gold[..., ch].mean()) < 1.5, (ch, c8[..., ch].mean(), gold[..., ch].mean())
    # box-filter both to 25x25 to beat the noise down, then compare structure
    a = c8.reshape(25, 5, 25, 5, 3).mean(axis=(1, 3)); b = gold.reshape(25, 5, 25, 5, 3).mean(axis=(1, 3))
    rmse = float(np.sqrt(np.mean((a - b) ** 2)))
    assert rmse < 6.0, rmse


def test_cameras_and_fixed_geometry_match_published_renders(oracle_scene):
    """A second and third reference-held artefact: finalimage.png (In One Weekend final scene, 600x400) and finalScene400.png (The
    Next Week final scene, 500x500), README.md:5,18.  Their RANDOM content cannot pin anything — they were rendered by earlier
    revisions with other generators, and the sphere placement from `randGen 1024` does not coincide (tools/pin_scenes_vs_png.py:
    edge correlation at the null of other seeds; profiles/r2_pin_scenes_vs_png.json) — but the deterministic objects and the
    cameras do: the silhouette of every fixed sphere, projected with the oracle's camera block (newCamera, randomSceneCamera /
    nextWeekFinalSceneCamera), must land on an edge of the published image — stronger than on any copy of the contour shifted by
    8, 16 or 24 pixels in 8 directions.  Fixture: tests/golden/published_render_edges.npz (edge maps, tests/golden/make_fixtures.py)."""
    E = np.load(os.path.join(GOLD, "published_render_edges.npz"))

    def project(cam, P, W, H):
        o, llc, hz, vt = cam[0:3], cam[3:6], cam[6:9], cam[9:12]
        s, t, _ = np.linalg.solve(np.stack([hz, vt, -(P - o)], axis=1), o - llc)
        return s * W, (1.0 - t) * H

    def silhouette(cam, c, r, W, H, n=180):
        o = cam[0:3]
        v = c - o; d = np.linalg.norm(v); w = v / d
        a = np.cross(w, [0.0, 1.0, 0.0]); a /= np.linalg.norm(a); b = np.cross(w, a)
        cc = o + w * (d - r * r / d); rr = r * np.sqrt(1.0 - r * r / (d * d))
        return np.array([project(cam, cc + rr * (np.cos(th) * a + np.sin(th) * b), W, H) for th in 2 * np.pi * np.arange(n) / n])

    def score(e, pts, dx=0.0, dy=0.0):
        x = np.round(pts[:, 0] + dx).astype(int); y = np.round(pts[:, 1] + dy).astype(int)
        ok = (x >= 0) & (x < e.shape[1]) & (y >= 0) & (y < e.shape[0])
        return e[y[ok], x[ok]].astype(np.float64).mean()
    cases = {"finalimage": ("random_book1", [((0, 1, 0), 1.0), ((-4, 1, 0), 1.0), ((4, 1, 0), 1.0)]),
             "finalScene400": ("next_week_final", [((400, 200, 400), 100.0), ((260, 150, 45), 50.0), ((0, 150, 145), 50.0), ((360, 150, 145), 70.0), ((220, 280, 300), 80.0)])}
    for key, (name, objs) in cases.items():
        e = E[key]
        H, W = e.shape
        cam = oracle_scene(name).camera(W, H)
        for c, r in objs:
            pts = silhouette(cam, np.array(c, float), r, W, H)
            on = score(e, pts)
            null = [score(e, pts, R * np.cos(a), R * np.sin(a)) for R in (8, 16, 24) for a in np.arange(8) * np.pi / 4]
            assert on > max(null), (key, c, on, max(null))
