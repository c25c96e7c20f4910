"""CPU check of the scan-stream traversal LOGIC of ray-tracing_b200/csrc/rt_coop.cuh (window of 32 records, first state-changing
lane committed, in-window subtree skips) — a transliteration of coopTraverse's control flow in numpy float64, run against the
oracle's recursion on the cfg5 scene family: the committed records must be exactly the oracle's visits, in order (same count, same
hash, same hit).  The CUDA kernels themselves are held to the same bar on the GPU (tests/test_gpu_coop.py); this test pins the
algorithm where no GPU is available."""
import numpy as np

K_SPHERE, K_RECT_XY, K_RECT_YZ, K_BVH = 1, 3, 5, 7


def _stream(d):
    """pre-order expansion of the world BVH: (kind, ref, a[4], b[4], skip) per record — rt_abi.cu buildDevice"""
    recs = []

    def emit(ref):
        k, i = ref >> 28, ref & 0x07FFFFFF
        at = len(recs)
        if k == K_BVH:
            recs.append([k, ref, d["node_box"][i, :3], d["node_box"][i, 3:], 0])
            emit(int(d["node_left"][i])); emit(int(d["node_right"][i]))
            recs[at][4] = len(recs)
        elif k == K_SPHERE:
            recs.append([k, ref, d["sphere_cr"][i], None, 0])
        else:
            recs.append([k, ref, d["rect"][i], None, 0])
    emit(int(d["world_root"]))
    return recs


def _hmax(x, y): return y if x <= y else x
def _hmin(x, y): return x if x <= y else y


def _box(lo, hi, o, dd, tmin, tmax):  # Lib.hs:798-814, three independent slabs
    for a in range(3):
        with np.errstate(divide="ignore", invalid="ignore"):
            ta, tb = (lo[a] - o[a]) / dd[a], (hi[a] - o[a]) / dd[a]
        t0, t1 = (ta, tb) if ta < tb else (tb, ta)
        if not (_hmin(t1, tmax) > _hmax(t0, tmin)):
            return False
    return True


def _sphere(c, o, dd, tmin, tmax):  # Lib.hs:1081-1095
    oc = o - c[:3]
    dot = lambda p, q: p[0] * q[0] + p[1] * q[1] + p[2] * q[2]  # the reference's dot, left to right (Lib.hs:256)
    a, b, cc = dot(dd, dd), dot(oc, dd), dot(oc, oc) - c[3] * c[3]
    disc = b * b - a * cc
    if disc > 0:
        sd = np.sqrt(disc)
        for t in ((-b - sd) / a, (-b + sd) / a):
            if tmin < t < tmax:
                return True, t
    return False, 0.0


def _rect(kind, r, o, dd, tmin, tmax):  # Lib.hs:1014-1020
    ci, cj, ck = {3: (0, 1, 2), 4: (0, 2, 1), 5: (1, 2, 0)}[kind]
    with np.errstate(divide="ignore", invalid="ignore"):
        t = (r[4] - o[ck]) / dd[ck]
    if t < tmin or t > tmax:
        return False, 0.0
    i, j = o[ci] + t * dd[ci], o[cj] + t * dd[cj]
    if i < r[0] or i > r[1] or j < r[2] or j > r[3]:
        return False, 0.0
    return True, t


def _coop_traverse(recs, o, dd, tmin, tmax):
    """coopTraverse, lane by lane: returns (visited refs in commit order, hit, tbest, bprim)"""
    n, c = len(recs), 0
    tbest, bprim, hit, visits = tmax, 0, False, []
    while c < n:
        lanes = range(c, min(c + 32, n))
        ev, th = {}, {}
        for i in lanes:  # every lane evaluates its record against the CURRENT tbest
            k, ref, a, b, skip = recs[i]
            if k == K_BVH:
                ev[i] = not _box(a, b, o, dd, tmin, tbest)
            elif k == K_SPHERE:
                ev[i], th[i] = _sphere(a, o, dd, tmin, tbest)
            else:
                ev[i], th[i] = _rect(k, a, o, dd, tmin, tbest)
        alive = set(lanes)
        pos = c
        while True:
            evs = [i for i in sorted(alive) if i >= pos and ev[i]]
            if not evs:
                visits += [recs[i][1] for i in sorted(alive) if i >= pos]
                c += 32
                break
            f = evs[0]
            visits += [recs[i][1] for i in sorted(alive) if pos <= i <= f]
            if recs[f][0] == K_BVH:
                s = recs[f][4]
                if s >= c + 32:
                    c = s
                    break
                alive -= set(range(f + 1, s))
                pos = s
            else:
                tbest, bprim, hit = th[f], recs[f][1], True
                c = f + 1
                break
    return visits, hit, tbest, bprim


def test_scan_stream_commits_exactly_the_reference_visits(R, O):
    n = 12
    hs = R.HostScene("synthetic", synth_n=n)
    d = R.desc_arrays(hs.desc(60, 40).contents)
    recs = _stream(d)
    assert len(recs) == 2 * d["node_box"].shape[0] + 1  # a binary tree in pre-order: N nodes, N + 1 leaf slots (BVHNode x x doubles its leaf, Lib.hs:948)
    osc = O.OracleScene("synthetic", synth_n=n)
    cam = osc.camera(60, 40)
    rng = np.random.default_rng(3)
    prim = O.primary_rays(cam, 60, 40)[::40]
    rnd = np.concatenate([cam[0:3] + rng.normal(size=(60, 3)) * 3.0, rng.normal(size=(60, 3)), rng.random((60, 1))], 1)
    rnd[:6, 4] = 0.0  # axis-parallel directions: zero divisors in the slab tests
    rays = np.concatenate([prim, rnd], 0)
    ref = osc.trace(rays, prec=O.PREC_F64, log_stride=4096)
    assert ref["nvisit"].max() <= 4096 and ref["hit"].sum() > 10
    for r in range(len(rays)):
        visits, hit, tbest, bprim = _coop_traverse(recs, rays[r, 0:3], rays[r, 3:6], 1e-4, np.inf)
        nv = int(ref["nvisit"][r])
        assert len(visits) == nv, (r, len(visits), nv)
        assert np.array_equal(np.array(visits, np.uint32), ref["vlog"][r, :nv]), r
        assert bool(ref["hit"][r]) == hit
        if hit:
            assert int(ref["prim"][r]) == bprim and ref["rec"][r, 0] == tbest
