#!/usr/bin/env python
"""First-divergence classifier: where and why a device path separates from the oracle's path on the same stream.

Both sides log, per camera path, the result of every main query (`hit world`, Lib.hs:1309) in bounce order — primitive ref and t
(rt_render_pathlog / oracle render(pathlog=...)).  For every path the first bounce whose (primitive, t) differs is classified:

  identical        the whole log agrees bit for bit
  value_drift      same primitives at every bounce, some t differs in value (no decision changed)
  self_hit_flip    at the first differing bounce ONE side re-hit the primitive it had just left (t_min = 1e-4 self-intersection,
                   Lib.hs:77,1309) and the other did not: the hit point's last-bit rounding decides that comparison
  hit_miss_flip    one side hit something, the other missed everything
  other_flip       different primitives otherwise (a grazing silhouette, a Schlick / rejection comparison upstream)

and attributed to the material that scattered the ray into the differing bounce and to whether the PREVIOUS bounce's t was still
bit-identical (then the difference was born in that scatter's own arithmetic — sin / cos / atan2 / asin / log / pow — not inherited).
With --strict both sides use the shared strict-math functions and the expectation is `identical` everywhere.

usage: python tools/parity_classify.py out.json [--strict] [--spp 8] [--scenes a,b]      (test infrastructure: imports oracle/)"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import oracle as O  # noqa: E402
import rtb200_loader  # noqa: E402

pkg = rtb200_loader.load()
MAT = ["Lambertian", "Metal", "Dielectric", "DiffuseLight", "Isotropic"]
UNREACHED, MISS = 0xFFFFFFFE, 0xFFFFFFFF


def prim_material(d):
    """ref -> material kind name, from the flattened scene arrays"""
    def of(ref):
        k, i = int(ref) >> 28, int(ref) & 0x07FFFFFF
        m = {1: "sphere_mat", 2: "moving_mat", 3: "rect_mat", 4: "rect_mat", 5: "rect_mat", 6: "cuboid_mat", 10: "medium_mat"}.get(k)
        return MAT[int(d["mat_kind"][int(d[m][i])])] if m else "?"
    return of


def classify(gp, gt, op, ot, mat_of):
    n = gp.shape[0]
    out = {"paths": int(n), "identical": 0, "value_drift": 0, "self_hit_flip": 0, "hit_miss_flip": 0, "other_flip": 0,
           "by_scatter_material": {}, "prev_t_identical": 0, "first_bounce_hist": {}}
    same_prim = gp == op
    same_t = (gt.view(np.uint64) == ot.view(np.uint64)) | ~(gp < UNREACHED)
    for i in np.nonzero(~(same_prim & same_t).all(axis=1))[0]:
        b = int(np.argmin(same_prim[i] & same_t[i]))
        if same_prim[i].all():
            out["value_drift"] += 1
            continue
        bp = int(np.argmin(same_prim[i]))
        g, o = int(gp[i, bp]), int(op[i, bp])
        prev = int(gp[i, bp - 1]) if bp > 0 else None
        if prev is not None and (g == prev) != (o == prev) and prev < UNREACHED:
            kind = "self_hit_flip"
        elif (g == MISS) != (o == MISS):
            kind = "hit_miss_flip"
        else:
            kind = "other_flip"
        out[kind] += 1
        m = mat_of(prev) if prev is not None and prev < UNREACHED else "camera"
        out["by_scatter_material"][m] = out["by_scatter_material"].get(m, 0) + 1
        if bp > 0 and gt[i, bp - 1].view(np.uint64) == ot[i, bp - 1].view(np.uint64):
            out["prev_t_identical"] += 1
        out["first_bounce_hist"][str(min(bp, 10))] = out["first_bounce_hist"].get(str(min(bp, 10)), 0) + 1
        del b
    flips = out["self_hit_flip"] + out["hit_miss_flip"] + out["other_flip"]
    out["identical"] = int(n - flips - out["value_drift"])
    out["diverged_fraction"] = flips / max(n, 1)
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("out")
    ap.add_argument("--strict", action="store_true")
    ap.add_argument("--spp", type=int, default=8)
    ap.add_argument("--size", default="60x40")
    ap.add_argument("--scenes", default="cornell,cornell_smoke,simple_light+lit,two_spheres+lit,random_book1+lit,random+lit,next_week_final+lit,quirks+lit,synthetic")
    args = ap.parse_args()
    W, H = (int(x) for x in args.size.split("x"))
    earth = pkg.load_earth()
    flags = pkg.FLAG_STRICT_MATH if args.strict else 0
    res = {"config": {"width": W, "height": H, "spp": args.spp, "depth": 50, "seed": 21, "strict_math": bool(args.strict),
                      "oracle_order": "beta" if args.strict else "reference", "doc": __doc__.split("usage:")[0].strip()}}
    for key in args.scenes.split(","):
        name, lit = key.split("+")[0], key.endswith("+lit")
        hs = pkg.HostScene(name, lit=lit, synth_n=30)
        d = pkg.desc_arrays(hs.desc(W, H).contents)
        ds = hs.upload(W, H)
        osc = O.OracleScene(name, earth=earth if name in ("earth", "random", "next_week_final") else None, lights_mode=int(lit), synth_n=30)
        res[key] = {}
        for pn, gprec, oprec in (("f32", pkg.PREC_F32, O.PREC_F32), ("mixed", pkg.PREC_MIXED, O.PREC_MIXED), ("f64", pkg.PREC_F64, O.PREC_F64)):
            _, gp, gt = ds.render_pathlog(W, H, args.spp, 50, depth=50, seed=21, precision=gprec, flags=flags)
            _, op, ot = osc.render(W, H, args.spp, depth=50, prec=oprec, mode=1, seed=21, pathlog=50, strict=args.strict, order="beta" if args.strict else "reference")
            res[key][pn] = classify(gp, gt, op, ot, prim_material(d))
            r = res[key][pn]
            print("%-22s %-5s paths %d identical %d drift %d self-hit %d hit/miss %d other %d  prev-t-identical %d  by %s" % (
                key, pn, r["paths"], r["identical"], r["value_drift"], r["self_hit_flip"], r["hit_miss_flip"], r["other_flip"], r["prev_t_identical"],
                r["by_scatter_material"]), flush=True)
        ds.destroy()
    json.dump(res, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
