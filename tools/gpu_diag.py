#!/usr/bin/env python
"""Stage-by-stage mismatch report: ours vs the CPU oracle vs the reference CUDA build.
Development aid for the GPU box (tests/ hold the actual gates).  Writes gpurun_out/diag.json."""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tests import util  # noqa: E402
import guava_renderer_b200 as grb  # noqa: E402
import diff_gaussian_rasterization_32 as ours  # noqa: E402

DEV = "cuda:0"
CAP = 1 << 22


def bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


def cmp_exact(a, b):
    a, b = np.asarray(a), np.asarray(b)
    if a.shape != b.shape:
        return {"shape_mismatch": [list(a.shape), list(b.shape)]}
    ne = a != b
    out = {"n": int(a.size), "mismatch": int(ne.sum())}
    if ne.any():
        idx = np.argwhere(ne)[:5].tolist()
        out["first"] = idx
        out["a"] = [float(a[tuple(i)]) for i in idx]
        out["b"] = [float(b[tuple(i)]) for i in idx]
    return out


def cmp_float(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    d = np.abs(a - b)
    return {"max_abs": float(d.max()) if d.size else 0.0, "frac_gt_1e-5": float((d > 1e-5).mean()) if d.size else 0.0,
            "rel": util.rel_err(a, b)}


def main():
    grb.set_option("emit_sorted_keys", 1)
    grb.set_option("capacity_hint", CAP)
    ref = util.import_reference()
    report = {"reference_build": ref is not None, "device": torch.cuda.get_device_name(0)}
    cases = [dict(P=10_000, H=512, W=512, seed=1000, scale_mult=1.0),
             dict(P=3_000, H=200, W=152, seed=21, scale_mult=1.5),
             dict(P=20_000, H=128, W=256, seed=33, scale_mult=3.0)]
    for ci, case in enumerate(cases):
        for aa in (False, True):
            inp = util.inputs_for(**case)
            inp["bg"] = torch.linspace(-1, 1, 32)
            P = case["P"]
            t0 = time.time()
            R, color, radii, geom, binning, img, invd = util.c_forward(ours, inp, DEV, aa)
            torch.cuda.synchronize()
            o = util.decode_ours(geom, binning, img, 1, P, inp["W"], inp["H"], CAP)
            o.update(radii=radii.cpu().numpy(), color=color.cpu().numpy(), invdepth=invd.cpu().numpy())
            f = util.oracle_forward(inp, antialiasing=aa)
            vis = f["radii"] > 0
            rep = {"R_ours": R, "R_oracle": f["num_rendered"], "time_s": time.time() - t0}
            rep["vs_oracle"] = {
                "radii": cmp_exact(o["radii"], f["radii"]),
                "tiles_touched": cmp_exact(o["tiles_touched"], f["tiles_touched"]),
                "depths": cmp_exact(bits(o["depths"][vis]), bits(f["depths"][vis])),
                "means2D": cmp_exact(bits(o["means2D"][vis]), bits(f["means2D"][vis])),
                "conic_opacity": cmp_exact(bits(o["conic_opacity"][vis]), bits(f["conic_opacity"][vis])),
                "ranges": cmp_exact(o["ranges"], f["ranges"]),
                "point_list": cmp_exact(o["point_list"], f["point_list"]),
                "keys": cmp_exact(o.get("keys", np.zeros(0)), f["keys"]),
                "n_contrib": cmp_exact(o["n_contrib"], f["n_contrib"]),
                "color": cmp_float(o["color"], f["color"]),
                "invdepth": cmp_float(o["invdepth"], f["invdepth"]),
            }
            if ref is not None:
                Rr, colr, radr, gr, br, ir, invr = util.c_forward(ref, inp, DEV, aa)
                torch.cuda.synchronize()
                r = util.decode_reference(gr, br, ir, P, inp["H"], inp["W"], Rr)
                radr = radr.cpu().numpy()
                visr = radr > 0
                rep["R_ref"] = Rr
                rep["vs_reference"] = {
                    "radii": cmp_exact(o["radii"], radr),
                    "tiles_touched": cmp_exact(o["tiles_touched"], r["tiles_touched"]),
                    "depths": cmp_exact(bits(o["depths"][visr]), bits(r["depths"][visr])),
                    "means2D": cmp_exact(bits(o["means2D"][visr]), bits(r["means2D"][visr])),
                    "conic_opacity": cmp_exact(bits(o["conic_opacity"][visr]), bits(r["conic_opacity"][visr])),
                    "ranges": cmp_exact(o["ranges"], r["ranges"]),
                    "point_list": cmp_exact(o["point_list"], r["point_list"]),
                    "keys": cmp_exact(o.get("keys", np.zeros(0)), r["keys"]),
                    "n_contrib": cmp_exact(o["n_contrib"], r["n_contrib"]),
                    "final_T": cmp_exact(bits(o["final_T"]), bits(r["final_T"])),
                    "color": cmp_float(o["color"], colr.cpu().numpy()),
                    "invdepth": cmp_float(o["invdepth"], invr.cpu().numpy()),
                }
                rep["oracle_vs_reference"] = {
                    "radii": cmp_exact(f["radii"], radr),
                    "conic_opacity": cmp_exact(bits(f["conic_opacity"][visr]), bits(r["conic_opacity"][visr])),
                    "means2D": cmp_exact(bits(f["means2D"][visr]), bits(r["means2D"][visr])),
                    "cov3D": cmp_exact(bits(f["cov3D"][visr]), bits(r["cov3D"][visr])),
                    "keys": cmp_exact(f["keys"], r["keys"]),
                    "point_list": cmp_exact(f["point_list"], r["point_list"]),
                    "n_contrib": cmp_exact(f["n_contrib"], r["n_contrib"]),
                    "color": cmp_float(f["color"], colr.cpu().numpy()),
                }
            report[f"case{ci}_aa{int(aa)}"] = rep
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "diag.json"), "w") as fh:
        json.dump(report, fh, indent=1)
    # compact console summary
    for k, v in report.items():
        if not isinstance(v, dict):
            print(k, v)
            continue
        for grp in ("vs_oracle", "vs_reference", "oracle_vs_reference"):
            if grp in v:
                bad = {n: (x.get("mismatch", x.get("max_abs"))) for n, x in v[grp].items()}
                print(k, grp, bad)


if __name__ == "__main__":
    main()
