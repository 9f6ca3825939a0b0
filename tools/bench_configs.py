#!/usr/bin/env python
"""Timing of the other BASELINE.json configs (C1, C2, C4, C5): ours vs the unmodified reference build,
CUDA-event timed on one B200.  Informational table for BASELINE.md / DESIGN.md (bench.py owns the headline)."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tests import util  # noqa: E402
import guava_renderer_b200 as grb  # noqa: E402
import diff_gaussian_rasterization_32 as ours  # noqa: E402
from guava_renderer_b200 import synthetic  # noqa: E402

DEV = "cuda:0"


def timeit(fn, warm=3, reps=10):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def per_view_loop(mod, views, dL, backward):
    """gaussian_render.py:37-67: one rasterizer call per view."""
    def run():
        for inp, d in zip(views, dL):
            leaf = {k: inp[k].requires_grad_(backward) for k in ("xyz", "features", "opacity", "scaling", "rotation")}
            m2 = torch.zeros_like(leaf["xyz"], requires_grad=backward)
            rast = mod.GaussianRasterizer_32(raster_settings=inp["settings_" + mod.__name__])
            with torch.set_grad_enabled(backward):
                color, radii, invd = rast(means3D=leaf["xyz"], means2D=m2, shs=None, colors_precomp=leaf["features"],
                                          opacities=leaf["opacity"], scales=leaf["scaling"], rotations=leaf["rotation"],
                                          cov3D_precomp=None)
            if backward:
                color.backward(d)
                for v in leaf.values():
                    v.grad = None
    return run


def main():
    ref = util.import_reference()
    rows = []
    cfgs = [("C1", 10_000, 512, 512, 1, 1.0, True), ("C2", 100_000, 512, 512, 8, 1.0, False),
            ("C4", 300_000, 1024, 1024, 4, 1.0, False), ("C5", 1_000_000, 1080, 1920, 1, 4.0, True)]
    for name, P, H, W, V, sm, bwd in cfgs:
        views = []
        for v in range(V):
            inp = util.inputs_for(P, H, W, 1000 * int(name[1]) + v, view=v, num_views=max(V, 2), scale_mult=sm)
            d = {k: (t.to(DEV) if torch.is_tensor(t) else t) for k, t in inp.items()}
            for mod in [m for m in (ours, ref) if m is not None]:
                d["settings_" + mod.__name__] = util.settings_for(mod, inp, DEV)
            views.append(d)
        dL = [synthetic.upstream_grads(H, W, 3000 + v)[0].to(DEV) for v in range(V)]
        row = dict(config=name, P=P, H=H, W=W, views=V, passes="fwd+bwd" if bwd else "fwd")
        t = timeit(per_view_loop(ours, views, dL, bwd))
        row["ours_per_view_loop_ms"] = round(t, 4)
        if ref is not None:
            row["reference_ms"] = round(timeit(per_view_loop(ref, views, dL, bwd)), 4)
        if V > 1 and not bwd:   # batched entry: one launch chain for all views
            st = grb.BatchRasterizationSettings(
                H, W, torch.tensor([[x["tanfovx"], x["tanfovy"]] for x in views], device=DEV),
                torch.zeros(V, 32, device=DEV), 1.0, torch.stack([x["viewmatrix"] for x in views]),
                torch.stack([x["projmatrix"] for x in views]))
            stack = lambda k: torch.stack([x[k] for x in views]).detach()
            args = (stack("xyz"), torch.zeros(V, P, 3, device=DEV), stack("features"), stack("opacity"),
                    stack("scaling"), stack("rotation"), st)
            with torch.no_grad():
                row["ours_batched_ms"] = round(timeit(lambda: grb.rasterize_gaussians_batch(*args)), 4)
            step = grb.GraphedRenderStep(args[0], args[2], args[3], args[4], args[5], st)
            row["ours_graphed_ms"] = round(timeit(step.replay), 4)
            step.check()
        best = min(v for k, v in row.items() if k.startswith("ours_"))
        row["Mpix_per_s_ours"] = round(V * H * W / best / 1e3, 1)
        if ref is not None:
            row["Mpix_per_s_reference"] = round(V * H * W / row["reference_ms"] / 1e3, 1)
            row["speedup"] = round(row["reference_ms"] / best, 2)
        rows.append(row)
        print(json.dumps(row), flush=True)


if __name__ == "__main__":
    main()
