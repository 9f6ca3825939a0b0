#!/usr/bin/env python
"""Small end-to-end exercise of every kernel variant, meant to run under compute-sanitizer:
    compute-sanitizer --tool memcheck python tools/sanitize_case.py
single-view call (odd P: scalar input loads), batched call (V = 2, P % 4 == 0: vector loads), both binning modes,
taken marks on/off, TMA staging, backward with and without an inverse-depth gradient, graph replay."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import guava_renderer_b200 as grb  # noqa: E402
import diff_gaussian_rasterization_32 as dgr  # noqa: E402
from tests import util  # noqa: E402

dev = "cuda:0"


def single(P, H, W, with_inv):
    inp = util.inputs_for(P, H, W, seed=5, scale_mult=1.5)
    leaf = {k: inp[k].to(dev).requires_grad_(True) for k in ("xyz", "features", "opacity", "scaling", "rotation")}
    m2 = torch.zeros_like(leaf["xyz"], requires_grad=True)
    rast = dgr.GaussianRasterizer_32(util.settings_for(dgr, inp, dev))
    color, radii, invd = rast(means3D=leaf["xyz"], means2D=m2, opacities=leaf["opacity"], colors_precomp=leaf["features"],
                              scales=leaf["scaling"], rotations=leaf["rotation"])
    loss = color.sum() + (invd.sum() if with_inv else 0.0)
    loss.backward()
    torch.cuda.synchronize()
    return float(color.sum())


def batched(V, P, H, W):
    views = [util.inputs_for(P, H, W, seed=70 + v, view=v, num_views=V, scale_mult=1.3) for v in range(V)]
    stack = lambda k: torch.stack([x[k] for x in views], 0).to(dev)
    tanfov = torch.tensor([[x["tanfovx"], x["tanfovy"]] for x in views], device=dev)
    st = grb.BatchRasterizationSettings(H, W, tanfov, torch.zeros(V, 32, device=dev), 1.0, stack("viewmatrix"), stack("projmatrix"))
    dL = torch.randn(V, 32, H, W, device=dev)
    step = grb.GraphedRenderStep(stack("xyz"), stack("features"), stack("opacity"), stack("scaling"), stack("rotation"), st, dL_dcolor=dL)
    step.replay()
    step.check()
    return float(step.packed.abs().sum())


for masks in (1, 0):
    for tma in (0, 1):
        for mode in (0, 1):
            grb.set_option("taken_masks", masks)
            grb.set_option("tma_staging", tma)
            grb.set_option("sort_mode", mode)
            a = single(1237, 72, 100, with_inv=(mode == 1))
            b = batched(2, 1500, 64, 80)
            print("masks", masks, "tma", tma, "sort_mode", mode, "ok", round(a, 3), round(b, 3), flush=True)
grb.set_option("sort_mode", -1)
grb.set_option("tma_staging", 0)
grb.set_option("taken_masks", 1)
single(30000, 64, 64, False)      # long tile lists: mid / large bucket sorts
print("done")
