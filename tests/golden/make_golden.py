#!/usr/bin/env python
"""Generate the golden vectors of tests/golden/*.npz.

Run ON THE GPU BOX (needs a CUDA device and oracle/_ref, the unmodified reference CUDA build made by
oracle/build_ref.py from /root/reference):

    python tests/golden/make_golden.py gpurun_out/golden      # then copy the .npz files to tests/golden/

Every file holds the outputs of the REFERENCE rasterizer (its `_C.rasterize_gaussians` and autograd
backward) for one small seeded synthetic case; inputs are regenerated from the seed by
tests/util.inputs_for, so only outputs are stored.  The CPU suite checks oracle/splat_oracle.c against
these files (tests/test_oracle_cpu.py::test_oracle_against_reference_golden); that is what pins the oracle.
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from tests import util  # noqa: E402
from guava_renderer_b200 import synthetic  # noqa: E402

CASES = [
    dict(name="p400_40x56", P=400, H=40, W=56, seed=101, scale_mult=0.35, antialiasing=False),
    dict(name="p400_40x56_aa", P=400, H=40, W=56, seed=101, scale_mult=0.35, antialiasing=True),
    dict(name="p1500_64x48", P=1500, H=64, W=48, seed=202, scale_mult=0.3, antialiasing=False),
]
DL_SEED, DL_SCALE = 4242, 1.0e4


def main(out_dir):
    ref = util.import_reference()
    assert ref is not None, "oracle/_ref is missing: run oracle/build_ref.py where /root/reference exists"
    os.makedirs(out_dir, exist_ok=True)
    dev = "cuda:0"
    for c in CASES:
        inp = util.inputs_for(c["P"], c["H"], c["W"], c["seed"], scale_mult=c["scale_mult"])
        inp["bg"] = torch.linspace(-0.5, 0.5, 32)
        R, color, radii, geom, binning, img, invd = util.c_forward(ref, inp, dev, c["antialiasing"])
        torch.cuda.synchronize()
        d = util.decode_reference(geom, binning, img, c["P"], c["H"], c["W"], R)
        # gradients through the reference's own autograd surface
        leaf = {k: inp[k].to(dev).clone().requires_grad_(True) for k in ("xyz", "features", "opacity", "scaling", "rotation")}
        m2 = torch.zeros_like(leaf["xyz"], requires_grad=True)
        rast = ref.GaussianRasterizer_32(raster_settings=util.settings_for(ref, inp, dev, c["antialiasing"]))
        col, _, _ = rast(means3D=leaf["xyz"], means2D=m2, shs=None, colors_precomp=leaf["features"],
                         opacities=leaf["opacity"], scales=leaf["scaling"], rotations=leaf["rotation"], cov3D_precomp=None)
        dc, _ = synthetic.upstream_grads(c["H"], c["W"], DL_SEED)
        (col * (dc * DL_SCALE).to(dev)).sum().backward()
        torch.cuda.synchronize()
        np.savez_compressed(
            os.path.join(out_dir, c["name"] + ".npz"),
            P=c["P"], H=c["H"], W=c["W"], seed=c["seed"], scale_mult=c["scale_mult"], antialiasing=c["antialiasing"],
            bg=inp["bg"].numpy(), num_rendered=R, radii=radii.cpu().numpy(), tiles_touched=d["tiles_touched"],
            depths=d["depths"], means2D=d["means2D"], conic_opacity=d["conic_opacity"], keys=d["keys"],
            point_list=d["point_list"], ranges=d["ranges"], n_contrib=d["n_contrib"], final_T=d["final_T"],
            color=color.cpu().numpy(), invdepth=invd.cpu().numpy(),
            dL_dcolor_seed=DL_SEED, dL_scale=DL_SCALE, with_invdepth_grad=True,
            dL_dmeans2D=m2.grad.cpu().numpy(), dL_dcolors=leaf["features"].grad.cpu().numpy(),
            dL_dopacity=leaf["opacity"].grad.cpu().numpy(), dL_dmeans3D=leaf["xyz"].grad.cpu().numpy(),
            dL_dscales=leaf["scaling"].grad.cpu().numpy(), dL_drotations=leaf["rotation"].grad.cpu().numpy(),
            generator="tests/golden/make_golden.py on " + torch.cuda.get_device_name(0))
        print("wrote", c["name"], "R =", R, "visible =", int((radii > 0).sum()))


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "golden"))
