"""GPU parity at settings GUAVA leaves at their defaults but the reference's interface accepts
(`GaussianRasterizationSettings`, DGR32/diff_gaussian_rasterization_32/__init__.py:143-156): `scale_modifier != 1`,
`antialiasing=True` with gradients, a close wide-angle camera (the +-1.3 tanfov clamp of computeCov2D and its
clamp-aware gradient, forward.cu:84-87 / backward.cu:182-183; the near-plane cull of in_frustum, auxiliary.h:151-176)
and rotations that are not unit quaternions (computeCov3D uses q as it comes, forward.cu:123).  Each case: the
bit-exact forward gates and the gradient gates of tests/test_gpu_parity.py against the unmodified reference build.
"""
import numpy as np
import pytest
import torch

from tests import util

pytestmark = pytest.mark.gpu

import guava_renderer_b200 as grb  # noqa: E402
import diff_gaussian_rasterization_32 as ours  # noqa: E402
from guava_renderer_b200 import synthetic  # noqa: E402

DEV = "cuda:0"
CAP = 1 << 22


@pytest.fixture(autouse=True)
def _options():
    grb.set_option("emit_sorted_keys", 1)
    grb.set_option("capacity_hint", CAP)
    yield
    grb.set_option("emit_sorted_keys", 0)
    grb.set_option("capacity_hint", 0)
    grb.set_option("sort_mode", -1)


@pytest.fixture(scope="module")
def ref():
    m = util.import_reference()
    if m is None:
        pytest.skip("oracle/_ref (reference CUDA build) not present")
    return m


def _bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


def _close_camera(inp, radius, tanfov):
    """Replace the orbit camera of `inp` by one `radius` away from the cloud centre with a wide field of view."""
    H, W = inp["H"], inp["W"]
    cam = synthetic.camera(0, 1, tanfov * W / H, tanfov, radius=radius)
    inp.update(viewmatrix=cam["viewmatrix"], projmatrix=cam["projmatrix"], campos=cam["campos"],
               tanfovx=tanfov * W / H, tanfovy=tanfov)
    return inp


def _forward_gates(ref, inp, antialiasing, scale_modifier):
    P, H, W = inp["xyz"].shape[0], inp["H"], inp["W"]
    R, color, radii, geom, binning, img, invd = util.c_forward(ours, inp, DEV, antialiasing, scale_modifier)
    torch.cuda.synchronize()
    o = util.decode_ours(geom, binning, img, 1, P, W, H, CAP)
    Rr, color_r, radii_r, geom_r, binning_r, img_r, invd_r = util.c_forward(ref, inp, DEV, antialiasing, scale_modifier)
    torch.cuda.synchronize()
    r = util.decode_reference(geom_r, binning_r, img_r, P, H, W, Rr)
    radii, radii_r = radii.cpu().numpy(), radii_r.cpu().numpy()
    vis = radii_r > 0
    assert R == Rr
    assert np.array_equal(radii, radii_r)
    assert np.array_equal(o["keys"], r["keys"])
    assert np.array_equal(o["point_list"], r["point_list"])
    assert np.array_equal(o["ranges"], r["ranges"])
    assert np.array_equal(o["n_contrib"], r["n_contrib"])
    assert np.array_equal(_bits(o["means2D"][vis]), _bits(r["means2D"][vis]))
    assert np.array_equal(_bits(o["conic_opacity"][vis]), _bits(r["conic_opacity"][vis]))
    assert np.array_equal(_bits(o["final_T"]), _bits(r["final_T"]))
    assert np.abs(color.cpu().numpy() - color_r.cpu().numpy()).max() <= 1e-5
    assert np.abs(invd.cpu().numpy() - invd_r.cpu().numpy()).max() <= 1e-5
    return int(vis.sum()), Rr


def _gradients(mod, inp, dL, dLi, antialiasing, scale_modifier):
    """The call pattern of models/UbodyAvatar/gaussian_render.py:30-63 for one view."""
    leaf = {k: inp[k].to(DEV).clone().requires_grad_(True) for k in ("xyz", "features", "opacity", "scaling", "rotation")}
    m2 = torch.zeros_like(leaf["xyz"], requires_grad=True)
    st = util.settings_for(mod, inp, DEV, antialiasing, scale_modifier=scale_modifier)
    color, radii, invd = mod.GaussianRasterizer_32(raster_settings=st)(
        means3D=leaf["xyz"], means2D=m2, shs=None, colors_precomp=leaf["features"], opacities=leaf["opacity"],
        scales=leaf["scaling"], rotations=leaf["rotation"], cov3D_precomp=None)
    ((color * dL).sum() + (invd * dLi).sum()).backward()
    torch.cuda.synchronize()
    g = {k: v.grad.cpu().numpy() for k, v in leaf.items()}
    g["means2D"] = m2.grad.cpu().numpy()
    return g


def _gradient_gates(ref, inp, antialiasing, scale_modifier, seed=17):
    g = torch.Generator().manual_seed(seed)
    dL = torch.randn(32, inp["H"], inp["W"], generator=g).to(DEV)
    dLi = torch.randn(1, inp["H"], inp["W"], generator=g).to(DEV)
    go = _gradients(ours, inp, dL, dLi, antialiasing, scale_modifier)
    gr = _gradients(ref, inp, dL, dLi, antialiasing, scale_modifier)
    for k in gr:
        assert np.isfinite(go[k]).all() == np.isfinite(gr[k]).all(), k
        assert util.rel_err(go[k], gr[k]) <= 1e-4, (k, util.rel_err(go[k], gr[k]))
        util.grad_gate(k, go[k], gr[k])
    return gr


@pytest.mark.parametrize("sort_mode", [0, 1], ids=["buckets", "onesweep"])
@pytest.mark.parametrize("antialiasing", [False, True], ids=["plain", "antialiasing"])
@pytest.mark.parametrize("scale_modifier", [0.6, 1.7])
def test_scale_modifier_and_antialiasing_vs_reference_build(ref, scale_modifier, antialiasing, sort_mode):
    grb.set_option("sort_mode", sort_mode)
    inp = util.inputs_for(P=6000, H=144, W=208, seed=61, scale_mult=1.2)
    inp["bg"] = torch.linspace(1, -1, 32)
    visible, R = _forward_gates(ref, inp, antialiasing, scale_modifier)
    assert visible > 3000 and R > visible
    gr = _gradient_gates(ref, inp, antialiasing, scale_modifier)
    assert np.abs(gr["scaling"]).max() > 0 and np.abs(gr["rotation"]).max() > 0


@pytest.mark.parametrize("antialiasing", [False, True], ids=["plain", "antialiasing"])
def test_close_wide_camera_vs_reference_build(ref, antialiasing):
    """Camera inside the shell (0.1 from its centre), tanfov 0.9, large splats: most Gaussians are behind the near
    plane (culled: radius 0, no gradient) and several hundred of the visible ones project outside +-1.3 tanfov
    (clamped Jacobian, gradient multiplier 0 on that axis) while their footprint still reaches the image."""
    inp = util.inputs_for(P=4000, H=128, W=160, seed=62, scale_mult=1.5)
    inp["bg"] = torch.linspace(-0.5, 0.5, 32)
    inp = _close_camera(inp, radius=0.1, tanfov=0.9)
    view = inp["viewmatrix"].numpy()            # row-vector convention: p_view = [p, 1] @ view
    pv = np.concatenate([inp["xyz"].numpy(), np.ones((4000, 1), np.float32)], 1) @ view
    culled = int((pv[:, 2] <= 0.2).sum())
    lim = 1.3 * np.array([inp["tanfovx"], inp["tanfovy"]])
    front = pv[:, 2] > 0.2
    clamped = np.zeros(4000, bool)
    clamped[front] = (np.abs(pv[front, :2] / pv[front, 2:3]) > lim).any(1)
    assert culled > 2000 and clamped.sum() > 500, (culled, clamped.sum())     # the case does what its name says
    visible, R = _forward_gates(ref, inp, antialiasing, 1.0)
    assert 500 < visible <= 4000 - culled
    gr = _gradient_gates(ref, inp, antialiasing, 1.0)
    behind = pv[:, 2] <= 0.2
    assert not gr["xyz"][behind].any() and not gr["features"][behind].any()
    assert (np.abs(gr["xyz"][clamped]).sum(1) > 0).sum() > 100        # clamped splats that reach the image


def test_rotations_that_are_not_unit_quaternions_vs_reference_build(ref):
    inp = util.inputs_for(P=5000, H=128, W=128, seed=63, scale_mult=1.0)
    g = torch.Generator().manual_seed(3)
    inp["rotation"] = inp["rotation"] * (0.5 + 1.5 * torch.rand(5000, 1, generator=g))      # |q| in [0.5, 2]
    inp["bg"] = torch.linspace(0, 1, 32)
    _forward_gates(ref, inp, False, 1.0)
    _gradient_gates(ref, inp, False, 1.0)
